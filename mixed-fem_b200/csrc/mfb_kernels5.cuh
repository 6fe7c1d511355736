// k_update_shell : MixedStretch<3>::update() on a TriMesh (triangles in 3D with a rest normal, src/mesh/tri_mesh.cpp;
// BASELINE config C5).  Same per-element sequence as the tet kernel (mixed_stretch.cpp:158-243, 345-366) with
//   F    = J0 x + n(x) (x) N                                   (tri_mesh.cpp:104-126)
//   Jloc = Jloc0 + N (I - n n^T) dn_dq / |v1 x v2|  (9 x 9)    (tri_mesh.cpp:128-190, update_jacobian)
//   vol  = doublearea * thickness / 2                          (tri_mesh.cpp:38-41)
// and 9 x 9 element blocks (3 vertices).  One thread per triangle, B = (Js Jloc)^T rows of vertices 1, 2 in registers
// (vertex 0 = -(1 + 2): both parts of Jloc have zero column-block sums), direct FP64 reductions into the
// reference-layout CSR.  Shell meshes of the reference are 1e4-1e5 triangles: this kernel is not the headline path.
#pragma once
#include "mfb_kernels2.cuh"

namespace mfb {

template <bool FUSE_RHS>
__global__ void __launch_bounds__(128) k_update_shell(DevMesh m, DevState st, const double* __restrict__ x, double h2) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m.nE) return;
  const int64_t ld = m.ld;
  int status = 0, sweeps = 0;
  ShellGeo geo;
  shell_geo(m, e, x, geo);
  double F[9], R[9], Y[9], s[6];
  shell_F(geo, F);
#pragma unroll
  for (int i = 0; i < 9; ++i) R[i] = st.R[i * ld + e];
#pragma unroll
  for (int k = 0; k < 6; ++k) s[k] = st.s[k * ld + e];
  const double vol = m.vol[e];
  const int iters = procrustes3(R, F, Y, 1e-6, 100, &status);
#pragma unroll
  for (int i = 0; i < 9; ++i) st.R[i * ld + e] = R[i];
  st.iters[e] = iters;
  double S[6];
  stretch6(R, F, S);
#pragma unroll
  for (int k = 0; k < 6; ++k) st.S[k * ld + e] = S[k];
  {
    double Sc[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) Sc[k] = S[k];
    const double q0 = Y[5] - Y[7], q1 = Y[6] - Y[2], q2 = Y[1] - Y[3];
    if (!(sqrt(q0 * q0 + q1 * q1 + q2 * q2) < 1e-6)) {     // evaluate_constraint's own polar (mixed_stretch.cpp:358-360)
      M33 Rc, Fc;
#pragma unroll
      for (int i = 0; i < 9; ++i) { Rc.a[i] = R[i]; Fc.a[i] = F[i]; }
      const V6 r = polar_from_copy_slow(Rc, Fc);
#pragma unroll
      for (int k = 0; k < 6; ++k) Sc[k] = r.a[k];
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) st.grad_la[k * ld + e] = ((k < 3) ? 1.0 : 2.0) * (Sc[k] - s[k]) * vol;
  }
  // dSdF_ and B = (Js Jloc)^T : rows [vertex 1: 0..2, vertex 2: 3..5] x 6
  double B[6][6];
  {
    double Hinv3[6];
    if (!procrustes_hinv(Y, Hinv3)) status |= ST_POLAR_HFAIL;
#pragma unroll
    for (int n = 0; n < 6; ++n) {
      double col[9], a1[3], a2[3];
      dsdf3_column(n, R, F, Hinv3, col);
      const double sy = (n < 3) ? 1.0 : 2.0;
#pragma unroll
      for (int k = 0; k < 9; ++k) st.dSdF[(n * 9 + k) * ld + e] = sy * col[k];
      shell_JT(geo, col, a1, a2);
#pragma unroll
      for (int r = 0; r < 3; ++r) { B[r][n] = a1[r]; B[3 + r][n] = a2[r]; }
    }
    if (!(B[0][0] == B[0][0])) status |= ST_NONFINITE;
  }
  // material, PSD fix (fast path, else the Jacobi eigen-solver inline)
  double g[6], Hf[21], Hinv[21];
  {
    const MaterialDev mat = m.mats[m.matid ? m.matid[e] : 0];
    double Hm[21], L[21], lmax;
    material_grad_hess6(mat.model, mat.mu, mat.la, s, h2, g, Hm);
    int nit;
    if (psd_fast6(Hm, L, Hinv, 1e-8, 1e-5, &lmax, &nit)) {
#pragma unroll
      for (int i = 0; i < 21; ++i) Hf[i] = Hm[i];
    } else {
#pragma unroll
      for (int i = 0; i < 21; ++i) Hf[i] = Hm[i];
      lmax = psd_fix_invert<6>(Hf, Hinv, 1e-8, &sweeps);
    }
    st.evals[e] = h2 * lmax;
    if (!(lmax == lmax) || lmax > 1e300) status |= ST_NONFINITE;
    const double iv = 1.0 / vol;
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int j = 0; j <= i; ++j) {
        const double w = ((i < 3) ? 1.0 : 0.5) * ((j < 3) ? 1.0 : 0.5);
        st.H[symidx(i, j) * ld + e] = iv * w * Hf[symidx(i, j)];
        st.Hinv[symidx(i, j) * ld + e] = Hinv[symidx(i, j)];
      }
#pragma unroll
    for (int k = 0; k < 6; ++k) st.g[k * ld + e] = g[k];
  }
  double u[6];
  if (FUSE_RHS) {
#pragma unroll
    for (int i = 0; i < 6; ++i) {
      double a = g[i];
#pragma unroll
      for (int j = 0; j < 6; ++j) a += Hf[symidx(i, j)] * (S[j] - s[j]);
      u[i] = a;
      st.gl[i * ld + e] = ((i < 3) ? 1.0 : 0.5) * a;
    }
  }
  int f[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) f[k] = m.fidx[k * ld + e];
  // vertex vectors
  {
    double tl[6];
#pragma unroll
    for (int n = 0; n < 6; ++n) tl[n] = ((n < 3) ? 1.0 : 2.0) * st.la[n * ld + e];
    double gx0[3] = {0, 0, 0}, rh0[3] = {0, 0, 0};
#pragma unroll
    for (int k = 0; k < 2; ++k)
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        double a = 0.0, b = 0.0;
#pragma unroll
        for (int n = 0; n < 6; ++n) { a += B[3 * k + r][n] * tl[n]; if (FUSE_RHS) b += B[3 * k + r][n] * u[n]; }
        a *= vol; b *= -vol;
        gx0[r] -= a; rh0[r] -= b;
        if (f[k + 1] >= 0) {
          red_add(&st.grad_x[3 * (int64_t)f[k + 1] + r], a);
          if (FUSE_RHS) red_add(&st.rhs[3 * (int64_t)f[k + 1] + r], b);
        }
      }
    if (f[0] >= 0) {
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        red_add(&st.grad_x[3 * (int64_t)f[0] + r], gx0[r]);
        if (FUSE_RHS) red_add(&st.rhs[3 * (int64_t)f[0] + r], rh0[r]);
      }
    }
  }
  // matrix: rows Bf = [-(B1 + B2); B1; B2] (9 x 6), W = vol Bf Hfix, block (j,k) = W_j Bf_k^T
  {
    double Bf[9][6], W[9][6];
#pragma unroll
    for (int n = 0; n < 6; ++n) {
#pragma unroll
      for (int r = 0; r < 3; ++r) { Bf[r][n] = -(B[r][n] + B[3 + r][n]); Bf[3 + r][n] = B[r][n]; Bf[6 + r][n] = B[3 + r][n]; }
    }
#pragma unroll
    for (int q = 0; q < 9; ++q)
#pragma unroll
      for (int n = 0; n < 6; ++n) {
        double a = 0.0;
#pragma unroll
        for (int i = 0; i < 6; ++i) a += Bf[q][i] * Hf[symidx(i, n)];
        W[q][n] = vol * a;
      }
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const int stride = m.sstride[j * ld + e];
#pragma unroll
      for (int k = 0; k < 3; ++k) {
        const int idx = m.smap[(j * 3 + k) * ld + e];
        if (idx < 0) continue;
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            double a = 0.0;
#pragma unroll
            for (int n = 0; n < 6; ++n) a += W[3 * j + r][n] * Bf[3 * k + c][n];
            red_add(&st.Avals[(int64_t)idx + (int64_t)r * stride + c], a);
          }
      }
    }
  }
  count_status(st.counters, status, iters, sweeps);
}

}  // namespace mfb
