// k_update_tet_sm : same per-element work as k_update_tet (phases B, C, D of MixedStretch<3>::update, see
// mfb_kernels.cuh) with the 9x6 matrix B = (Js Jloc)^T staged in SHARED memory instead of registers.
//
// Why: B (54 doubles = 108 registers) made the register-resident variant sit at 255 registers with spills and 2
// blocks/SM, and that variant spent 2/3 of its issue slots waiting on the scoreboard (ncu profiles/r1c).  Each thread
// owns one column of a [54][128] double tile (thread t touches sB[i*128 + t] only: conflict-free, no barriers), which
// brings the kernel under 168 registers => 3 blocks/SM, and lets the scatter indices be loaded late, where their
// latency overlaps the B L product.
#pragma once
#include "mfb_kernels.cuh"

namespace mfb {

constexpr int kSmThreads = 128;
constexpr int kSmBytes = 54 * kSmThreads * (int)sizeof(double);   // 55 296 B per block

template <bool FUSE_RHS>
__device__ __forceinline__ void update_tet_sm_body(const DevMesh& m, const DevState& st, const double* __restrict__ x, int e, double* sB) {
  double* myB = sB + threadIdx.x;          // element (row, n) at myB[(6*row + n) * kSmThreads]
  const int64_t ld = m.ld;
  int status = 0;

  // ---- B : deformation gradient first (T -> x is the only dependent load chain, start it before anything else)
  double dphi[9], F[9], R[9], Y[9];
  tet_F(m, e, x, dphi, F);
#pragma unroll
  for (int i = 0; i < 9; ++i) R[i] = st.R[i * ld + e];
  double s[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) s[k] = st.s[k * ld + e];
  const double vol = m.vol[e];
  const int iters = procrustes3(R, F, Y, 1e-6, 100, &status);
#pragma unroll
  for (int i = 0; i < 9; ++i) st.R[i * ld + e] = R[i];
  st.iters[e] = iters;
  double u[6];    // S - s first, then Hfix (S - s) + g
  {
    double S[6];
    stretch6(R, F, S);
#pragma unroll
    for (int k = 0; k < 6; ++k) st.S[k * ld + e] = S[k];
    double Sc[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) Sc[k] = S[k];
    double Y2[9];
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
      for (int r = 0; r < 3; ++r) Y2[3 * c + r] = R[r] * F[c] + R[3 + r] * F[3 + c] + R[6 + r] * F[6 + c];
    const double q0 = Y2[5] - Y2[7], q1 = Y2[6] - Y2[2], q2 = Y2[1] - Y2[3];
    if (!(sqrt(q0 * q0 + q1 * q1 + q2 * q2) < 1e-6)) {     // evaluate_constraint's own polar (mixed_stretch.cpp:358-360)
      M33 Rc, Fc;
#pragma unroll
      for (int i = 0; i < 9; ++i) { Rc.a[i] = R[i]; Fc.a[i] = F[i]; }
      const V6 r = polar_from_copy_slow(Rc, Fc);
#pragma unroll
      for (int k = 0; k < 6; ++k) Sc[k] = r.a[k];
    }
#pragma unroll
    for (int k = 0; k < 6; ++k) {
      st.grad_la[k * ld + e] = ((k < 3) ? 1.0 : 2.0) * (Sc[k] - s[k]) * vol;
      u[k] = S[k] - s[k];
    }
  }

  // ---- C : dSdF_ and B (to shared memory)
  {
    double Hinv3[6];
    if (!procrustes_hinv(Y, Hinv3)) status |= ST_POLAR_HFAIL;
#pragma unroll
    for (int n = 0; n < 6; ++n) {
      double col[9];
      dsdf3_column(n, R, F, Hinv3, col);
      const double sy = (n < 3) ? 1.0 : 2.0;
#pragma unroll
      for (int k = 0; k < 9; ++k) st.dSdF[(n * 9 + k) * ld + e] = sy * col[k];
#pragma unroll
      for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int r = 0; r < 3; ++r)
          myB[(6 * (3 * k + r) + n) * kSmThreads] = dphi[3 * k] * col[r] + dphi[3 * k + 1] * col[3 + r] + dphi[3 * k + 2] * col[6 + r];
      if (n == 0 && !(col[0] == col[0])) status |= ST_NONFINITE;
    }
  }

  // ---- A's products: Cholesky factor of the PSD-fixed Hessian and g (k_material_psd6)
  double L[21];
#pragma unroll
  for (int i = 0; i < 21; ++i) L[i] = st.Lfac[i * ld + e];
  int f[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) f[k] = m.fidx[k * ld + e];
  double tl[6];
#pragma unroll
  for (int n = 0; n < 6; ++n) tl[n] = ((n < 3) ? 1.0 : 2.0) * st.la[n * ld + e];
  if (FUSE_RHS) {
    double t[6];
#pragma unroll
    for (int j = 0; j < 6; ++j) { double a = 0.0;
#pragma unroll
      for (int i = j; i < 6; ++i) a += L[symidx(i, j)] * u[i];
      t[j] = a; }
#pragma unroll
    for (int i = 0; i < 6; ++i) { double a = st.g[i * ld + e];
#pragma unroll
      for (int j = 0; j <= i; ++j) a += L[symidx(i, j)] * t[j];
      u[i] = a; st.gl[i * ld + e] = ((i < 3) ? 1.0 : 0.5) * a; }
  }

  // ---- D : vectors, and B <- sqrt(vol) B L in place (one row at a time through registers)
  int stride[4], sidx[16];
#pragma unroll
  for (int k = 0; k < 4; ++k) stride[k] = m.sstride[k * ld + e];
#pragma unroll
  for (int q = 0; q < 16; ++q) sidx[q] = m.smap[q * ld + e];
  const double sv = sqrt(vol);
  double B0[3][6];
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int n = 0; n < 6; ++n) B0[r][n] = 0.0;
  {
    double gx0[3] = {0, 0, 0}, rh0[3] = {0, 0, 0};
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        double row[6];
#pragma unroll
        for (int n = 0; n < 6; ++n) row[n] = myB[(6 * (3 * k + r) + n) * kSmThreads];
        double a = 0.0, b = 0.0;
#pragma unroll
        for (int n = 0; n < 6; ++n) { a += row[n] * tl[n]; if (FUSE_RHS) b += row[n] * u[n]; }
        a *= vol; b *= -vol;
        gx0[r] -= a; rh0[r] -= b;
        if (f[k + 1] >= 0) {
          red_add(&st.grad_x[3 * (int64_t)f[k + 1] + r], a);
          if (FUSE_RHS) red_add(&st.rhs[3 * (int64_t)f[k + 1] + r], b);
        }
#pragma unroll
        for (int j = 0; j < 6; ++j) {
          double t = 0.0;
#pragma unroll
          for (int i = j; i < 6; ++i) t += row[i] * L[symidx(i, j)];
          t *= sv;
          myB[(6 * (3 * k + r) + j) * kSmThreads] = t;
          B0[r][j] -= t;
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

  // ---- D : matrix blocks, upper triangle of the global matrix only (k_mirror_lower completes it)
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    double bj[3][6];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int n = 0; n < 6; ++n) bj[r][n] = (j == 0) ? B0[r][n] : myB[(6 * (3 * (j - 1) + r) + n) * kSmThreads];
#pragma unroll
    for (int k = j; k < 4; ++k) {
      const int idx_jk = sidx[j * 4 + k];
      if (idx_jk < 0) continue;
      const bool up = (j == k) || (f[j] < f[k]);
      const int64_t base = up ? (int64_t)idx_jk : (int64_t)sidx[k * 4 + j];
      const int64_t rs = up ? stride[j] : stride[k];
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        double bk[6];
#pragma unroll
        for (int n = 0; n < 6; ++n) bk[n] = (k == j) ? bj[c][n] : myB[(6 * (3 * (k - 1) + c) + n) * kSmThreads];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
          if (j == k && c < r) continue;
          double a = 0.0;
#pragma unroll
          for (int n = 0; n < 6; ++n) a += bj[r][n] * bk[n];
          red_add(&st.Avals[up ? base + r * rs + c : base + c * rs + r], a);
        }
      }
    }
  }
  count_status(st.counters, status, iters, 0);
}

template <bool FUSE_RHS>
__global__ void __launch_bounds__(kSmThreads, 3) k_update_tet_sm(DevMesh m, DevState st, const double* __restrict__ x, double h2) {
  extern __shared__ double sB[];
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m.nE) return;
  update_tet_sm_body<FUSE_RHS>(m, st, x, e, sB);
}

// Whole update() in one launch: the FP64-bound PSD phase and the reduction-bound assembly phase of different warps
// overlap on the SM (they saturate different units), and the g / Cholesky hand-off stays in L2.
template <bool FUSE_RHS>
__global__ void __launch_bounds__(kSmThreads, 3) k_update_tet_fused(DevMesh m, DevState st, const double* __restrict__ x, double h2) {
  extern __shared__ double sB[];
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m.nE) return;
  material_psd6_body(m, st, e, h2);
  update_tet_sm_body<FUSE_RHS>(m, st, x, e, sB);
}

}  // namespace mfb
