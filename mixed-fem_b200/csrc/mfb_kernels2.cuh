// Remaining kernels of the path: reset, non-fused rhs, solve, energy / constraint reductions, evaluate_constraint,
// matrix-free products and layout conversions.  Templated on DIM (3: tets, N=6, M=9; 2: triangles, N=3, M=4).
#pragma once
#include "mfb_kernels.cuh"

namespace mfb {

// Element kinds as template values: 3 = tetrahedra (TetrahedralMesh), 2 = 2D triangles (Tri2DMesh), K_SHELL = triangles in
// 3D with a rest normal (TriMesh, src/mesh/tri_mesh.cpp).  SD = space dimension, NPE = vertices per element.
constexpr int K_SHELL = 5;
template <int DIM> struct Dims {
  static constexpr int SD = DIM == 2 ? 2 : 3, N = SD == 3 ? 6 : 3, M = SD * SD, NPE = DIM == 3 ? 4 : 3, NP = N * (N + 1) / 2;
};

template <int DIM>
__device__ __forceinline__ double symw(int k) { return k < Dims<DIM>::SD ? 1.0 : 2.0; }
template <int DIM>
__device__ __forceinline__ double syminvw(int k) { return k < Dims<DIM>::SD ? 1.0 : 0.5; }

// ---- shells: F = J0 x + n(x) (x) N, Jloc = Jloc0 + N (I - n n^T) dn_dq / |v1 x v2|  (tri_mesh.cpp:104-190).
// dphi rows 1..2 are stored (6 components: (k-1)*3 + c), row 0 = -(row1 + row2); Nr = rest normal of the face.
struct ShellGeo { double dphi[6], Nr[3], v1[3], v2[3], n[3], il; };
__device__ __forceinline__ void shell_geo(const DevMesh& m, int e, const double* __restrict__ x, ShellGeo& g) {
  const int64_t ld = m.ld;
  int v[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) v[k] = m.T[k * ld + e];
#pragma unroll
  for (int i = 0; i < 6; ++i) g.dphi[i] = m.dphi[i * ld + e];
#pragma unroll
  for (int i = 0; i < 3; ++i) g.Nr[i] = m.nrest[i * ld + e];
#pragma unroll
  for (int l = 0; l < 3; ++l) {
    const double x0 = __ldg(&x[3 * (int64_t)v[0] + l]);
    g.v1[l] = __ldg(&x[3 * (int64_t)v[1] + l]) - x0;
    g.v2[l] = __ldg(&x[3 * (int64_t)v[2] + l]) - x0;
  }
  cross3(g.v1, g.v2, g.n);
  const double len = sqrt(g.n[0] * g.n[0] + g.n[1] * g.n[1] + g.n[2] * g.n[2]);
  g.il = 1.0 / len;
#pragma unroll
  for (int l = 0; l < 3; ++l) g.n[l] *= g.il;
}
__device__ __forceinline__ void shell_F(const ShellGeo& g, double* F) {
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int r = 0; r < 3; ++r) F[3 * c + r] = g.v1[r] * g.dphi[c] + g.v2[r] * g.dphi[3 + c] + g.n[r] * g.Nr[c];
}
// Jloc^T t for the 9-vector t: contributions a1, a2 to vertices 1, 2 (vertex 0 gets -(a1 + a2))
__device__ __forceinline__ void shell_JT(const ShellGeo& g, const double* t, double* a1, double* a2) {
  double w[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    a1[r] = g.dphi[0] * t[r] + g.dphi[1] * t[3 + r] + g.dphi[2] * t[6 + r];
    a2[r] = g.dphi[3] * t[r] + g.dphi[4] * t[3 + r] + g.dphi[5] * t[6 + r];
    w[r] = g.Nr[0] * t[r] + g.Nr[1] * t[3 + r] + g.Nr[2] * t[6 + r];
  }
  const double nw = g.n[0] * w[0] + g.n[1] * w[1] + g.n[2] * w[2];
  double z[3], c1[3], c2[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) z[r] = (w[r] - g.n[r] * nw) * g.il;
  cross3(g.v2, z, c1);      // (-[v2]x)^T z
  cross3(g.v1, z, c2);      // ([v1]x)^T z = -(v1 x z)
#pragma unroll
  for (int r = 0; r < 3; ++r) { a1[r] += c1[r]; a2[r] -= c2[r]; }
}
// Jloc d for vertex displacements d0, d1, d2 -> 9-vector
__device__ __forceinline__ void shell_J(const ShellGeo& g, const double* d0, const double* d1, const double* d2, double* out) {
  double e1[3], e2[3], q1[3], q2[3], a[3];
#pragma unroll
  for (int r = 0; r < 3; ++r) { e1[r] = d1[r] - d0[r]; e2[r] = d2[r] - d0[r]; }
  cross3(g.v2, e1, q1);     // v2 x (d1 - d0)  -> enters with a minus sign
  cross3(g.v1, e2, q2);     // v1 x (d2 - d0)
#pragma unroll
  for (int r = 0; r < 3; ++r) a[r] = q2[r] - q1[r];
  const double na = g.n[0] * a[0] + g.n[1] * a[1] + g.n[2] * a[2];
#pragma unroll
  for (int r = 0; r < 3; ++r) a[r] = (a[r] - g.n[r] * na) * g.il;
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int r = 0; r < 3; ++r) out[3 * c + r] = g.dphi[c] * e1[r] + g.dphi[3 + c] * e2[r] + g.Nr[c] * a[r];
}

// F for either element kind; dphi gets the d*d stored rows
template <int DIM>
__device__ __forceinline__ void elem_F(const DevMesh& m, int e, const double* __restrict__ x, double* dphi, double* F) {
  if constexpr (DIM == K_SHELL) {
    ShellGeo g;
    shell_geo(m, e, x, g);
    shell_F(g, F);
#pragma unroll
    for (int i = 0; i < 9; ++i) dphi[i] = i < 6 ? g.dphi[i] : 0.0;
    return;
  } else {
  const int64_t ld = m.ld;
  int v[DIM + 1];
#pragma unroll
  for (int k = 0; k <= DIM; ++k) v[k] = m.T[k * ld + e];
#pragma unroll
  for (int i = 0; i < DIM * DIM; ++i) dphi[i] = m.dphi[i * ld + e];
  double x0[DIM], dx[DIM * DIM];
#pragma unroll
  for (int l = 0; l < DIM; ++l) x0[l] = __ldg(&x[DIM * (int64_t)v[0] + l]);
#pragma unroll
  for (int k = 0; k < DIM; ++k)
#pragma unroll
    for (int l = 0; l < DIM; ++l) dx[DIM * k + l] = __ldg(&x[DIM * (int64_t)v[k + 1] + l]) - x0[l];
#pragma unroll
  for (int c = 0; c < DIM; ++c)
#pragma unroll
    for (int r = 0; r < DIM; ++r) {
      double a = 0.0;
#pragma unroll
      for (int k = 0; k < DIM; ++k) a += dx[DIM * k + r] * dphi[DIM * k + c];
      F[DIM * c + r] = a;
    }
  }
}

// polar without gradients from a given R (3D warm start; 2D closed form ignores R): stretch vector only
template <int DIM>
__device__ __forceinline__ void polar_stretch(double* R, const double* F, double* S, int* status, int* iters) {
  if constexpr (Dims<DIM>::SD == 3) {
    double Y[9];
    *iters = procrustes3(R, F, Y, 1e-6, 100, status);
    stretch6(R, F, S);
  } else {
    double Js[3][4];
    polar2(F, R, S, Js);
    *iters = 0;
  }
}

// scatter  sign * vol * Jloc^T t  into a free-vertex vector (t is the M-vector of an element)
template <int DIM>
__device__ __forceinline__ void scatter_vertex(const DevMesh& m, int e, const double* dphi, const double* t, double scale, double* out) {
  if constexpr (DIM == K_SHELL) {   // Jloc depends on the positions of the last update (m.xlast)
    ShellGeo g;
    shell_geo(m, e, m.xlast, g);
    double a1[3], a2[3];
    shell_JT(g, t, a1, a2);
    const int f0 = m.fidx[e], f1 = m.fidx[m.ld + e], f2 = m.fidx[2 * m.ld + e];
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      if (f1 >= 0) red_add(&out[3 * (int64_t)f1 + r], scale * a1[r]);
      if (f2 >= 0) red_add(&out[3 * (int64_t)f2 + r], scale * a2[r]);
      if (f0 >= 0) red_add(&out[3 * (int64_t)f0 + r], -scale * (a1[r] + a2[r]));
    }
    return;
  } else {
  const int64_t ld = m.ld;
  double v0[DIM];
#pragma unroll
  for (int r = 0; r < DIM; ++r) v0[r] = 0.0;
#pragma unroll
  for (int k = 0; k < DIM; ++k) {
    const int f = m.fidx[(k + 1) * ld + e];
#pragma unroll
    for (int r = 0; r < DIM; ++r) {
      double a = 0.0;
#pragma unroll
      for (int c = 0; c < DIM; ++c) a += dphi[DIM * k + c] * t[DIM * c + r];
      a *= scale;
      v0[r] -= a;
      if (f >= 0) red_add(&out[DIM * (int64_t)f + r], a);
    }
  }
  const int f0 = m.fidx[e];
  if (f0 >= 0) {
#pragma unroll
    for (int r = 0; r < DIM; ++r) red_add(&out[DIM * (int64_t)f0 + r], v0[r]);
  }
  }
}

// (PJW^T dx)_e : M-vector, = vol * sum_k dphi(k,c) dx[f_k][r] over free vertices
template <int DIM>
__device__ __forceinline__ void gather_Jdx(const DevMesh& m, int e, const double* __restrict__ dx, double vol, double* Jdx) {
  if constexpr (DIM == K_SHELL) {
    ShellGeo g;
    shell_geo(m, e, m.xlast, g);
    double d[3][3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
      const int f = m.fidx[k * m.ld + e];
#pragma unroll
      for (int r = 0; r < 3; ++r) d[k][r] = (f >= 0) ? __ldg(&dx[3 * (int64_t)f + r]) : 0.0;
    }
    shell_J(g, d[0], d[1], d[2], Jdx);
#pragma unroll
    for (int i = 0; i < 9; ++i) Jdx[i] *= vol;
    return;
  } else {
  const int64_t ld = m.ld;
  double d[DIM + 1][DIM], dp[DIM * DIM];
#pragma unroll
  for (int i = 0; i < DIM * DIM; ++i) dp[i] = m.dphi[i * ld + e];
#pragma unroll
  for (int k = 0; k <= DIM; ++k) {
    const int f = m.fidx[k * ld + e];
#pragma unroll
    for (int r = 0; r < DIM; ++r) d[k][r] = (f >= 0) ? __ldg(&dx[DIM * (int64_t)f + r]) : 0.0;
  }
#pragma unroll
  for (int c = 0; c < DIM; ++c) {
    double p0 = 0.0;
#pragma unroll
    for (int k = 0; k < DIM; ++k) p0 -= dp[DIM * k + c];
#pragma unroll
    for (int r = 0; r < DIM; ++r) {
      double a = p0 * d[0][r];
#pragma unroll
      for (int k = 0; k < DIM; ++k) a += dp[DIM * k + c] * d[k + 1][r];
      Jdx[DIM * c + r] = vol * a;
    }
  }
  }
}

// reset()  mixed_stretch.cpp:294-338 : R = I, polar with gradients, s_ = S_, la_ = 0, H_ = Hinv_ = I, g_ = 0
template <int DIM>
__global__ void __launch_bounds__(128) k_reset(DevMesh m, DevState st, const double* __restrict__ x) {
  using D = Dims<DIM>;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m.nE) return;
  const int64_t ld = m.ld;
  double dphi[D::M], F[D::M], R[D::M], S[D::N];
  int status = 0, iters = 0;
  elem_F<DIM>(m, e, x, dphi, F);
#pragma unroll
  for (int i = 0; i < D::M; ++i) R[i] = 0.0;
#pragma unroll
  for (int i = 0; i < D::SD; ++i) R[i * D::SD + i] = 1.0;
  if constexpr (D::SD == 3) {
    double Y[9], Hinv3[6];
    iters = procrustes3(R, F, Y, 1e-6, 100, &status);
    stretch6(R, F, S);
    if (!procrustes_hinv(Y, Hinv3)) status |= ST_POLAR_HFAIL;
#pragma unroll
    for (int n = 0; n < 6; ++n) {
      double col[9];
      dsdf3_column(n, R, F, Hinv3, col);
#pragma unroll
      for (int k = 0; k < 9; ++k) st.dSdF[(n * 9 + k) * ld + e] = symw<3>(n) * col[k];
    }
  } else {
    double Js[3][4];
    polar2(F, R, S, Js);
#pragma unroll
    for (int n = 0; n < 3; ++n)
#pragma unroll
      for (int k = 0; k < 4; ++k) st.dSdF[(n * 4 + k) * ld + e] = symw<2>(n) * Js[n][k];
  }
#pragma unroll
  for (int i = 0; i < D::M; ++i) st.R[i * ld + e] = R[i];
#pragma unroll
  for (int k = 0; k < D::N; ++k) { st.S[k * ld + e] = S[k]; st.s[k * ld + e] = S[k]; st.la[k * ld + e] = 0.0; st.ds[k * ld + e] = 0.0; st.g[k * ld + e] = 0.0; st.gl[k * ld + e] = 0.0; st.grad_la[k * ld + e] = 0.0; }
#pragma unroll
  for (int i = 0; i < D::N; ++i)
#pragma unroll
    for (int j = 0; j <= i; ++j) { st.H[symidx(i, j) * ld + e] = (i == j) ? 1.0 : 0.0; st.Hinv[symidx(i, j) * ld + e] = (i == j) ? 1.0 : 0.0; }
  st.evals[e] = 0.0;
  st.iters[e] = iters;
  count_status(st.counters, status, iters, 0);
}

// rhs()  mixed_stretch.cpp:245-275 from stored fields (used when the fused value is stale)
template <int DIM>
__global__ void __launch_bounds__(128) k_rhs(DevMesh m, DevState st) {
  using D = Dims<DIM>;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m.nE) return;
  const int64_t ld = m.ld;
  const double vol = m.vol[e];
  double diff[D::N], gl[D::N];
#pragma unroll
  for (int k = 0; k < D::N; ++k) diff[k] = symw<DIM>(k) * (st.S[k * ld + e] - st.s[k * ld + e]);
#pragma unroll
  for (int r = 0; r < D::N; ++r) {
    double a = 0.0;
#pragma unroll
    for (int k = 0; k < D::N; ++k) a += st.H[symidx(r, k) * ld + e] * diff[k];
    gl[r] = vol * a + syminvw<DIM>(r) * st.g[r * ld + e];
    st.gl[r * ld + e] = gl[r];
  }
  double t[D::M], dphi[D::M];
#pragma unroll
  for (int q = 0; q < D::M; ++q) {
    double a = 0.0;
#pragma unroll
    for (int n = 0; n < D::N; ++n) a += st.dSdF[(n * D::M + q) * ld + e] * gl[n];
    t[q] = a;
    dphi[q] = m.dphi[q * ld + e];
  }
  scatter_vertex<DIM>(m, e, dphi, t, -vol, st.rhs);
}

// grad_x_ only (mixed_stretch.cpp:224-232), for callers that changed la_ after update()
template <int DIM>
__global__ void __launch_bounds__(128) k_grad_x(DevMesh m, DevState st) {
  using D = Dims<DIM>;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m.nE) return;
  const int64_t ld = m.ld;
  double t[D::M], dphi[D::M];
#pragma unroll
  for (int q = 0; q < D::M; ++q) {
    double a = 0.0;
#pragma unroll
    for (int n = 0; n < D::N; ++n) a += st.dSdF[(n * D::M + q) * ld + e] * st.la[n * ld + e];
    t[q] = a;
    dphi[q] = m.dphi[q * ld + e];
  }
  scatter_vertex<DIM>(m, e, dphi, t, m.vol[e], st.grad_x);
}

// solve(dx)  mixed_stretch.cpp:277-292
template <int DIM>
__global__ void __launch_bounds__(128) k_solve(DevMesh m, DevState st, const double* __restrict__ dx) {
  using D = Dims<DIM>;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m.nE) return;
  const int64_t ld = m.ld;
  double Jdx[D::M];
  gather_Jdx<DIM>(m, e, dx, m.vol[e], Jdx);
  double t[D::N], la[D::N];
#pragma unroll
  for (int n = 0; n < D::N; ++n) {
    double a = 0.0;
#pragma unroll
    for (int q = 0; q < D::M; ++q) a += st.dSdF[(n * D::M + q) * ld + e] * Jdx[q];
    t[n] = a;
  }
#pragma unroll
  for (int r = 0; r < D::N; ++r) {
    double a = st.gl[r * ld + e];
#pragma unroll
    for (int k = 0; k < D::N; ++k) a += st.H[symidx(r, k) * ld + e] * t[k];
    la[r] = a;
    st.la[r * ld + e] = a;
  }
  double v[D::N];
#pragma unroll
  for (int k = 0; k < D::N; ++k) v[k] = st.g[k * ld + e] - symw<DIM>(k) * la[k];
#pragma unroll
  for (int r = 0; r < D::N; ++r) {
    double a = 0.0;
#pragma unroll
    for (int k = 0; k < D::N; ++k) a += st.Hinv[symidx(r, k) * ld + e] * v[k];
    st.ds[r * ld + e] = -a;
  }
}

// block-level sum, deterministic: per-block partials, then k_sum_partials adds them in index order
__device__ __forceinline__ double block_sum_128(double v, double* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  const int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) sh[w] = v;
  __syncthreads();
  double r = 0.0;
  if (threadIdx.x == 0) { for (int i = 0; i < (int)(blockDim.x >> 5); ++i) r += sh[i]; }
  return r;
}
static __global__ void k_sum_partials(const double* __restrict__ partials, int n, double* out) {
  __shared__ double sh[32];
  double a = 0.0;
  for (int i = threadIdx.x; i < n; i += blockDim.x) a += partials[i];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) a += __shfl_down_sync(0xffffffffu, a, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = a;
  __syncthreads();
  if (threadIdx.x == 0) { double r = 0.0; for (int i = 0; i < (int)(blockDim.x >> 5); ++i) r += sh[i]; *out = r; }
}

// energy()  mixed_stretch.cpp:93-107 ; s given component-major with leading dimension ld
template <int DIM>
__global__ void __launch_bounds__(128) k_energy(DevMesh m, const double* __restrict__ s_soa, double* partials) {
  using D = Dims<DIM>;
  __shared__ double sh[4];
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  double val = 0.0;
  if (e < m.nE) {
    double s[D::N];
#pragma unroll
    for (int k = 0; k < D::N; ++k) s[k] = s_soa[k * m.ld + e];
    const MaterialDev mat = m.mats[m.matid ? m.matid[e] : 0];
    const double psi = (D::SD == 3) ? material_energy6(mat.model, mat.mu, mat.la, s) : material_energy3(mat.model, mat.mu, mat.la, s);
    val = psi * m.vol[e];
  }
  const double r = block_sum_128(val, sh);
  if (threadIdx.x == 0) partials[blockIdx.x] = r;
}

// constraint_value()  mixed_stretch.cpp:109-136  (WRITE_C: evaluate_constraint :345-366 writes the vector instead)
template <int DIM, bool WRITE_C>
__global__ void __launch_bounds__(128) k_constraint(DevMesh m, DevState st, const double* __restrict__ x,
                                                    const double* __restrict__ s_soa, double* partials, double* c_out) {
  using D = Dims<DIM>;
  __shared__ double sh[4];
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  double val = 0.0;
  if (e < m.nE) {
    const int64_t ld = m.ld;
    double dphi[D::M], F[D::M], R[D::M], S[D::N];
    int status = 0, iters = 0;
    elem_F<DIM>(m, e, x, dphi, F);
#pragma unroll
    for (int i = 0; i < D::M; ++i) R[i] = st.R[i * ld + e];
    polar_stretch<DIM>(R, F, S, &status, &iters);
    const double vol = m.vol[e];
#pragma unroll
    for (int k = 0; k < D::N; ++k) {
      const double d = symw<DIM>(k) * (S[k] - s_soa[k * ld + e]);
      if (WRITE_C) c_out[k * ld + e] = d * vol;
      else val += st.la[k * ld + e] * d;
    }
    val *= vol;
  }
  if (!WRITE_C) {
    const double r = block_sum_128(val, sh);
    if (threadIdx.x == 0) partials[blockIdx.x] = r;
  }
}

// product_* operators  mixed_stretch.cpp:406-468 ; element vectors are component-major (ld), vertex vectors d*nF
template <int DIM>
__global__ void __launch_bounds__(128) k_product(DevMesh m, DevState st, int which, int transposed,
                                                 const double* __restrict__ xin, double* out) {
  using D = Dims<DIM>;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m.nE) return;
  const int64_t ld = m.ld;
  const double vol = m.vol[e];
  if (which == 0 || which == 1) {
    double xv[D::N];
#pragma unroll
    for (int k = 0; k < D::N; ++k) xv[k] = xin[k * ld + e];
#pragma unroll
    for (int r = 0; r < D::N; ++r) {
      double a = 0.0;
#pragma unroll
      for (int k = 0; k < D::N; ++k) {
        // Hloc_ = vol * Hfix = vol^2 Sym H_ Sym ; product_hessian_inv uses vol * Hinv_
        const double hv = (which == 0) ? vol * vol * symw<DIM>(r) * symw<DIM>(k) * st.H[symidx(r, k) * ld + e]
                                       : vol * st.Hinv[symidx(r, k) * ld + e];
        a += hv * xv[k];
      }
      out[r * ld + e] += a;
    }
  } else if (which == 2) {
#pragma unroll
    for (int k = 0; k < D::N; ++k) out[k * ld + e] += vol * symw<DIM>(k) * xin[k * ld + e];
  } else if (transposed) {   // out (d*nF) -= PJW * (dSdF x_e)
    double t[D::M], dphi[D::M];
#pragma unroll
    for (int q = 0; q < D::M; ++q) {
      double a = 0.0;
#pragma unroll
      for (int n = 0; n < D::N; ++n) a += st.dSdF[(n * D::M + q) * ld + e] * xin[n * ld + e];
      t[q] = a; dphi[q] = m.dphi[q * ld + e];
    }
    scatter_vertex<DIM>(m, e, dphi, t, -vol, out);
  } else {                   // out_e -= dSdF^T (PJW^T x)_e
    double Jdx[D::M];
    gather_Jdx<DIM>(m, e, xin, vol, Jdx);
#pragma unroll
    for (int n = 0; n < D::N; ++n) {
      double a = 0.0;
#pragma unroll
      for (int q = 0; q < D::M; ++q) a += st.dSdF[(n * D::M + q) * ld + e] * Jdx[q];
      out[n * ld + e] -= a;
    }
  }
}

// jacobian_x() = -PJW_ * blockdiag(dSdF_)  (mixed_stretch.cpp:385-390) as COO triplets, one (element, stretch dof n,
// local vertex j, component r) entry each: row = d*free(v_j) + r (-1 when v_j is pinned), col = N*e + n,
// value = -vol * sum_c Jloc(j; r, c-th column of F) dSdF(., n).  Entry index = ((e*N + n)*NPE + j)*SD + r.
template <int DIM>
__global__ void __launch_bounds__(128) k_jacobian_x_coo(DevMesh m, DevState st, int* __restrict__ rows, int* __restrict__ cols,
                                                        double* __restrict__ vals) {
  using D = Dims<DIM>;
  constexpr int SD = D::SD, NPE = D::NPE;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m.nE) return;
  const int64_t ld = m.ld;
  const double vol = m.vol[e];
  int f[NPE];
#pragma unroll
  for (int j = 0; j < NPE; ++j) f[j] = m.fidx[j * ld + e];
  [[maybe_unused]] ShellGeo geo;
  [[maybe_unused]] double dphi[D::M];
  if constexpr (DIM == K_SHELL) shell_geo(m, e, m.xlast, geo);
  else {
#pragma unroll
    for (int q = 0; q < D::M; ++q) dphi[q] = m.dphi[q * ld + e];
  }
#pragma unroll
  for (int n = 0; n < D::N; ++n) {
    double t[D::M], a[NPE][SD];
#pragma unroll
    for (int q = 0; q < D::M; ++q) t[q] = st.dSdF[(n * D::M + q) * ld + e];
    if constexpr (DIM == K_SHELL) {
      shell_JT(geo, t, a[1], a[2]);
#pragma unroll
      for (int r = 0; r < 3; ++r) a[0][r] = -(a[1][r] + a[2][r]);
    } else {
#pragma unroll
      for (int r = 0; r < SD; ++r) a[0][r] = 0.0;
#pragma unroll
      for (int k = 0; k < SD; ++k)
#pragma unroll
        for (int r = 0; r < SD; ++r) {
          double v = 0.0;
#pragma unroll
          for (int c = 0; c < SD; ++c) v += dphi[SD * k + c] * t[SD * c + r];
          a[k + 1][r] = v; a[0][r] -= v;
        }
    }
#pragma unroll
    for (int j = 0; j < NPE; ++j)
#pragma unroll
      for (int r = 0; r < SD; ++r) {
        const int64_t i = (((int64_t)e * D::N + n) * NPE + j) * SD + r;
        rows[i] = f[j] >= 0 ? SD * f[j] + r : -1;
        cols[i] = D::N * e + n;
        vals[i] = f[j] >= 0 ? -vol * a[j][r] : 0.0;
      }
  }
}

// -----------------------------------------------------------------------------------------------------------------
// k_update_tri2d : MixedStretch<2>::update() for Tri2DMesh elements, same fusion as k_update_tet (N=3, M=4, 6x6 blocks).
// The reference's 2D "volume" is the SIGNED double area (tri2d_mesh.cpp:32-34), so vol multiplies at the end and no
// square root of it is taken.
// -----------------------------------------------------------------------------------------------------------------
template <bool FUSE_RHS, bool ASSEMBLE>
__global__ void __launch_bounds__(128) k_update_tri2d(DevMesh m, DevState st, const double* __restrict__ x, double h2) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m.nE) return;
  const int64_t ld = m.ld;
  int status = 0;
  double s[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) s[k] = st.s[k * ld + e];
  const double vol = m.vol[e];
  const MaterialDev mat = m.mats[m.matid ? m.matid[e] : 0];
  double g[3], Hf[6], Hinv[6];
  int sweeps;
  material_grad_hess3(mat.model, mat.mu, mat.la, s, h2, g, Hf);
  const double lmax = psd_fix_invert<3>(Hf, Hinv, 1e-8, &sweeps);
  st.evals[e] = h2 * lmax;
  if (!(lmax == lmax) || lmax > 1e300) status |= ST_NONFINITE;
  const double iv = 1.0 / vol;
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j <= i; ++j) {
      const double w = ((i < 2) ? 1.0 : 0.5) * ((j < 2) ? 1.0 : 0.5);
      st.H[symidx(i, j) * ld + e] = iv * w * Hf[symidx(i, j)];
      st.Hinv[symidx(i, j) * ld + e] = Hinv[symidx(i, j)];
    }
#pragma unroll
  for (int k = 0; k < 3; ++k) st.g[k * ld + e] = g[k];

  double dphi[4], F[4], R[4], S[3], Js[3][4];
  elem_F<2>(m, e, x, dphi, F);
  polar2(F, R, S, Js);
#pragma unroll
  for (int i = 0; i < 4; ++i) st.R[i * ld + e] = R[i];
  st.iters[e] = 0;
#pragma unroll
  for (int k = 0; k < 3; ++k) {
    st.S[k * ld + e] = S[k];
    st.grad_la[k * ld + e] = ((k < 2) ? 1.0 : 2.0) * (S[k] - s[k]) * vol;
  }
  double u[3];
  if (FUSE_RHS) {
#pragma unroll
    for (int i = 0; i < 3; ++i) {
      double a = g[i];
#pragma unroll
      for (int j = 0; j < 3; ++j) a += Hf[symidx(i, j)] * (S[j] - s[j]);
      u[i] = a;
      st.gl[i * ld + e] = ((i < 2) ? 1.0 : 0.5) * a;
    }
  }
  double B[4][3];   // rows (k-1)*2 + r for vertices k = 1..2
#pragma unroll
  for (int n = 0; n < 3; ++n) {
    const double sy = (n < 2) ? 1.0 : 2.0;
#pragma unroll
    for (int k = 0; k < 4; ++k) st.dSdF[(n * 4 + k) * ld + e] = sy * Js[n][k];
#pragma unroll
    for (int k = 0; k < 2; ++k)
#pragma unroll
      for (int r = 0; r < 2; ++r) B[2 * k + r][n] = dphi[2 * k] * Js[n][r] + dphi[2 * k + 1] * Js[n][2 + r];
  }
  if (!(B[0][0] == B[0][0])) status |= ST_NONFINITE;
  int f[3];
#pragma unroll
  for (int k = 0; k < 3; ++k) f[k] = m.fidx[k * ld + e];
  {
    double tl[3];
#pragma unroll
    for (int n = 0; n < 3; ++n) tl[n] = ((n < 2) ? 1.0 : 2.0) * st.la[n * ld + e];
    double gx0[2] = {0, 0}, rh0[2] = {0, 0};
#pragma unroll
    for (int k = 0; k < 2; ++k)
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        double a = 0.0, b = 0.0;
#pragma unroll
        for (int n = 0; n < 3; ++n) { a += B[2 * k + r][n] * tl[n]; if (FUSE_RHS) b += B[2 * k + r][n] * u[n]; }
        a *= vol; b *= -vol;
        gx0[r] -= a; rh0[r] -= b;
        if (f[k + 1] >= 0) {
          red_add(&st.grad_x[2 * (int64_t)f[k + 1] + r], a);
          if (FUSE_RHS) red_add(&st.rhs[2 * (int64_t)f[k + 1] + r], b);
        }
      }
    if (f[0] >= 0) {
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        red_add(&st.grad_x[2 * (int64_t)f[0] + r], gx0[r]);
        if (FUSE_RHS) red_add(&st.rhs[2 * (int64_t)f[0] + r], rh0[r]);
      }
    }
  }
  if (ASSEMBLE) {
    // W = vol * B * Hfix (4x3, plus vertex-0 rows), block(j,k) = W_j B_k^T
    double Bf[6][3], W[6][3];
#pragma unroll
    for (int n = 0; n < 3; ++n) {
      Bf[0][n] = -(B[0][n] + B[2][n]); Bf[1][n] = -(B[1][n] + B[3][n]);
#pragma unroll
      for (int q = 0; q < 4; ++q) Bf[2 + q][n] = B[q][n];
    }
#pragma unroll
    for (int q = 0; q < 6; ++q)
#pragma unroll
      for (int n = 0; n < 3; ++n) {
        double a = 0.0;
#pragma unroll
        for (int i = 0; i < 3; ++i) a += Bf[q][i] * Hf[symidx(i, n)];
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
        for (int r = 0; r < 2; ++r)
#pragma unroll
          for (int c = 0; c < 2; ++c) {
            double a = 0.0;
#pragma unroll
            for (int n = 0; n < 3; ++n) a += W[2 * j + r][n] * Bf[2 * k + c][n];
            red_add(&st.Avals[(int64_t)idx + (int64_t)r * stride + c], a);
          }
      }
    }
  }
  count_status(st.counters, status, 0, sweeps);
}

// ---- layout conversions (getters / setters; not on the hot path) ----
static __global__ void k_soa_to_aos(const double* __restrict__ soa, double* __restrict__ aos, int ncomp, int nE, int64_t ld) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)nE * ncomp) return;
  const int e = (int)(i / ncomp), c = (int)(i % ncomp);
  aos[i] = soa[c * ld + e];
}
static __global__ void k_aos_to_soa(const double* __restrict__ aos, double* __restrict__ soa, int ncomp, int nE, int64_t ld) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)nE * ncomp) return;
  const int e = (int)(i / ncomp), c = (int)(i % ncomp);
  soa[c * ld + e] = aos[i];
}
// packed symmetric N x N (component-major) -> full column-major N x N per element; mode 0 plain, 1 Hloc scaling
static __global__ void k_sym_to_full(const double* __restrict__ soa, const double* __restrict__ vol, double* __restrict__ aos,
                              int N, int dim, int nE, int64_t ld, int mode) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)nE * N * N) return;
  const int e = (int)(i / (N * N)), q = (int)(i % (N * N));
  const int r = q % N, c = q / N;
  double v = soa[(int64_t)symidx(r, c) * ld + e];
  if (mode == 1) { const double w = ((r < dim) ? 1.0 : 2.0) * ((c < dim) ? 1.0 : 2.0); v *= vol[e] * vol[e] * w; }
  aos[i] = v;
}
static __global__ void k_grad_mixed(const double* __restrict__ g, const double* __restrict__ vol, double* __restrict__ aos, int N, int nE, int64_t ld) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)nE * N) return;
  const int e = (int)(i / N), c = (int)(i % N);
  aos[i] = vol[e] * g[c * ld + e];     // grad_ = vol * g_   (mixed_stretch.cpp:237-242)
}
static __global__ void k_int_to_double(const int* __restrict__ a, double* __restrict__ o, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) o[i] = (double)a[i];
}

}  // namespace mfb
