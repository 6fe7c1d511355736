// CUDA kernels of the mixed-stretch hot path for sm_100a (FP64, one thread per element, structure-of-arrays state).
//
// Data layout in HBM: every per-element quantity is stored component-major, component c of element e at
// base[c*ld + e] with ld = nE rounded up to 32, so that a warp's 32 threads read/write 256 contiguous bytes per
// component (fully coalesced, 128-byte aligned rows).  Vertex vectors keep the reference layout x[d*v + l].
// The assembled matrix keeps the reference's scalar CSR value order (src/utils/sparse_utils.cpp:155-163).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "mfb_math.cuh"

namespace mfb {

struct MaterialDev { int model; double mu, la; };

struct DevMesh {
  int dim, npe, nV, nE, nF;
  int64_t ld;
  const int* T;          // [npe][ld]   vertex ids
  const int* fidx;       // [npe][ld]   free index of each local vertex (-1 = pinned)
  const double* dphi;    // [d*d][ld]   rows 1..d of dphi/dX: component (k-1)*d + c
  const double* vol;     // [ld]
  const int* matid;      // [ld] or nullptr
  const MaterialDev* mats;
  int exp;               // experiment bits (what-if runs, option "exp"): 1 skip dSdF stores, 2 skip REDs, 4 skip H/Hinv stores, 8 skip lambda_max
  MaterialDev mat0;      // materials[0] by value: lives in the kernel-parameter constant bank (no load when matid == nullptr)
  const int* smap;       // [npe*npe][ld]  value index of entry (0,0) of block (j,k), -1 if pinned
  const int* sstride;    // [npe][ld]      CSR row length of block-row j
  const double* nrest;   // [3][ld]        shells: rest normal of the face (TriMesh::N_), else nullptr
  const double* xlast;   // shells: positions of the last update (their local Jacobian depends on x), else nullptr
};

struct DevState {
  double *R, *s, *la, *ds, *S, *dSdF, *g, *H, *Hinv, *evals, *gl, *grad_la;
  double* Lfac;          // [21][ld] scratch: Cholesky factor of the PSD-fixed Hessian (k_material_psd6 -> k_update_tet)
  int* iters;
  double *grad_x, *rhs, *Avals;
  unsigned long long* counters;   // [5]: maxiter, hfail, nonfinite, polar iters, eig sweeps
};

__device__ __forceinline__ void red_add(double* addr, double v) { atomicAdd(addr, v); }

__device__ __forceinline__ void count_status(unsigned long long* counters, int status, int iters, int sweeps) {
  const unsigned full = 0xffffffffu;
  const unsigned mask = __activemask();
  (void)full;
  int c0 = (status & ST_POLAR_MAXITER) ? 1 : 0, c1 = (status & ST_POLAR_HFAIL) ? 1 : 0, c2 = (status & ST_NONFINITE) ? 1 : 0;
  c0 = __reduce_add_sync(mask, c0); c1 = __reduce_add_sync(mask, c1); c2 = __reduce_add_sync(mask, c2);
  const int it = __reduce_add_sync(mask, iters), sw = __reduce_add_sync(mask, sweeps);
  const int leader = __ffs(mask) - 1;
  if ((threadIdx.x & 31) == leader) {
    if (c0) atomicAdd(&counters[0], (unsigned long long)c0);
    if (c1) atomicAdd(&counters[1], (unsigned long long)c1);
    if (c2) atomicAdd(&counters[2], (unsigned long long)c2);
    atomicAdd(&counters[3], (unsigned long long)it);
    atomicAdd(&counters[4], (unsigned long long)sw);
  }
}

struct M33 { double a[9]; };
struct V6 { double a[6]; };

// evaluate_constraint()'s polar from a COPY of R_ (mixed_stretch.cpp:358-360), only reached when the freshly formed
// R F^T does not pass the tolerance test (never in practice after update_rotations converged): kept out of line.
static __device__ __noinline__ V6 polar_from_copy_slow(M33 R, M33 F) {
  double Y[9]; int st = 0;
  procrustes3(R.a, F.a, Y, 1e-6, 100, &st);
  V6 s; stretch6(R.a, F.a, s.a);
  return s;
}

// F = sum_k x_k (x) dphi_k  with dphi_0 = -(dphi_1+dphi_2+dphi_3)  (tet_mesh.cpp:17-27, linear_tet_dphi_dX.cpp:5-17)
__device__ __forceinline__ void tet_F(const DevMesh& m, int e, const double* __restrict__ x, double* dphi, double* F) {
  const int64_t ld = m.ld;
  int v[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) v[k] = __ldcs(&m.T[k * ld + e]);
#pragma unroll
  for (int i = 0; i < 9; ++i) dphi[i] = __ldcs(&m.dphi[i * ld + e]);
  double x0[3], dx[9];
#pragma unroll
  for (int l = 0; l < 3; ++l) x0[l] = __ldg(&x[3 * (int64_t)v[0] + l]);
#pragma unroll
  for (int k = 0; k < 3; ++k)
#pragma unroll
    for (int l = 0; l < 3; ++l) dx[3 * k + l] = __ldg(&x[3 * (int64_t)v[k + 1] + l]) - x0[l];
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int r = 0; r < 3; ++r) F[3 * c + r] = dx[r] * dphi[c] + dx[3 + r] * dphi[3 + c] + dx[6 + r] * dphi[6 + c];
}

// in-place Cholesky of a packed-lower SPD 6x6: H = L L^T
__device__ __forceinline__ void cholesky6(double* H) {
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    double d = H[symidx(j, j)];
#pragma unroll
    for (int k = 0; k < j; ++k) d -= H[symidx(j, k)] * H[symidx(j, k)];
    d = sqrt(d);
    H[symidx(j, j)] = d;
    const double id = 1.0 / d;
#pragma unroll
    for (int i = j + 1; i < 6; ++i) {
      double a = H[symidx(i, j)];
#pragma unroll
      for (int k = 0; k < j; ++k) a -= H[symidx(i, k)] * H[symidx(j, k)];
      H[symidx(i, j)] = a * id;
    }
  }
}

// -----------------------------------------------------------------------------------------------------------------
// k_update_tet : MixedStretch<3>::update() in ONE pass per element (mixed_stretch.cpp:158-243, 345-366):
//   A. g = h^2 dpsi, H = h^2 d2psi, psd_fix_invert, evals_, H_ = Syminv H Syminv / vol, Hinv_        (:193-204)
//   B. F, warm-started polar, R_, S_; grad_la_ = Sym (S - s) vol                                      (:166-185, :345-366)
//   C. dSdF_ = Js^T Sym ; B = (Js Jloc)^T                                                             (:182)
//   D. grad_x_ += vol B Sym la ; [fused rhs: gl_, rhs_ -= vol B (H (S-s) + g)]                        (:224-232, :245-275)
//      A += vol B Hfix B^T scattered through the element->nonzero map                                 (:209-220)
// Aloc_ is never materialised: with Hfix = L L^T the 12x12 block is vol (B L)(B L)^T, formed 3x3 block by block and
// sent to the CSR values with FP64 reductions (RED.ADD.F64).
// -----------------------------------------------------------------------------------------------------------------
// Phase A as its own kernel (k_material_psd6): g_ = h^2 dpsi, H = h^2 d2psi, psd_fix_invert, evals_, H_, Hinv_ and the
// Cholesky factor of the fixed Hessian (scratch, read back by k_update_tet).  Splitting it off keeps each kernel's
// SASS inside the SM instruction cache and lets this latency-bound Jacobi run at 3 blocks/SM instead of 2.
__device__ __forceinline__ void material_psd6_body(const DevMesh& m, const DevState& st, int e, double h2) {
  const int64_t ld = m.ld;
  double s[6];
#pragma unroll
  for (int k = 0; k < 6; ++k) s[k] = st.s[k * ld + e];
  const double vol = m.vol[e];
  const MaterialDev mat = m.mats[m.matid ? m.matid[e] : 0];
  double g[6], L[21], Hinv[21];
  int sweeps;
  material_grad_hess6(mat.model, mat.mu, mat.la, s, h2, g, L);
#pragma unroll
  for (int k = 0; k < 6; ++k) st.g[k * ld + e] = g[k];
  const double lmax = psd_fix_invert<6>(L, Hinv, 1e-8, &sweeps);
  st.evals[e] = h2 * lmax;
  const double iv = 1.0 / vol;
#pragma unroll
  for (int i = 0; i < 6; ++i)
#pragma unroll
    for (int j = 0; j <= i; ++j) {
      const double w = ((i < 3) ? 1.0 : 0.5) * ((j < 3) ? 1.0 : 0.5);
      st.H[symidx(i, j) * ld + e] = iv * w * L[symidx(i, j)];
      st.Hinv[symidx(i, j) * ld + e] = Hinv[symidx(i, j)];
    }
  cholesky6(L);
#pragma unroll
  for (int i = 0; i < 21; ++i) st.Lfac[i * ld + e] = L[i];
  const int bad = (!(lmax == lmax) || lmax > 1e300) ? ST_NONFINITE : 0;
  count_status(st.counters, bad, 0, sweeps);
}
static __global__ void __launch_bounds__(128, 3) k_material_psd6(DevMesh m, DevState st, double h2) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m.nE) return;
  material_psd6_body(m, st, e, h2);
}

// Fill the strict lower triangle of the assembled matrix from the accumulated upper triangle: block (gi,gj), gi > gj,
// is the transpose of block (gj,gi) (index bT[b]); inside diagonal blocks entry (r,c), r > c, copies (c,r).
// One thread per scalar entry; A is symmetric by construction (Aloc = vol^2 Jloc^T K Jloc, mixed_stretch.cpp:212).
static __global__ void __launch_bounds__(256) k_mirror_lower(const int4* __restrict__ list, int n, int dim, double* __restrict__ A) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int dd = dim * dim;
  if (i >= (int64_t)n * dd) return;
  const int q = (int)(i / dd), rc = (int)(i % dd), r = rc / dim, c = rc % dim;
  const int4 v = __ldg(&list[q]);        // (dst0, dst row stride, src0, src row stride)
  if (v.x == v.z && r <= c) return;      // diagonal block: only its strict lower half is copied
  A[(int64_t)v.x + (int64_t)r * v.y + c] = A[(int64_t)v.z + (int64_t)c * v.w + r];
}

#ifndef MFB_K2_MINBLOCKS
#define MFB_K2_MINBLOCKS 2
#endif
template <bool FUSE_RHS, bool ASSEMBLE>
__global__ void __launch_bounds__(128, MFB_K2_MINBLOCKS) k_update_tet(DevMesh m, DevState st, const double* __restrict__ x, double h2) {
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m.nE) return;
  const int64_t ld = m.ld;
  int status = 0;

  // ---- A (done by k_material_psd6): pick up s, g and the Cholesky factor of the fixed Hessian
  double s[6], g[6], L[21];
#pragma unroll
  for (int k = 0; k < 6; ++k) { s[k] = st.s[k * ld + e]; g[k] = st.g[k * ld + e]; }
#pragma unroll
  for (int i = 0; i < 21; ++i) L[i] = st.Lfac[i * ld + e];
  const double vol = m.vol[e];
  const int sweeps = 0;
  // scatter indices are fetched up front so their latency hides behind the polar / dSdF work
  int f[4], stride[4], sidx[16];
#pragma unroll
  for (int k = 0; k < 4; ++k) { f[k] = m.fidx[k * ld + e]; stride[k] = m.sstride[k * ld + e]; }
#pragma unroll
  for (int q = 0; q < 16; ++q) sidx[q] = m.smap[q * ld + e];
  double la6[6];
#pragma unroll
  for (int n = 0; n < 6; ++n) la6[n] = st.la[n * ld + e];

  // ---- B
  double dphi[9], F[9], R[9], Y[9];
  tet_F(m, e, x, dphi, F);
#pragma unroll
  for (int i = 0; i < 9; ++i) R[i] = st.R[i * ld + e];
  const int iters = procrustes3(R, F, Y, 1e-6, 100, &status);
#pragma unroll
  for (int i = 0; i < 9; ++i) st.R[i * ld + e] = R[i];
  st.iters[e] = iters;
  double S[6];
  stretch6(R, F, S);
#pragma unroll
  for (int k = 0; k < 6; ++k) st.S[k * ld + e] = S[k];
  {
    // evaluate_constraint: polar again from a copy of R_ ; converged => zero iterations => same S
    double Sc[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) Sc[k] = S[k];
    double Y2[9];
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
      for (int r = 0; r < 3; ++r) Y2[3 * c + r] = R[r] * F[c] + R[3 + r] * F[3 + c] + R[6 + r] * F[6 + c];
    const double q0 = Y2[5] - Y2[7], q1 = Y2[6] - Y2[2], q2 = Y2[1] - Y2[3];
    if (!(sqrt(q0 * q0 + q1 * q1 + q2 * q2) < 1e-6)) {
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
  // u = Hfix (S - s) + g ; gl = Syminv u   (rhs(): gl = vol H_ Sym (S-s) + Syminv g)
  double u[6];
  if (FUSE_RHS) {
    double d[6], t[6];
#pragma unroll
    for (int k = 0; k < 6; ++k) d[k] = S[k] - s[k];
#pragma unroll
    for (int j = 0; j < 6; ++j) { double a = 0.0;
#pragma unroll
      for (int i = j; i < 6; ++i) a += L[symidx(i, j)] * d[i];
      t[j] = a; }
#pragma unroll
    for (int i = 0; i < 6; ++i) { double a = g[i];
#pragma unroll
      for (int j = 0; j <= i; ++j) a += L[symidx(i, j)] * t[j];
      u[i] = a; st.gl[i * ld + e] = ((i < 3) ? 1.0 : 0.5) * a; }
  }

  // ---- C
  double Hinv3[6];
  if (!procrustes_hinv(Y, Hinv3)) status |= ST_POLAR_HFAIL;
  double B[9][6];   // rows (k-1)*3 + r for vertices k = 1..3
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
      for (int r = 0; r < 3; ++r) B[3 * k + r][n] = dphi[3 * k] * col[r] + dphi[3 * k + 1] * col[3 + r] + dphi[3 * k + 2] * col[6 + r];
  }
  if (!(B[0][0] == B[0][0])) status |= ST_NONFINITE;

  // ---- D : vector contributions
  {
    double tl[6];
#pragma unroll
    for (int n = 0; n < 6; ++n) tl[n] = ((n < 3) ? 1.0 : 2.0) * la6[n];
    double gx0[3] = {0, 0, 0}, rh0[3] = {0, 0, 0};
#pragma unroll
    for (int k = 0; k < 3; ++k)
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

  // ---- D : matrix.  B <- sqrt(vol) B L ; block(j,k) = B_j B_k^T
  if (ASSEMBLE) {
    const double sv = sqrt(vol);
#pragma unroll
    for (int row = 0; row < 9; ++row) {
      double t[6];
#pragma unroll
      for (int j = 0; j < 6; ++j) { double a = 0.0;
#pragma unroll
        for (int i = j; i < 6; ++i) a += B[row][i] * L[symidx(i, j)];
        t[j] = a * sv; }
#pragma unroll
      for (int j = 0; j < 6; ++j) B[row][j] = t[j];
    }
    double B0[3][6];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int n = 0; n < 6; ++n) B0[r][n] = -(B[r][n] + B[3 + r][n] + B[6 + r][n]);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
#pragma unroll
      for (int k = j; k < 4; ++k) {
        const int idx_jk = sidx[j * 4 + k];
        if (idx_jk < 0) continue;             // either vertex pinned (then (k,j) is pinned too)
        // Only the upper triangle of the GLOBAL matrix is accumulated (block (gi,gj) with gi < gj, and c >= r inside
        // diagonal blocks); k_mirror_lower fills the rest afterwards.  78 reductions per tet instead of 144.
        const bool up = (j == k) || (f[j] < f[k]);
        const int64_t base = up ? (int64_t)idx_jk : (int64_t)sidx[k * 4 + j];
        const int64_t rs = up ? stride[j] : stride[k];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            if (j == k && c < r) continue;
            double a = 0.0;
#pragma unroll
            for (int n = 0; n < 6; ++n) {
              const double bj = (j == 0) ? B0[r][n] : B[3 * (j - 1) + r][n];
              const double bk = (k == 0) ? B0[c][n] : B[3 * (k - 1) + c][n];
              a += bj * bk;
            }
            red_add(&st.Avals[up ? base + r * rs + c : base + c * rs + r], a);
          }
      }
    }
  }
  count_status(st.counters, status, iters, sweeps);
}

}  // namespace mfb
