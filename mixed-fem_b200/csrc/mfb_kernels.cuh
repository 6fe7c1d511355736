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

}  // namespace mfb
