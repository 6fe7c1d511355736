// Linear-solver side of the path: y = (M + A) x on the assembled CSR (block-structured: every block-row has d scalar
// rows of equal length, sparse_utils.cpp:101-116) and a preconditioned conjugate-gradient loop that restates Eigen's
// ConjugateGradient recurrence used by EigenIterativeSolver (src/linear_solvers/eigen_iterative_solver.h:27-65,
// registered with Diagonal / BlockJacobi preconditioners in src/factories/linear_solver_factory.cpp:107-173; the same
// recurrence is hand-written in src/linear_solvers/pcg.h:11-80).
//
// SpMV is HBM-bound: 8 B per stored value (+8 B when M is given) + 4/d^2 B of column index per value (one block-column
// index per d x d block, not one per scalar entry) + the vectors.  One warp owns one block-row: its d*d*nbr values
// are contiguous in the CSR value array, so the warp streams them with coalesced 256-byte reads.
#include "mfb_ctx.cuh"

using namespace mfb;

#define CHECK_CTX(c) do { if (!(c)) { g_last_error = "null context"; return MFB_ERR_INVALID; } cudaSetDevice((c)->device); } while (0)
static inline unsigned nblk(int64_t n, int t = 256) { return (unsigned)((n + t - 1) / t); }

namespace {

template <int DIM, bool HAS_M>
__global__ void __launch_bounds__(256) k_spmv(const int* __restrict__ blockptr, const int* __restrict__ bcol, int nF,
                                              const double* __restrict__ A, const double* __restrict__ Mv,
                                              const double* __restrict__ x, double* __restrict__ y) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int lane = threadIdx.x & 31;
  if (warp >= nF) return;
  const int b0 = blockptr[warp], nbr = blockptr[warp + 1] - b0;
  const int rowlen = DIM * nbr;
  const int64_t base = (int64_t)DIM * DIM * b0;
  double acc[DIM];
#pragma unroll
  for (int r = 0; r < DIM; ++r) acc[r] = 0.0;
  for (int q = lane; q < rowlen; q += 32) {     // q = position inside one scalar row
    const int col = DIM * bcol[b0 + q / DIM] + (q % DIM);
    const double xv = __ldg(&x[col]);
#pragma unroll
    for (int r = 0; r < DIM; ++r) {
      double a = A[base + (int64_t)r * rowlen + q];
      if (HAS_M) a += Mv[base + (int64_t)r * rowlen + q];
      acc[r] += a * xv;
    }
  }
#pragma unroll
  for (int r = 0; r < DIM; ++r) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) acc[r] += __shfl_down_sync(0xffffffffu, acc[r], o);
  }
  if (lane == 0) {
#pragma unroll
    for (int r = 0; r < DIM; ++r) y[DIM * warp + r] = acc[r];
  }
}

// inverse of the d x d diagonal blocks (mode 2) or of the scalar diagonal (mode 1); dinv holds d*d per block-row
template <int DIM>
__global__ void k_build_precond(const int* __restrict__ blockptr, const int* __restrict__ diagblk, int nF, const double* __restrict__ A,
                                const double* __restrict__ Mv, int mode, double* __restrict__ dinv) {
  const int gi = blockIdx.x * blockDim.x + threadIdx.x;
  if (gi >= nF) return;
  const int b0 = blockptr[gi], nbr = blockptr[gi + 1] - b0, rank = diagblk[gi] - b0;
  const int64_t base = (int64_t)DIM * DIM * b0;
  double D[DIM][DIM];
#pragma unroll
  for (int r = 0; r < DIM; ++r)
#pragma unroll
    for (int c = 0; c < DIM; ++c) {
      const int64_t i = base + (int64_t)r * DIM * nbr + DIM * rank + c;
      D[r][c] = A[i] + (Mv ? Mv[i] : 0.0);
    }
  double I[DIM][DIM];
  if (mode == 1) {
#pragma unroll
    for (int r = 0; r < DIM; ++r)
#pragma unroll
      for (int c = 0; c < DIM; ++c) I[r][c] = (r == c) ? ((D[r][r] != 0.0) ? 1.0 / D[r][r] : 1.0) : 0.0;   // Eigen: zero diagonal -> 1
  } else if constexpr (DIM == 2) {
    const double id = 1.0 / (D[0][0] * D[1][1] - D[0][1] * D[1][0]);
    I[0][0] = D[1][1] * id; I[0][1] = -D[0][1] * id; I[1][0] = -D[1][0] * id; I[1][1] = D[0][0] * id;
  } else {
    const double c00 = D[1][1] * D[2][2] - D[1][2] * D[2][1];
    const double c01 = D[1][2] * D[2][0] - D[1][0] * D[2][2];
    const double c02 = D[1][0] * D[2][1] - D[1][1] * D[2][0];
    const double id = 1.0 / (D[0][0] * c00 + D[0][1] * c01 + D[0][2] * c02);
    I[0][0] = c00 * id;
    I[0][1] = (D[0][2] * D[2][1] - D[0][1] * D[2][2]) * id;
    I[0][2] = (D[0][1] * D[1][2] - D[0][2] * D[1][1]) * id;
    I[1][0] = c01 * id;
    I[1][1] = (D[0][0] * D[2][2] - D[0][2] * D[2][0]) * id;
    I[1][2] = (D[0][2] * D[1][0] - D[0][0] * D[1][2]) * id;
    I[2][0] = c02 * id;
    I[2][1] = (D[0][1] * D[2][0] - D[0][0] * D[2][1]) * id;
    I[2][2] = (D[0][0] * D[1][1] - D[0][1] * D[1][0]) * id;
  }
#pragma unroll
  for (int r = 0; r < DIM; ++r)
#pragma unroll
    for (int c = 0; c < DIM; ++c) dinv[(int64_t)gi * DIM * DIM + r * DIM + c] = I[r][c];
}

// deterministic two-stage dot products: up to 2 dots per pass
__global__ void __launch_bounds__(256) k_dot2(const double* __restrict__ a, const double* __restrict__ b, const double* __restrict__ c2,
                                              const double* __restrict__ d2, int n, double* partials, int np) {
  __shared__ double sh[2][8];
  double s0 = 0.0, s1 = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    s0 += a[i] * b[i];
    if (c2) s1 += c2[i] * d2[i];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { s0 += __shfl_down_sync(0xffffffffu, s0, o); s1 += __shfl_down_sync(0xffffffffu, s1, o); }
  if ((threadIdx.x & 31) == 0) { sh[0][threadIdx.x >> 5] = s0; sh[1][threadIdx.x >> 5] = s1; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double r0 = 0.0, r1 = 0.0;
    for (int i = 0; i < 8; ++i) { r0 += sh[0][i]; r1 += sh[1][i]; }
    partials[blockIdx.x] = r0; partials[np + blockIdx.x] = r1;
  }
}
__global__ void k_finish2(const double* __restrict__ partials, int np, double* out0, double* out1) {
  __shared__ double sh[2][8];
  double s0 = 0.0, s1 = 0.0;
  for (int i = threadIdx.x; i < np; i += blockDim.x) { s0 += partials[i]; s1 += partials[np + i]; }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) { s0 += __shfl_down_sync(0xffffffffu, s0, o); s1 += __shfl_down_sync(0xffffffffu, s1, o); }
  if ((threadIdx.x & 31) == 0) { sh[0][threadIdx.x >> 5] = s0; sh[1][threadIdx.x >> 5] = s1; }
  __syncthreads();
  if (threadIdx.x == 0) {
    double r0 = 0.0, r1 = 0.0;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) { r0 += sh[0][i]; r1 += sh[1][i]; }
    *out0 = r0; if (out1) *out1 = r1;
  }
}
// r = b - t
__global__ void k_residual(const double* __restrict__ b, const double* __restrict__ t, double* __restrict__ r, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) r[i] = b[i] - t[i];
}
// z = Dinv r (block-diagonal d x d)
template <int DIM>
__global__ void k_apply_precond(const double* __restrict__ dinv, const double* __restrict__ r, double* __restrict__ z, int nF, int mode) {
  const int gi = blockIdx.x * blockDim.x + threadIdx.x;
  if (gi >= nF) return;
  if (mode == 0) {
#pragma unroll
    for (int k = 0; k < DIM; ++k) z[DIM * gi + k] = r[DIM * gi + k];
    return;
  }
  double rv[DIM];
#pragma unroll
  for (int k = 0; k < DIM; ++k) rv[k] = r[DIM * gi + k];
#pragma unroll
  for (int a = 0; a < DIM; ++a) {
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < DIM; ++k) s += dinv[(int64_t)gi * DIM * DIM + a * DIM + k] * rv[k];
    z[DIM * gi + a] = s;
  }
}
// scal[0]=absNew, scal[1]=pAp  ->  alpha ; x += alpha p ; r -= alpha t
__global__ void k_cg_update(const double* __restrict__ scal, const double* __restrict__ p, const double* __restrict__ t,
                            double* __restrict__ x, double* __restrict__ r, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double alpha = scal[0] / scal[1];
  x[i] += alpha * p[i];
  r[i] -= alpha * t[i];
}
// scal[0]=absOld, scal[2]=absNew -> beta ; p = z + beta p
__global__ void k_cg_dir(const double* __restrict__ scal, const double* __restrict__ z, double* __restrict__ p, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const double beta = scal[2] / scal[0];
  p[i] = z[i] + beta * p[i];
}
__global__ void k_copy_scalar(double* scal, int dst, int src) { scal[dst] = scal[src]; }

int spmv(mfb_ctx* c, const double* Mv, const double* x, double* y) {
  const int64_t threads = (int64_t)c->nF * 32;
  if (c->nF == 0) return MFB_OK;
  if (c->dim == 3) {
    if (Mv) k_spmv<3, true><<<nblk(threads), 256, 0, c->stream>>>(c->blockptr.p, c->bcol.p, c->nF, c->Avals.p, Mv, x, y);
    else k_spmv<3, false><<<nblk(threads), 256, 0, c->stream>>>(c->blockptr.p, c->bcol.p, c->nF, c->Avals.p, Mv, x, y);
  } else {
    if (Mv) k_spmv<2, true><<<nblk(threads), 256, 0, c->stream>>>(c->blockptr.p, c->bcol.p, c->nF, c->Avals.p, Mv, x, y);
    else k_spmv<2, false><<<nblk(threads), 256, 0, c->stream>>>(c->blockptr.p, c->bcol.p, c->nF, c->Avals.p, Mv, x, y);
  }
  c->launches += 1;
  MFB_CUDA(cudaGetLastError());
  return MFB_OK;
}

const int kDotBlocks = 592;   // 4 x 148 SMs

int dot2(mfb_ctx* c, const double* a, const double* b, const double* c2, const double* d2, int n, int out0, int out1) {
  k_dot2<<<kDotBlocks, 256, 0, c->stream>>>(a, b, c2, d2, n, c->partials.p, kDotBlocks);
  k_finish2<<<1, 256, 0, c->stream>>>(c->partials.p, kDotBlocks, c->scal.p + out0, out1 >= 0 ? c->scal.p + out1 : nullptr);
  c->launches += 2;
  MFB_CUDA(cudaGetLastError());
  return MFB_OK;
}

}  // namespace

extern "C" int mfb_spmv_dev(mfb_ctx* c, const double* M_vals_dev, const double* x_dev, double* y_dev) {
  CHECK_CTX(c);
  if (!x_dev || !y_dev) return MFB_ERR_INVALID;
  return spmv(c, M_vals_dev, x_dev, y_dev);
}

extern "C" int mfb_pcg_dev(mfb_ctx* c, const double* Mv, const double* b, double* x, int precond, double tol, int max_iters,
                           int* iters_out, double* relres_out) {
  CHECK_CTX(c);
  if (!b || !x || precond < 0 || precond > 2) return MFB_ERR_INVALID;
  const int n = c->dim * c->nF;
  const int dd = c->dim * c->dim;
  int rc;
  if (c->pcg_r.n < (size_t)n + 1) {
    if ((rc = c->pcg_r.alloc((size_t)n + 1)) || (rc = c->pcg_p.alloc((size_t)n + 1)) || (rc = c->pcg_z.alloc((size_t)n + 1)) ||
        (rc = c->pcg_t.alloc((size_t)n + 1)) || (rc = c->pcg_dinv.alloc((size_t)c->nF * dd + 1)))
      return rc;
  }
  if (c->partials.n < (size_t)2 * kDotBlocks) { if ((rc = c->partials.alloc((size_t)2 * kDotBlocks + c->grid()))) return rc; }
  cudaStream_t st = c->stream;
  double *r = c->pcg_r.p, *p = c->pcg_p.p, *z = c->pcg_z.p, *t = c->pcg_t.p;
  if (precond) {
    if (c->dim == 3) k_build_precond<3><<<nblk(c->nF), 256, 0, st>>>(c->blockptr.p, c->diagblk.p, c->nF, c->Avals.p, Mv, precond, c->pcg_dinv.p);
    else k_build_precond<2><<<nblk(c->nF), 256, 0, st>>>(c->blockptr.p, c->diagblk.p, c->nF, c->Avals.p, Mv, precond, c->pcg_dinv.p);
    c->launches += 1;
  }
  auto apply_prec = [&](const double* rr, double* zz) {
    if (c->dim == 3) k_apply_precond<3><<<nblk(c->nF), 256, 0, st>>>(c->pcg_dinv.p, rr, zz, c->nF, precond);
    else k_apply_precond<2><<<nblk(c->nF), 256, 0, st>>>(c->pcg_dinv.p, rr, zz, c->nF, precond);
    c->launches += 1;
  };
  // residual = b - A x ; rhsNorm2
  if ((rc = spmv(c, Mv, x, t))) return rc;
  k_residual<<<nblk(n), 256, 0, st>>>(b, t, r, n); c->launches += 1;
  if ((rc = dot2(c, b, b, r, r, n, 4, 5))) return rc;                 // scal[4] = |b|^2, scal[5] = |r|^2
  MFB_CUDA(cudaMemcpyAsync(c->h_scal, c->scal.p, 8 * sizeof(double), cudaMemcpyDeviceToHost, st));
  MFB_CUDA(cudaStreamSynchronize(st));
  const double rhsNorm2 = c->h_scal[4];
  double resNorm2 = c->h_scal[5];
  int it = 0;
  if (rhsNorm2 == 0.0) {
    MFB_CUDA(cudaMemsetAsync(x, 0, (size_t)n * 8, st));
    if (iters_out) *iters_out = 0; if (relres_out) *relres_out = 0.0;
    return MFB_OK;
  }
  const double threshold = tol * tol * rhsNorm2;
  if (resNorm2 >= threshold) {
    apply_prec(r, p);
    if ((rc = dot2(c, r, p, nullptr, nullptr, n, 0, -1))) return rc;  // scal[0] = absNew
    while (it < max_iters) {
      if ((rc = spmv(c, Mv, p, t))) return rc;
      if ((rc = dot2(c, p, t, nullptr, nullptr, n, 1, -1))) return rc;     // scal[1] = p.Ap
      k_cg_update<<<nblk(n), 256, 0, st>>>(c->scal.p, p, t, x, r, n); c->launches += 1;
      apply_prec(r, z);
      if ((rc = dot2(c, r, r, r, z, n, 5, 2))) return rc;               // scal[5] = |r|^2, scal[2] = r.z
      MFB_CUDA(cudaMemcpyAsync(c->h_scal, c->scal.p, 8 * sizeof(double), cudaMemcpyDeviceToHost, st));
      MFB_CUDA(cudaStreamSynchronize(st));
      resNorm2 = c->h_scal[5];
      if (resNorm2 < threshold) break;
      k_cg_dir<<<nblk(n), 256, 0, st>>>(c->scal.p, z, p, n);
      k_copy_scalar<<<1, 1, 0, st>>>(c->scal.p, 0, 2);
      c->launches += 2;
      ++it;
    }
  }
  MFB_CUDA(cudaGetLastError());
  if (iters_out) *iters_out = it;
  if (relres_out) *relres_out = sqrt(resNorm2 / rhsNorm2);
  return MFB_OK;
}
