// Linear-solver side of the path: y = (M + A) x on the assembled CSR (block-structured: every block-row has d scalar
// rows of equal length, sparse_utils.cpp:101-116) and a preconditioned conjugate-gradient loop that restates Eigen's
// ConjugateGradient recurrence used by EigenIterativeSolver (src/linear_solvers/eigen_iterative_solver.h:27-65,
// registered with Diagonal / BlockJacobi preconditioners in src/factories/linear_solver_factory.cpp:107-173; the same
// recurrence is hand-written in src/linear_solvers/pcg.h:11-80).
//
// SpMV is HBM-bound: 8 B per stored value (+8 B when M is given) + 4/d^2 B of column index per value (one block-column
// index per d x d block, not one per scalar entry) + the vectors.
//
// Two forms of the loop: three launches per iteration (k_pcg_spmv_dot, k_pcg_update, k_pcg_dir[_dist]) and, for small
// systems, ONE persistent cooperative kernel per solve (k_pcg_persist).  On several GPUs both exchange the halo rows of
// t = L p and the dot products through peer mailboxes from inside the kernels (peer_post / peer_collect below, set up by
// mfb_dist.cu); NCCL (halo_exchange / allreduce_scalars between the launches) is the fallback.
#include "mfb_ctx.cuh"
#include <cooperative_groups.h>
#include <algorithm>
#include <cmath>
#include <vector>

using namespace mfb;

#define CHECK_CTX(c) do { if (!(c)) { g_last_error = "null context"; return MFB_ERR_INVALID; } cudaSetDevice((c)->device); } while (0)
static inline unsigned nblk(int64_t n, int t = 256) { return (unsigned)((n + t - 1) / t); }

namespace {

// Half a warp per block-row, one lane per d x d block (an interior vertex of a tet mesh has ~15 blocks): the lane reads
// its d contiguous values of each of the d scalar rows, adjacent lanes adjacent 8d-byte pieces (tools/ubench_spmv.cu:
// 50 us per product at 1M tets against 64 us for a warp per block-row; a plain read of the values takes 43 us).
template <int DIM, bool HAS_M>
__global__ void __launch_bounds__(256) k_spmv(const int* __restrict__ blockptr, const int* __restrict__ bcol, int nF,
                                              const double* __restrict__ A, const double* __restrict__ Mv,
                                              const double* __restrict__ x, double* __restrict__ y) {
  const int lane = threadIdx.x & 31, hl = lane & 15;
  const int row = (blockIdx.x * 8 + (threadIdx.x >> 5)) * 2 + (lane >> 4);
  int b0 = 0, nbr = 0;
  if (row < nF) { b0 = blockptr[row]; nbr = blockptr[row + 1] - b0; }
  const int rowlen = DIM * nbr;
  const int64_t base = (int64_t)DIM * DIM * b0;
  double acc[DIM];
#pragma unroll
  for (int r = 0; r < DIM; ++r) acc[r] = 0.0;
  for (int k = hl; k < nbr; k += 16) {
    const int col = DIM * bcol[b0 + k];
    double xv[DIM];
#pragma unroll
    for (int c = 0; c < DIM; ++c) xv[c] = __ldg(&x[col + c]);
#pragma unroll
    for (int r = 0; r < DIM; ++r)
#pragma unroll
      for (int c = 0; c < DIM; ++c) {
        const int64_t i = base + (int64_t)r * rowlen + DIM * k + c;
        double a = A[i];
        if (HAS_M) a += Mv[i];
        acc[r] = fma(a, xv[c], acc[r]);
      }
  }
#pragma unroll
  for (int r = 0; r < DIM; ++r) {
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) acc[r] += __shfl_down_sync(0xffffffffu, acc[r], o, 16);
  }
  if (hl == 0 && row < nF) {
#pragma unroll
    for (int r = 0; r < DIM; ++r) y[DIM * row + r] = acc[r];
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
// ---------------------------------------------------------------------------------------------------------------
// Fused CG iteration: three launches, no host round trip.
//   k_pcg_spmv_dot : t = L p (L = M + A summed once per solve), block partials of p.t
//   k_pcg_update   : alpha = absOld / p.t ; x += alpha p ; r -= alpha t ; z = Dinv r ; block partials of r.r, r.z;
//                    |r|^2 < threshold -> done (the reference breaks before counting that iteration)
//   k_pcg_dir      : beta = r.z / absOld ; p = z + beta p
// (forming p on the fly inside the SpMV gather instead of k_pcg_dir was measured: the second gathered vector costs
// more than the 11 us launch it saves at 1M tets.)
// The LAST block of a launch to finish (ticket counter) sums the block partials in fixed order and writes the scalars,
// so no single-block "finish" launch sits between the kernels and the sums do not depend on block scheduling.
// scal[]: [0],[1] r.z by iteration parity (iteration k reads k-1's while writing its own), [2] done flag,
// [3] completed iterations, [4] |b|^2, [5] |r|^2, [6] threshold, [7] p.t.  Once `done` is set every later launch of
// the batch returns at once; the host looks at the flag once per batch of iterations.
// ---------------------------------------------------------------------------------------------------------------
constexpr int kPcgBlocks = 1184;       // 8 x 148 SMs, 256 threads each (vector kernels)
constexpr int kSpmvBlocks = 740;       // 5 x 148: what is resident at the SpMV's 48 registers (tools/ubench_spmv.cu)

__device__ __forceinline__ double block_sum_256(double v, double* sh) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
  __syncthreads();                                  // sh may still be read from a previous call
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = v;
  __syncthreads();
  double r = 0.0;
#pragma unroll
  for (int i = 0; i < 8; ++i) r += sh[i];           // same order in every thread and block
  return r;
}
// sum of this launch's block partials in fixed order
__device__ __forceinline__ double sum_partials_all(const double* partials, double* sh) {
  double v = 0.0;
  for (int i = threadIdx.x; i < (int)gridDim.x; i += 256) v += __ldcg(&partials[i]);     // L2: written by other blocks
  return block_sum_256(v, sh);
}

// Half a warp per block-row, one lane per d x d block (an interior vertex of a tet mesh has ~15 blocks): the lane reads
// its d contiguous values of each of the d scalar rows (adjacent lanes read adjacent 8d-byte pieces) and the d entries
// of the direction.  The row pointers of the half-warp's NEXT block-row are fetched before the current one is
// consumed, so the blockptr -> values dependency does not leave the memory pipe idle between rows.
// true in exactly one block per launch: the last one to arrive; its view of every block's partials is complete
__device__ __forceinline__ bool last_block_done(unsigned* ticket) {
  __shared__ bool last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned tk = atomicAdd(ticket, 1u);
    last = (tk == gridDim.x - 1);
    if (last) *ticket = 0u;                           // ready for the next launch
  }
  __syncthreads();
  if (last) __threadfence();
  return last;
}

// ---- peer mailboxes (PeerView, mfb_ctx.cuh) ----
// ONE block (the last to arrive of its kernel; every block has fenced its peer stores at system scope before taking its
// ticket): thread q hands this rank's partial sums to rank q, fences, then raises the sequence number there.
__device__ __forceinline__ void peer_post(const PeerView* pv, int stage, int par, unsigned long long seq, double v0, double v1) {
  const int q = threadIdx.x;
  if (q < pv->world) {
    volatile double* d = pv->sc_at[stage][q] + par * kPeerMax * 2;
    d[0] = v0; d[1] = v1;
    // release at system scope: the two values above, and (cumulatively) the halo rows the other blocks stored and fenced
    // before the barrier / ticket this thread has passed
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(pv->flag_at[stage][q]), "l"(seq) : "memory");
  }
}
// EVERY block calls it; only BLOCK 0 looks at the mailbox: its threads q < world wait (bounded) for rank q's sequence
// number with system-scope acquire loads, thread 0 adds the partial sums in rank order (identical bits on all ranks) and
// publishes (sum0, sum1, ok) in a local slot, then releases a local go-word at GPU scope; the other blocks watch that word.
// (All blocks polling the mailbox themselves was measured: 1.2 k pollers at 2 GPUs made an exchange cost 5.5 us, 4.7 k at
// 8 GPUs 19 us - the peers' flag writes queue behind the polls on the same line.)  False on a timeout.
__device__ __forceinline__ bool peer_collect(const PeerView* pv, int stage, int par, unsigned long long seq, double* v0, double* v1) {
  __shared__ int ok_sh;
  __shared__ double bc_sh[2];
  double* bc = pv->my_sc + kPeerScWords + (stage * 2 + par) * 4;                  // local broadcast slot: sum0, sum1, ok
  unsigned long long* go = pv->my_flag + kPeerFlagWords + kPeerScWords + 16 + stage;
  if (threadIdx.x == 0) ok_sh = 1;
  __syncthreads();
  if (blockIdx.x == 0) {
    if (threadIdx.x < pv->world) {
      const unsigned long long* f = pv->my_flag + stage * kPeerMax + threadIdx.x;
      long long spins = 0;
      for (;;) {
        unsigned long long v;
        asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(f) : "memory");
        if (v >= seq) break;
        if (++spins > 64) __nanosleep(64);              // the first polls back to back: the usual wait is a few microseconds
        if (spins > kPeerSpinLimit) { ok_sh = 0; break; }
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      const volatile double* sc = pv->my_sc + (size_t)stage * 2 * kPeerMax * 2 + (size_t)par * kPeerMax * 2;
      double a = 0.0, b = 0.0;
      for (int q = 0; q < pv->world; ++q) { a += sc[2 * q]; b += sc[2 * q + 1]; }
      bc_sh[0] = a; bc_sh[1] = b;
      ((volatile double*)bc)[0] = a; ((volatile double*)bc)[1] = b; ((volatile double*)bc)[2] = (double)ok_sh;
      asm volatile("st.release.gpu.global.u64 [%0], %1;" ::"l"(go), "l"(seq) : "memory");
    }
    __syncthreads();
  } else {
    if (threadIdx.x == 0) {
      long long spins = 0;
      for (;;) {
        unsigned long long v;
        asm volatile("ld.acquire.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(go) : "memory");
        if (v >= seq) break;
        __nanosleep(100);
        if (++spins > kPeerSpinLimit) { ok_sh = 0; break; }
      }
      if (ok_sh) {
        bc_sh[0] = ((volatile double*)bc)[0]; bc_sh[1] = ((volatile double*)bc)[1];
        if (((volatile double*)bc)[2] == 0.0) ok_sh = 0;
      }
    }
    __syncthreads();
  }
  *v0 = bc_sh[0]; *v1 = bc_sh[1];
  return ok_sh != 0;
}

template <int DIM>
__global__ void __launch_bounds__(256) k_pcg_spmv_dot(const int* __restrict__ blockptr, const int* __restrict__ bcol, int nF,
                                                      const double* __restrict__ A, const double* __restrict__ p,
                                                      double* __restrict__ t, double* __restrict__ scal,
                                                      double* __restrict__ partials, unsigned* __restrict__ ticket,
                                                      const int* __restrict__ rowif_ptr, const int* __restrict__ rowif_idx,
                                                      double* __restrict__ send, const PeerView* __restrict__ pv = nullptr,
                                                      unsigned long long seq = 0, int par = 0) {
  // multi-GPU (rowif_ptr != nullptr): an interface row also goes straight into the send buffer of the halo exchange -
  // or, with peer mailboxes (pv), over NVLink into the neighbour's receive segment itself
  __shared__ double sh[8];
  if (scal[2] != 0.0) return;
  const int lane = threadIdx.x & 31, hl = lane & 15;
  const int stride = gridDim.x * 16;                  // half-warps in the grid
  double dot = 0.0;
  int row = (blockIdx.x * 8 + (threadIdx.x >> 5)) * 2 + (lane >> 4);
  int b0 = 0, b1 = 0;
  if (row < nF) { b0 = blockptr[row]; b1 = blockptr[row + 1]; }
  for (int row0 = row - (lane >> 4); row0 < nF; row0 += stride, row += stride) {     // warp-uniform trip count
    const int nxt = row + stride;
    int n0 = 0, n1 = 0;
    if (nxt < nF) { n0 = blockptr[nxt]; n1 = blockptr[nxt + 1]; }
    const int nbr = (row < nF) ? b1 - b0 : 0;
    const int rowlen = DIM * nbr;
    const double* Ar = A + (int64_t)DIM * DIM * b0;
    double acc[DIM];
#pragma unroll
    for (int r = 0; r < DIM; ++r) acc[r] = 0.0;
    for (int k = hl; k < nbr; k += 16) {
      const int col = DIM * bcol[b0 + k];
      double xv[DIM];
#pragma unroll
      for (int c = 0; c < DIM; ++c) xv[c] = __ldg(&p[col + c]);
#pragma unroll
      for (int r = 0; r < DIM; ++r)
#pragma unroll
        for (int c = 0; c < DIM; ++c) acc[r] = fma(Ar[(int64_t)r * rowlen + DIM * k + c], xv[c], acc[r]);
    }
#pragma unroll
    for (int r = 0; r < DIM; ++r) {
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) acc[r] += __shfl_down_sync(0xffffffffu, acc[r], o, 16);
    }
    if (hl == 0 && row < nF) {
#pragma unroll
      for (int r = 0; r < DIM; ++r) { t[DIM * row + r] = acc[r]; dot = fma(acc[r], p[DIM * row + r], dot); }
      if (rowif_ptr) {
        const int j0 = rowif_ptr[row], j1 = rowif_ptr[row + 1];
        for (int j = j0; j < j1; ++j) {
          double* d = pv ? pv->dst[(size_t)par * pv->n_ent + rowif_idx[j]] : send + (int64_t)DIM * rowif_idx[j];
#pragma unroll
          for (int r = 0; r < DIM; ++r) d[r] = acc[r];
        }
        // no system-scope fence here (14 k of them per SpMV at 10M tets on 2 GPUs cost 45 us): the block's __threadfence() and
        // ticket order these stores before the last block's st.release.sys, which is cumulative
      }
    }
    b0 = n0; b1 = n1;
  }
  const double s = block_sum_256(dot, sh);
  if (threadIdx.x == 0) partials[blockIdx.x] = s;
  if (last_block_done(ticket)) {
    const double pt = sum_partials_all(partials, sh);
    if (pv) {
      __shared__ double ptb;
      if (threadIdx.x == 0) ptb = pt;
      __syncthreads();
      peer_post(pv, 0, par, seq, ptb, 0.0);
    } else if (threadIdx.x == 0) scal[7] = pt;
  }
}

template <int DIM>
__global__ void __launch_bounds__(256) k_pcg_update(double* __restrict__ scal, int iter, unsigned* __restrict__ ticket,
                                                    const double* __restrict__ p, const double* __restrict__ t,
                                                    double* __restrict__ x, double* __restrict__ r, double* __restrict__ z,
                                                    const double* __restrict__ dinv, int nF, int mode,
                                                    double* __restrict__ part_out, const double* __restrict__ w, int dist,
                                                    const int* __restrict__ rowif_ptr, const int* __restrict__ rowif_idx,
                                                    const double* __restrict__ recv, const PeerView* __restrict__ pv = nullptr,
                                                    unsigned long long seq = 0) {
  // multi-GPU (rowif_ptr != nullptr): the neighbours' parts of an interface row of t = L p arrive in recv
  // peer mailboxes (pv, dist == 2): wait for exchange `seq` (halo rows + p.Lp of every rank), post r.r / r.z as seq + 1
  __shared__ double sh[8];
  if (scal[2] != 0.0) return;
  double pt = 0.0;
  if (pv) {
    double dummy;
    if (!peer_collect(pv, 0, iter & 1, seq, &pt, &dummy)) {
      if (threadIdx.x == 0) { scal[13] = 1.0; scal[2] = 2.0; }
      return;
    }
    recv = pv->my_halo + (size_t)(iter & 1) * 3 * pv->n_ent;
  } else pt = scal[7];
  const double alpha = scal[(iter + 1) & 1] / pt;
  double rr = 0.0, rz = 0.0;
  for (int gi = blockIdx.x * 256 + threadIdx.x; gi < nF; gi += gridDim.x * 256) {
    double rv[DIM], zv[DIM], tv[DIM];
#pragma unroll
    for (int k = 0; k < DIM; ++k) tv[k] = t[DIM * gi + k];
    if (rowif_ptr) {
      for (int j = rowif_ptr[gi]; j < rowif_ptr[gi + 1]; ++j) {
        const int64_t s0 = (int64_t)DIM * rowif_idx[j];
#pragma unroll
        for (int k = 0; k < DIM; ++k) tv[k] += __ldcg(&recv[s0 + k]);      // L2: written by NCCL or by a peer, never cached in L1
      }
    }
#pragma unroll
    for (int k = 0; k < DIM; ++k) {
      const int i = DIM * gi + k;
      x[i] += alpha * p[i];
      rv[k] = r[i] - alpha * tv[k];
      r[i] = rv[k];
    }
    if (mode == 0) {
#pragma unroll
      for (int k = 0; k < DIM; ++k) zv[k] = rv[k];
    } else {
#pragma unroll
      for (int a = 0; a < DIM; ++a) {
        double s = 0.0;
#pragma unroll
        for (int k = 0; k < DIM; ++k) s += dinv[(int64_t)gi * DIM * DIM + a * DIM + k] * rv[k];
        zv[a] = s;
      }
    }
#pragma unroll
    for (int k = 0; k < DIM; ++k) z[DIM * gi + k] = zv[k];
    double lr = 0.0, lz = 0.0;
#pragma unroll
    for (int k = 0; k < DIM; ++k) { lr += rv[k] * rv[k]; lz += rv[k] * zv[k]; }
    const double wi = w ? w[gi] : 1.0;               // multi-GPU: every global vertex is counted by exactly one rank
    rr += wi * lr; rz += wi * lz;
  }
  const double s0 = block_sum_256(rr, sh);
  const double s1 = block_sum_256(rz, sh);
  if (threadIdx.x == 0) { part_out[blockIdx.x] = s0; part_out[kPcgBlocks + blockIdx.x] = s1; }
  if (last_block_done(ticket)) {
    const double rr2 = sum_partials_all(part_out, sh);
    const double rz2 = sum_partials_all(part_out + kPcgBlocks, sh);
    if (dist == 2) {
      __shared__ double pb[2];
      if (threadIdx.x == 0) { pb[0] = rr2; pb[1] = rz2; }
      __syncthreads();
      peer_post(pv, 1, iter & 1, seq + 1, pb[0], pb[1]);
    } else if (threadIdx.x == 0) {
      if (dist) { scal[8] = rr2; scal[9] = rz2; }     // local sums: all-reduced, then k_pcg_dir_dist decides
      else {
        scal[5] = rr2;
        scal[iter & 1] = rz2;
        if (rr2 < scal[6]) scal[2] = 1.0;       // the reference breaks before counting this iteration
        else scal[3] += 1.0;                    // completed iterations, counted on the device (CUDA-graph replays)
      }
    }
  }
}

// ---------------------------------------------------------------------------------------------------------------
// The whole CG loop as ONE persistent cooperative kernel (option "cg_persist", default 1; preconditioners 0..2, one GPU or
// several with peer mailboxes): the three phases of an iteration are separated by grid barriers instead of kernel
// boundaries, every block adds the block partials itself (fixed order: the same bits in every block), and on several
// GPUs block 0 posts the rank's partial sums to every rank's mailbox while all blocks wait for the sequence numbers -
// compute and collective in one kernel, no launch and no NCCL call per iteration.  Measured on B200 (C3 split for 1 / 2 /
// 4 / 8 GPUs, per-rank sizes on ONE GPU): the three-launch iteration costs 66 / 44 / 30 / 28 us - a floor of ~9 us per
// dependent launch - see DESIGN.md section 6 for this kernel's numbers.
// Same recurrence and same scal[] protocol as the three kernels above (the host may continue a solve with either).
// A peer timeout in some block raises *abort; every block sees it after the next grid barrier and leaves.
// ---------------------------------------------------------------------------------------------------------------
struct PcgPersist {
  const int* blockptr; const int* bcol; int nF; const double* L;
  double *p, *t, *x, *r, *z;
  const double* dinv; int mode; const double* w;
  double* scal; double* part;                      // part: 3 x gridDim.x
  const int* rowif_ptr; const int* rowif_idx; const PeerView* pv; unsigned long long seq0;
  int iter0, niter; int* abort;
  long long* prof;                                 // option "profile": block 0's SM cycles per phase, 12 slots (null: off)
};
#define PCG_TICK(i) do { if (a.prof && blockIdx.x == 0 && tid == 0) { const long long c_ = clock64(); tacc[i] += c_ - tlast; tlast = c_; } } while (0)

template <int DIM>
__global__ void __launch_bounds__(256, 4) k_pcg_persist(const PcgPersist a) {
  namespace cg = cooperative_groups;
  cg::grid_group grid = cg::this_grid();
  __shared__ double sh[8];
  const int G = gridDim.x, tid = threadIdx.x;
  double* part0 = a.part; double* part1 = a.part + G; double* part2 = a.part + 2 * G;
  if (a.scal[2] != 0.0) return;                    // uniform: written before the launch
  const int lane = tid & 31, hl = lane & 15;
  const double thr = a.scal[6];
  double rz_prev = a.scal[(a.iter0 + 1) & 1], rr = a.scal[5];
  int completed = 0;
  bool conv = false, aborted = false;
  const PeerView* pv = a.pv;
  long long tacc[12] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0}, tlast = clock64();
  for (int k = a.iter0; k < a.iter0 + a.niter; ++k) {
    const int par = k & 1;
    const unsigned long long seqA = a.seq0 + 2ull * (unsigned long long)k + 1ull;
    // ---- phase 1: t = L p, block partials of p.t; interface rows also go to the neighbours' mailboxes
    {
      const int stride = G * 16;
      double dot = 0.0;
      int row = (blockIdx.x * 8 + (tid >> 5)) * 2 + (lane >> 4);
      int b0 = 0, b1 = 0;
      if (row < a.nF) { b0 = a.blockptr[row]; b1 = a.blockptr[row + 1]; }
      for (int row0 = row - (lane >> 4); row0 < a.nF; row0 += stride, row += stride) {
        const int nxt = row + stride;
        int n0 = 0, n1 = 0;
        if (nxt < a.nF) { n0 = a.blockptr[nxt]; n1 = a.blockptr[nxt + 1]; }
        const int nbr = (row < a.nF) ? b1 - b0 : 0;
        const int rowlen = DIM * nbr;
        const double* Ar = a.L + (int64_t)DIM * DIM * b0;
        double acc[DIM];
#pragma unroll
        for (int r = 0; r < DIM; ++r) acc[r] = 0.0;
        for (int q = hl; q < nbr; q += 16) {
          const int col = DIM * a.bcol[b0 + q];
          double xv[DIM];
#pragma unroll
          for (int c = 0; c < DIM; ++c) xv[c] = __ldcg(&a.p[col + c]);      // p changes inside this kernel: not through the read-only path
#pragma unroll
          for (int r = 0; r < DIM; ++r)
#pragma unroll
            for (int c = 0; c < DIM; ++c) acc[r] = fma(__ldg(&Ar[(int64_t)r * rowlen + DIM * q + c]), xv[c], acc[r]);
        }
#pragma unroll
        for (int r = 0; r < DIM; ++r) {
#pragma unroll
          for (int o = 8; o > 0; o >>= 1) acc[r] += __shfl_down_sync(0xffffffffu, acc[r], o, 16);
        }
        if (hl == 0 && row < a.nF) {
#pragma unroll
          for (int r = 0; r < DIM; ++r) { a.t[DIM * row + r] = acc[r]; dot = fma(acc[r], __ldcg(&a.p[DIM * row + r]), dot); }
          if (a.rowif_ptr) {
            const int j0 = a.rowif_ptr[row], j1 = a.rowif_ptr[row + 1];
            for (int j = j0; j < j1; ++j) {
              double* d = pv->dst[(size_t)par * pv->n_ent + a.rowif_idx[j]];
#pragma unroll
              for (int r = 0; r < DIM; ++r) d[r] = acc[r];
            }
            (void)j0; (void)j1;      // no fence here: the grid barrier orders these stores before block 0's release at system scope (cumulativity)
          }
        }
        b0 = n0; b1 = n1;
      }
      const double s = block_sum_256(dot, sh);
      if (tid == 0) part0[blockIdx.x] = s;
    }
    PCG_TICK(0);
    grid.sync();
    PCG_TICK(1);
    if (*(volatile int*)a.abort) { aborted = true; break; }
    double pt = sum_partials_all(part0, sh);
    PCG_TICK(2);
    if (pv) {
      if (blockIdx.x == 0) peer_post(pv, 0, par, seqA, pt, 0.0);
      PCG_TICK(3);
      double dummy;
      if (!peer_collect(pv, 0, par, seqA, &pt, &dummy) && tid == 0) *(volatile int*)a.abort = 1;
      PCG_TICK(4);
    }
    const double alpha = rz_prev / pt;
    // ---- phase 2: x += alpha p ; r -= alpha (t + halo) ; z = Dinv r ; block partials of r.r, r.z
    {
      const double* recv = pv ? pv->my_halo + (size_t)par * 3 * pv->n_ent : nullptr;
      double prr = 0.0, prz = 0.0;
      for (int gi = blockIdx.x * 256 + tid; gi < a.nF; gi += G * 256) {
        double rv[DIM], zv[DIM], tv[DIM];
#pragma unroll
        for (int q = 0; q < DIM; ++q) tv[q] = __ldcg(&a.t[DIM * gi + q]);
        if (a.rowif_ptr) {
          for (int j = a.rowif_ptr[gi]; j < a.rowif_ptr[gi + 1]; ++j) {
            const int64_t s0 = (int64_t)DIM * a.rowif_idx[j];
#pragma unroll
            for (int q = 0; q < DIM; ++q) tv[q] += __ldcg(&recv[s0 + q]);
          }
        }
#pragma unroll
        for (int q = 0; q < DIM; ++q) {
          const int i = DIM * gi + q;
          a.x[i] += alpha * __ldcg(&a.p[i]);
          rv[q] = a.r[i] - alpha * tv[q];
          a.r[i] = rv[q];
        }
        if (a.mode == 0) {
#pragma unroll
          for (int q = 0; q < DIM; ++q) zv[q] = rv[q];
        } else {
#pragma unroll
          for (int c = 0; c < DIM; ++c) {
            double s = 0.0;
#pragma unroll
            for (int q = 0; q < DIM; ++q) s += __ldg(&a.dinv[(int64_t)gi * DIM * DIM + c * DIM + q]) * rv[q];
            zv[c] = s;
          }
        }
#pragma unroll
        for (int q = 0; q < DIM; ++q) a.z[DIM * gi + q] = zv[q];
        double lr = 0.0, lz = 0.0;
#pragma unroll
        for (int q = 0; q < DIM; ++q) { lr += rv[q] * rv[q]; lz += rv[q] * zv[q]; }
        const double wi = a.w ? a.w[gi] : 1.0;
        prr += wi * lr; prz += wi * lz;
      }
      const double s0 = block_sum_256(prr, sh);
      const double s1 = block_sum_256(prz, sh);
      if (tid == 0) { part1[blockIdx.x] = s0; part2[blockIdx.x] = s1; }
    }
    PCG_TICK(5);
    grid.sync();
    PCG_TICK(6);
    if (*(volatile int*)a.abort) { aborted = true; break; }
    rr = sum_partials_all(part1, sh);
    double rz = sum_partials_all(part2, sh);
    PCG_TICK(7);
    if (pv) {
      if (blockIdx.x == 0) peer_post(pv, 1, par, seqA + 1ull, rr, rz);
      PCG_TICK(8);
      if (!peer_collect(pv, 1, par, seqA + 1ull, &rr, &rz) && tid == 0) *(volatile int*)a.abort = 1;
      PCG_TICK(9);
    }
    if (rr < thr) { conv = true; break; }           // the reference breaks before counting this iteration
    // ---- phase 3: p = z + beta p (the thread that wrote z[i] reads it back: same index map as phase 2)
    {
      const double beta = rz / rz_prev;
      for (int gi = blockIdx.x * 256 + tid; gi < a.nF; gi += G * 256) {
#pragma unroll
        for (int q = 0; q < DIM; ++q) { const int i = DIM * gi + q; a.p[i] = fma(beta, __ldcg(&a.p[i]), a.z[i]); }
      }
    }
    rz_prev = rz;
    ++completed;
    PCG_TICK(10);
    grid.sync();
    PCG_TICK(11);
    if (*(volatile int*)a.abort) { aborted = true; break; }
  }
  if (blockIdx.x == 0 && tid == 0) {
    if (a.prof) for (int i = 0; i < 12; ++i) a.prof[i] += tacc[i];
    a.scal[5] = rr;
    a.scal[3] += (double)completed;
    a.scal[(a.iter0 + completed + 1) & 1] = rz_prev;
    if (conv) a.scal[2] = 1.0;
    if (aborted) { a.scal[13] = 1.0; a.scal[2] = 2.0; }
  }
}

// beta = r.z of this iteration over r.z of the previous one ; p = z + beta p
__global__ void __launch_bounds__(256) k_pcg_dir(const double* __restrict__ scal, int iter, const double* __restrict__ z,
                                                 double* __restrict__ p, int n) {
  if (scal[2] != 0.0) return;
  const double beta = scal[iter & 1] / scal[(iter + 1) & 1];
  for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) p[i] = fma(beta, p[i], z[i]);
}
// multi-GPU: scal[8], scal[9] = global r.r, r.z of this iteration (all-reduced).  Every block takes the same decision; a
// block that starts after block 0 has set the flag returns early, which is harmless (the flag means: converged).
__global__ void __launch_bounds__(256) k_pcg_dir_dist(double* __restrict__ scal, int iter, const double* __restrict__ z,
                                                      double* __restrict__ p, int n, const PeerView* __restrict__ pv = nullptr,
                                                      unsigned long long seq = 0) {
  if (scal[2] != 0.0) return;
  double rr, rz;
  if (pv) {                                           // r.r / r.z of every rank, posted by k_pcg_update as exchange `seq`
    if (!peer_collect(pv, 1, iter & 1, seq, &rr, &rz)) {
      if (threadIdx.x == 0) { scal[13] = 1.0; scal[2] = 2.0; }
      return;
    }
  } else { rr = scal[8]; rz = scal[9]; }
  const double absOld = scal[(iter + 1) & 1];
  const bool conv = rr < scal[6];
  if (!conv) {
    const double beta = rz / absOld;
    for (int i = blockIdx.x * 256 + threadIdx.x; i < n; i += gridDim.x * 256) p[i] = fma(beta, p[i], z[i]);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    scal[5] = rr;
    if (conv) scal[2] = 1.0;
    else { scal[iter & 1] = rz; scal[3] += 1.0; }
  }
}
// ---------------------------------------------------------------------------------------------------------------
// precond 3: Chebyshev polynomial over block-Jacobi.  z = q(Dinv L) Dinv r with q the degree-(deg-1) Chebyshev
// polynomial for the interval [lmax / ratio, lmax] of the spectrum of Dinv L: every step is an SpMV plus a vertex-wise
// update, nothing sequential (the reference's stronger preconditioners - IncompleteCholesky, the Laplacian LLT, amgcl,
// src/factories/linear_solver_factory.cpp:119-173 - are triangular solves / hierarchies; this is the one that maps onto
// the SpMV this file already has).  NOT in the reference.  A fixed SPD operator for lmax >= lambda_max, hence a valid CG
// preconditioner; oracle/pcg.py::make_chebyshev restates it.
//   k_cheb_update : the x / r update of the CG iteration (k_pcg_update's part) + step 0: res = r, d = Dinv r / theta, z = d,
//                   local |r|^2 -> scal[8] (and r.z -> scal[9] when deg == 1)
//   k_cheb_step   : res -= t (t = L d from the SpMV before), d = c1 d + c2 Dinv res, z += d; last step: r.z -> scal[9]
// then k_pcg_dir_dist takes the decision and forms p (after the all-reduce of scal[8..9] on several GPUs).
// ---------------------------------------------------------------------------------------------------------------
template <int DIM>
__device__ __forceinline__ void dinv_apply(const double* __restrict__ dinv, int gi, const double* v, double* o) {
#pragma unroll
  for (int a = 0; a < DIM; ++a) {
    double s = 0.0;
#pragma unroll
    for (int k = 0; k < DIM; ++k) s += dinv[(int64_t)gi * DIM * DIM + a * DIM + k] * v[k];
    o[a] = s;
  }
}
// cg != 0: the CG update x += alpha p, r -= alpha t precedes step 0 (t gets the neighbours' halo parts first);
// cg == 0: plain step 0 on the given r (the application before the first iteration)
template <int DIM>
__global__ void __launch_bounds__(256) k_cheb_update(double* __restrict__ scal, int iter, unsigned* __restrict__ ticket, int cg,
                                                     const double* __restrict__ p, const double* __restrict__ t,
                                                     double* __restrict__ x, double* __restrict__ r, double* __restrict__ res,
                                                     double* __restrict__ d, double* __restrict__ z,
                                                     const double* __restrict__ dinv, int nF, double inv_theta, int want_rz,
                                                     double* __restrict__ part_out, const double* __restrict__ w,
                                                     const int* __restrict__ rowif_ptr, const int* __restrict__ rowif_idx,
                                                     const double* __restrict__ recv) {
  __shared__ double sh[8];
  if (cg && scal[2] != 0.0) return;
  const double alpha = cg ? scal[(iter + 1) & 1] / scal[7] : 0.0;
  double rr = 0.0, rz = 0.0;
  for (int gi = blockIdx.x * 256 + threadIdx.x; gi < nF; gi += gridDim.x * 256) {
    double rv[DIM], y[DIM];
    if (cg) {
      double tv[DIM];
#pragma unroll
      for (int k = 0; k < DIM; ++k) tv[k] = t[DIM * gi + k];
      if (rowif_ptr) {
        for (int j = rowif_ptr[gi]; j < rowif_ptr[gi + 1]; ++j) {
          const int64_t s0 = (int64_t)DIM * rowif_idx[j];
#pragma unroll
          for (int k = 0; k < DIM; ++k) tv[k] += recv[s0 + k];
        }
      }
#pragma unroll
      for (int k = 0; k < DIM; ++k) {
        const int i = DIM * gi + k;
        x[i] += alpha * p[i];
        rv[k] = r[i] - alpha * tv[k];
        r[i] = rv[k];
      }
    } else {
#pragma unroll
      for (int k = 0; k < DIM; ++k) rv[k] = r[DIM * gi + k];
    }
    dinv_apply<DIM>(dinv, gi, rv, y);
    double lr = 0.0, lz = 0.0;
#pragma unroll
    for (int k = 0; k < DIM; ++k) {
      const int i = DIM * gi + k;
      const double dv = y[k] * inv_theta;
      res[i] = rv[k]; d[i] = dv; z[i] = dv;
      lr += rv[k] * rv[k]; lz += rv[k] * dv;
    }
    const double wi = w ? w[gi] : 1.0;
    rr += wi * lr; rz += wi * lz;
  }
  const double s0 = block_sum_256(rr, sh);
  const double s1 = block_sum_256(rz, sh);
  if (threadIdx.x == 0) { part_out[blockIdx.x] = s0; part_out[kPcgBlocks + blockIdx.x] = s1; }
  if (last_block_done(ticket)) {
    const double rr2 = sum_partials_all(part_out, sh);
    const double rz2 = sum_partials_all(part_out + kPcgBlocks, sh);
    if (threadIdx.x == 0) { scal[8] = rr2; if (want_rz) scal[9] = rz2; }
  }
}
template <int DIM>
__global__ void __launch_bounds__(256) k_cheb_step(double* __restrict__ scal, unsigned* __restrict__ ticket, int check_flag,
                                                   const double* __restrict__ t, const double* __restrict__ r, double* __restrict__ res,
                                                   double* __restrict__ d, double* __restrict__ z, const double* __restrict__ dinv, int nF,
                                                   double c1, double c2, int want_rz, double* __restrict__ part_out,
                                                   const double* __restrict__ w, const int* __restrict__ rowif_ptr,
                                                   const int* __restrict__ rowif_idx, const double* __restrict__ recv) {
  __shared__ double sh[8];
  if (check_flag && scal[2] != 0.0) return;
  double rz = 0.0;
  for (int gi = blockIdx.x * 256 + threadIdx.x; gi < nF; gi += gridDim.x * 256) {
    double rs[DIM], y[DIM];
#pragma unroll
    for (int k = 0; k < DIM; ++k) rs[k] = res[DIM * gi + k] - t[DIM * gi + k];
    if (rowif_ptr) {
      for (int j = rowif_ptr[gi]; j < rowif_ptr[gi + 1]; ++j) {
        const int64_t s0 = (int64_t)DIM * rowif_idx[j];
#pragma unroll
        for (int k = 0; k < DIM; ++k) rs[k] -= recv[s0 + k];
      }
    }
    dinv_apply<DIM>(dinv, gi, rs, y);
    double lz = 0.0;
#pragma unroll
    for (int k = 0; k < DIM; ++k) {
      const int i = DIM * gi + k;
      const double dv = c1 * d[i] + c2 * y[k];
      const double zv = z[i] + dv;
      res[i] = rs[k]; d[i] = dv; z[i] = zv;
      if (want_rz) lz += r[i] * zv;
    }
    rz += (w ? w[gi] : 1.0) * lz;
  }
  if (!want_rz) return;
  const double s1 = block_sum_256(rz, sh);
  if (threadIdx.x == 0) part_out[blockIdx.x] = s1;
  if (last_block_done(ticket)) {
    const double rz2 = sum_partials_all(part_out, sh);
    if (threadIdx.x == 0) scal[9] = rz2;
  }
}
// power-iteration start: a rough vector (integer hash of the dof index, reproducible on the host), so that it has a
// component along the top of the spectrum - Dinv b is smooth (gravity) and made 10 iterations underestimate lambda_max
__global__ void k_hash_vec(double* __restrict__ v, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) v[i] = (double)((unsigned)i * 2654435761u) * (1.0 / 4294967296.0) - 0.5;
}
__global__ void k_scale_vec(double* __restrict__ v, double a, int n) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) v[i] *= a;
}

// weighted two-stage dots (w per vertex, dim entries each)
__global__ void __launch_bounds__(256) k_dot2w(const double* __restrict__ a, const double* __restrict__ b, const double* __restrict__ c2,
                                               const double* __restrict__ d2, const double* __restrict__ w, int dim, int n,
                                               double* partials, int np) {
  __shared__ double sh[2][8];
  double s0 = 0.0, s1 = 0.0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
    const double wi = w[i / dim];
    s0 += wi * a[i] * b[i];
    if (c2) s1 += wi * c2[i] * d2[i];
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
// diagonal d x d blocks of the (local, sub-assembled) operator, d*d per block-row: summed over ranks, then inverted
template <int DIM>
__global__ void k_extract_diag(const int* __restrict__ blockptr, const int* __restrict__ diagblk, int nF, const double* __restrict__ A,
                               double* __restrict__ D) {
  const int gi = blockIdx.x * blockDim.x + threadIdx.x;
  if (gi >= nF) return;
  const int b0 = blockptr[gi], nbr = blockptr[gi + 1] - b0, rank = diagblk[gi] - b0;
  const int64_t base = (int64_t)DIM * DIM * b0;
#pragma unroll
  for (int r = 0; r < DIM; ++r)
#pragma unroll
    for (int c = 0; c < DIM; ++c) D[(int64_t)gi * DIM * DIM + r * DIM + c] = A[base + (int64_t)r * DIM * nbr + DIM * rank + c];
}
template <int DIM>
__global__ void k_invert_diag(int nF, int mode, double* __restrict__ D) {
  const int gi = blockIdx.x * blockDim.x + threadIdx.x;
  if (gi >= nF) return;
  double* d = D + (int64_t)gi * DIM * DIM;
  double a[DIM][DIM], I[DIM][DIM];
#pragma unroll
  for (int r = 0; r < DIM; ++r)
#pragma unroll
    for (int c = 0; c < DIM; ++c) a[r][c] = d[r * DIM + c];
  if (mode == 1) {
#pragma unroll
    for (int r = 0; r < DIM; ++r)
#pragma unroll
      for (int c = 0; c < DIM; ++c) I[r][c] = (r == c) ? ((a[r][r] != 0.0) ? 1.0 / a[r][r] : 1.0) : 0.0;
  } else if constexpr (DIM == 2) {
    const double id = 1.0 / (a[0][0] * a[1][1] - a[0][1] * a[1][0]);
    I[0][0] = a[1][1] * id; I[0][1] = -a[0][1] * id; I[1][0] = -a[1][0] * id; I[1][1] = a[0][0] * id;
  } else {
    const double c00 = a[1][1] * a[2][2] - a[1][2] * a[2][1];
    const double c01 = a[1][2] * a[2][0] - a[1][0] * a[2][2];
    const double c02 = a[1][0] * a[2][1] - a[1][1] * a[2][0];
    const double id = 1.0 / (a[0][0] * c00 + a[0][1] * c01 + a[0][2] * c02);
    I[0][0] = c00 * id;
    I[0][1] = (a[0][2] * a[2][1] - a[0][1] * a[2][2]) * id;
    I[0][2] = (a[0][1] * a[1][2] - a[0][2] * a[1][1]) * id;
    I[1][0] = c01 * id;
    I[1][1] = (a[0][0] * a[2][2] - a[0][2] * a[2][0]) * id;
    I[1][2] = (a[0][2] * a[1][0] - a[0][0] * a[1][2]) * id;
    I[2][0] = c02 * id;
    I[2][1] = (a[0][1] * a[2][0] - a[0][0] * a[2][1]) * id;
    I[2][2] = (a[0][0] * a[1][1] - a[0][1] * a[1][0]) * id;
  }
#pragma unroll
  for (int r = 0; r < DIM; ++r)
#pragma unroll
    for (int c = 0; c < DIM; ++c) d[r * DIM + c] = I[r][c];
}
// lhs = A + M, once per solve (linear_system.h: lhs_ = M + A is formed once per Newton iteration)
__global__ void k_sum_values(const double* __restrict__ a, const double* __restrict__ b, double* __restrict__ o, int64_t n) {
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) o[i] = a[i] + b[i];
}
// scal[4] = |b|^2, scal[5] = |r|^2, scal[0] = r.z are in place; "iteration -1" has parity 1
__global__ void k_pcg_init(double* scal, double tol, unsigned* tickets) {
  scal[1] = scal[0]; scal[2] = 0.0; scal[3] = 0.0; scal[6] = tol * tol * scal[4]; scal[7] = 0.0; scal[13] = 0.0;
  tickets[0] = 0u; tickets[1] = 0u;
}

int spmv(mfb_ctx* c, const double* Mv, const double* x, double* y) {
  const int64_t threads = (int64_t)c->nF * 16;
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

// block 0's SM cycles per phase of k_pcg_persist, accumulated while option "profile" is 1 (12 slots: phase 1, barrier, sum,
// post, collect, phase 2, barrier, sums, post, collect, phase 3, barrier); reset != 0 clears them.  Debug aid, not in the header.
extern "C" int mfb_debug_pcg_cycles(mfb_ctx* c, long long* out12, int reset) {
  if (!c || !out12) return MFB_ERR_INVALID;
  cudaSetDevice(c->device);
  for (int i = 0; i < 12; ++i) out12[i] = 0;
  if (!c->pcg_prof.p) return MFB_OK;
  MFB_CUDA(cudaMemcpyAsync(out12, c->pcg_prof.p, 12 * sizeof(long long), cudaMemcpyDeviceToHost, c->stream));
  if (reset) MFB_CUDA(cudaMemsetAsync(c->pcg_prof.p, 0, 16 * sizeof(long long), c->stream));
  MFB_CUDA(cudaStreamSynchronize(c->stream));
  return MFB_OK;
}

extern "C" int mfb_pcg_dev(mfb_ctx* c, const double* Mv, const double* b, double* x, int precond, double tol, int max_iters,
                           int* iters_out, double* relres_out) {
  CHECK_CTX(c);
  if (!b || !x || precond < 0 || precond > 3) return MFB_ERR_INVALID;
  const bool cheb = precond == 3;
  const int pmode = cheb ? 2 : precond;          // what the diagonal-block kernels build: Chebyshev runs over block-Jacobi
  const int deg = cheb ? std::max(1, c->cheb_degree) : 1;
  const bool dist = c->world > 1;
  if (dist && !c->dot_w.p) { g_last_error = "mfb_pcg_dev: multi-GPU context without mfb_comm_set_interface"; return MFB_ERR_INVALID; }
  const double* w = dist ? c->dot_w.p : nullptr;
  const int n = c->dim * c->nF;
  const int dd = c->dim * c->dim;
  int rc;
  if (c->pcg_r.n < (size_t)n + 1) {
    if ((rc = c->pcg_r.alloc((size_t)n + 1)) || (rc = c->pcg_p.alloc((size_t)n + 1)) || (rc = c->pcg_z.alloc((size_t)n + 1)) ||
        (rc = c->pcg_t.alloc((size_t)n + 1)) || (rc = c->pcg_dinv.alloc((size_t)c->nF * dd + 1)))
      return rc;
  }
  if (cheb && c->pcg_res.n < (size_t)n + 1) {
    if ((rc = c->pcg_res.alloc((size_t)n + 1)) || (rc = c->pcg_d.alloc((size_t)n + 1))) return rc;
  }
  if (c->partials.n < (size_t)2 * kDotBlocks) { if ((rc = c->partials.alloc((size_t)2 * kDotBlocks + c->grid()))) return rc; }
  if (c->pcg_part.n < (size_t)5 * kPcgBlocks) { if ((rc = c->pcg_part.alloc((size_t)5 * kPcgBlocks))) return rc; }
  cudaStream_t st = c->stream;
  double *r = c->pcg_r.p, *p = c->pcg_p.p, *z = c->pcg_z.p, *t = c->pcg_t.p;
  // the operator of this solve: one value array, L = A + M (multi-GPU: this rank's sub-assembled part of it)
  const double* L = c->Avals.p;
  if (Mv) {
    if (c->pcg_lhs.n < (size_t)c->nnz + 1) { if ((rc = c->pcg_lhs.alloc((size_t)c->nnz + 1))) return rc; }
    k_sum_values<<<kPcgBlocks, 256, 0, st>>>(c->Avals.p, Mv, c->pcg_lhs.p, c->nnz);
    c->launches += 1;
    L = c->pcg_lhs.p;
  }
  auto spmv_L = [&](const double* xx, double* yy) {
    const int64_t threads = (int64_t)c->nF * 16;
    if (c->nF == 0) return;
    if (c->dim == 3) k_spmv<3, false><<<nblk(threads), 256, 0, st>>>(c->blockptr.p, c->bcol.p, c->nF, L, nullptr, xx, yy);
    else k_spmv<2, false><<<nblk(threads), 256, 0, st>>>(c->blockptr.p, c->bcol.p, c->nF, L, nullptr, xx, yy);
    c->launches += 1;
  };
  auto dots = [&](const double* a0, const double* b0, const double* a1, const double* b1, int out0, int out1) -> int {
    if (!dist) return dot2(c, a0, b0, a1, b1, n, out0, out1);
    k_dot2w<<<kDotBlocks, 256, 0, st>>>(a0, b0, a1, b1, w, c->dim, n, c->partials.p, kDotBlocks);
    k_finish2<<<1, 256, 0, st>>>(c->partials.p, kDotBlocks, c->scal.p + out0, out1 >= 0 ? c->scal.p + out1 : nullptr);
    c->launches += 2;
    int e = allreduce_scalars(c, c->scal.p + out0, 1, false);
    if (!e && out1 >= 0) e = allreduce_scalars(c, c->scal.p + out1, 1, false);
    return e;
  };
  if (precond && c->nF) {
    if (!dist) {
      if (c->dim == 3) k_build_precond<3><<<nblk(c->nF), 256, 0, st>>>(c->blockptr.p, c->diagblk.p, c->nF, L, nullptr, pmode, c->pcg_dinv.p);
      else k_build_precond<2><<<nblk(c->nF), 256, 0, st>>>(c->blockptr.p, c->diagblk.p, c->nF, L, nullptr, pmode, c->pcg_dinv.p);
      c->launches += 1;
    } else {       // the diagonal blocks of the GLOBAL operator: local blocks summed at the interface, then inverted
      if (c->dim == 3) k_extract_diag<3><<<nblk(c->nF), 256, 0, st>>>(c->blockptr.p, c->diagblk.p, c->nF, L, c->pcg_dinv.p);
      else k_extract_diag<2><<<nblk(c->nF), 256, 0, st>>>(c->blockptr.p, c->diagblk.p, c->nF, L, c->pcg_dinv.p);
      if ((rc = interface_sum(c, c->pcg_dinv.p, dd, nullptr, 0, nullptr, 0))) return rc;
      if (c->dim == 3) k_invert_diag<3><<<nblk(c->nF), 256, 0, st>>>(c->nF, pmode, c->pcg_dinv.p);
      else k_invert_diag<2><<<nblk(c->nF), 256, 0, st>>>(c->nF, pmode, c->pcg_dinv.p);
      c->launches += 2;
    }
  }
  // residual = b - L x ; |b|^2 ; |r|^2
  spmv_L(x, t);
  if (dist && (rc = interface_sum(c, t, c->dim, nullptr, 0, nullptr, 0))) return rc;
  if (n) { k_residual<<<nblk(n), 256, 0, st>>>(b, t, r, n); c->launches += 1; }
  if ((rc = dots(b, b, r, r, 4, 5))) return rc;                        // scal[4] = |b|^2, scal[5] = |r|^2
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
  if (resNorm2 >= threshold && max_iters > 0) {
    double* part = c->pcg_part.p;                   // pt[G] | rr[G] | rz[G] | two ticket counters
    unsigned* tickets = reinterpret_cast<unsigned*>(part + 3 * kPcgBlocks);
    // ---- Chebyshev set-up: lambda_max of Dinv L by power iteration from Dinv b, then the step coefficients
    double inv_theta = 1.0;
    std::vector<double> cc1, cc2;
    double *res = c->pcg_res.p, *dvec = c->pcg_d.p;
    auto plain_product = [&](const double* xx, double* yy) -> int {       // yy = L xx, consistent on every rank
      spmv_L(xx, yy);
      return dist ? interface_sum(c, yy, c->dim, nullptr, 0, nullptr, 0) : MFB_OK;
    };
    auto apply_dinv = [&](const double* in, double* out) {
      if (!c->nF) return;
      if (c->dim == 3) k_apply_precond<3><<<nblk(c->nF), 256, 0, st>>>(c->pcg_dinv.p, in, out, c->nF, pmode);
      else k_apply_precond<2><<<nblk(c->nF), 256, 0, st>>>(c->pcg_dinv.p, in, out, c->nF, pmode);
      c->launches += 1;
    };
    if (cheb) {
      double lam = 1.0;
      constexpr int kPower = 12;
      k_hash_vec<<<nblk(n), 256, 0, st>>>(z, n);                          // (z, t, p are free here)
      for (int itp = 0; itp <= kPower; ++itp) {
        if ((rc = dots(z, z, nullptr, nullptr, 10, -1))) return rc;
        MFB_CUDA(cudaMemcpyAsync(c->h_scal, c->scal.p + 10, sizeof(double), cudaMemcpyDeviceToHost, st));
        MFB_CUDA(cudaStreamSynchronize(st));
        const double nrm = sqrt(c->h_scal[0]);
        if (!(nrm > 0.0) || !std::isfinite(nrm)) { g_last_error = "mfb_pcg_dev: Chebyshev power iteration broke down"; return MFB_ERR_STATE; }
        if (itp > 0) lam = nrm;                                           // |Dinv L v| for |v| = 1
        if (itp == kPower) break;
        k_scale_vec<<<nblk(n), 256, 0, st>>>(z, 1.0 / nrm, n);
        if ((rc = plain_product(z, t))) return rc;
        apply_dinv(t, z);
        c->launches += 1;
      }
      // never above the rigorous bound: K_e <= npe blockdiag(K_e) for every PSD element matrix (and the mass blocks), hence
      // lambda_max(Dinv L) <= vertices per element
      const double lmax = std::min(1.2 * lam, (double)c->npe), lmin = lmax / (c->cheb_ratio > 1 ? (double)c->cheb_ratio : 30.0);
      const double theta = 0.5 * (lmax + lmin), delta = 0.5 * (lmax - lmin), sigma = theta / delta;
      inv_theta = 1.0 / theta;
      double rho = 1.0 / sigma;
      for (int j = 1; j < deg; ++j) {
        const double rho_new = 1.0 / (2.0 * sigma - rho);
        cc1.push_back(rho_new * rho); cc2.push_back(2.0 * rho_new / delta);
        rho = rho_new;
      }
    }
    const unsigned one = 1u;
    const unsigned gv = std::max(one, std::min(nblk(c->nF), (unsigned)kPcgBlocks));      // one vertex per thread
    const unsigned gn = std::max(one, std::min(nblk(n), (unsigned)kPcgBlocks));
    const unsigned gs = std::max(one, std::min(nblk((int64_t)c->nF * 16), (unsigned)kSpmvBlocks));
    // p = (preconditioner) r ; absNew = r.p
    if (!cheb) {
      apply_dinv(r, p);
    } else {                                       // z = Cheb(r) with plain launches (no flags yet), then p = z
      MFB_CUDA(cudaMemsetAsync(tickets, 0, 2 * sizeof(unsigned), st));
      if (c->dim == 3) k_cheb_update<3><<<gv, 256, 0, st>>>(c->scal.p, 0, tickets + 1, 0, nullptr, nullptr, nullptr, r, res, dvec, z, c->pcg_dinv.p, c->nF, inv_theta, 0, part + kPcgBlocks, w, nullptr, nullptr, nullptr);
      else k_cheb_update<2><<<gv, 256, 0, st>>>(c->scal.p, 0, tickets + 1, 0, nullptr, nullptr, nullptr, r, res, dvec, z, c->pcg_dinv.p, c->nF, inv_theta, 0, part + kPcgBlocks, w, nullptr, nullptr, nullptr);
      for (int j = 1; j < deg; ++j) {
        if ((rc = plain_product(dvec, t))) return rc;
        if (c->dim == 3) k_cheb_step<3><<<gv, 256, 0, st>>>(c->scal.p, tickets + 1, 0, t, r, res, dvec, z, c->pcg_dinv.p, c->nF, cc1[j - 1], cc2[j - 1], 0, part + kPcgBlocks, w, nullptr, nullptr, nullptr);
        else k_cheb_step<2><<<gv, 256, 0, st>>>(c->scal.p, tickets + 1, 0, t, r, res, dvec, z, c->pcg_dinv.p, c->nF, cc1[j - 1], cc2[j - 1], 0, part + kPcgBlocks, w, nullptr, nullptr, nullptr);
      }
      MFB_CUDA(cudaMemcpyAsync(p, z, (size_t)n * 8, cudaMemcpyDeviceToDevice, st));
      c->launches += deg;
    }
    if ((rc = dots(r, p, nullptr, nullptr, 0, -1))) return rc;         // scal[0] = absNew
    k_pcg_init<<<1, 1, 0, st>>>(c->scal.p, tol, tickets); c->launches += 1;
    // multi-GPU with a partition set: the halo of t = L p goes point-to-point, written / read inside the CG kernels
    const bool p2p = dist && c->use_p2p && c->ent_local.p;
    // ... and with peer mailboxes (all GPUs on one node, CUDA IPC available) without NCCL at all: the SpMV kernel stores the
    // interface rows into the neighbours' memory, the last block of each kernel posts the rank's partial sums to everybody,
    // the next kernel waits for the sequence number (two exchanges per iteration: seq0 + 2k + 1, seq0 + 2k + 2)
    bool peer = false;
    if (p2p && !cheb) { if ((rc = peer_setup(c))) return rc; peer = c->use_peer && c->peer_state == 1; }
    const PeerView* pvd = peer ? c->peer_view.p : nullptr;
    const unsigned long long seq0 = c->peer_seq;
    if (peer) c->peer_seq += 2ull * (unsigned long long)max_iters + 2ull;      // the same on every rank, whatever the iteration count
    double *hsend = nullptr, *hrecv = nullptr;
    const int *rif_ptr = nullptr, *rif_idx = nullptr;
    if (p2p) {
      if ((rc = halo_buffers(c, c->dim, &hsend, &hrecv))) return rc;
      rif_ptr = c->rowif_ptr.p; rif_idx = c->rowif_idx.p;
    }
    // One CG iteration = 3 launches (+ 2 NCCL groups on several GPUs).  The kernels only depend on the PARITY of the
    // iteration number (the count itself lives in scal[3]), so an even number of iterations is captured ONCE per solve into a
    // CUDA graph and replayed: no per-launch host cost, which is what a small per-GPU problem is made of (8 GPUs, 1M
    // tets: 68 us per iteration enqueued from the host against ~15 us of kernels).  The first iterations run eagerly (NCCL
    // sets up its connections on first use, which cannot be captured); a failed capture falls back to the eager loop.
    auto enqueue_iteration = [&](int k) -> int {
      const unsigned long long seqA = seq0 + 2ull * (unsigned long long)k + 1ull;
      if (c->dim == 3) k_pcg_spmv_dot<3><<<gs, 256, 0, st>>>(c->blockptr.p, c->bcol.p, c->nF, L, p, t, c->scal.p, part, tickets, rif_ptr, rif_idx, hsend, pvd, seqA, k & 1);
      else k_pcg_spmv_dot<2><<<gs, 256, 0, st>>>(c->blockptr.p, c->bcol.p, c->nF, L, p, t, c->scal.p, part, tickets, rif_ptr, rif_idx, hsend, pvd, seqA, k & 1);
      // multi-GPU: t <- sum over ranks at the interface and p.t <- sum of the local p.(L_r p): one NCCL group
      // (neighbour send / recv + the scalar), or one all-reduce over every shared vertex without a partition
      int e = MFB_OK;
      if (peer) e = MFB_OK;
      else if (p2p) e = halo_exchange(c, c->dim, c->scal.p + 7, 1);
      else if (dist) e = interface_sum(c, t, c->dim, nullptr, 0, c->scal.p + 7, 1);
      if (e) return e;
      if (cheb) {
        // CG update + Chebyshev step 0, then (deg - 1) x [SpMV of the correction, halo, step]; r.r and r.z end up in scal[8..9]
        if (c->dim == 3) k_cheb_update<3><<<gv, 256, 0, st>>>(c->scal.p, k, tickets + 1, 1, p, t, x, r, res, dvec, z, c->pcg_dinv.p, c->nF, inv_theta, deg == 1, part + kPcgBlocks, w, rif_ptr, rif_idx, hrecv);
        else k_cheb_update<2><<<gv, 256, 0, st>>>(c->scal.p, k, tickets + 1, 1, p, t, x, r, res, dvec, z, c->pcg_dinv.p, c->nF, inv_theta, deg == 1, part + kPcgBlocks, w, rif_ptr, rif_idx, hrecv);
        for (int j = 1; j < deg; ++j) {
          if (c->dim == 3) k_pcg_spmv_dot<3><<<gs, 256, 0, st>>>(c->blockptr.p, c->bcol.p, c->nF, L, dvec, t, c->scal.p, part, tickets, rif_ptr, rif_idx, hsend);
          else k_pcg_spmv_dot<2><<<gs, 256, 0, st>>>(c->blockptr.p, c->bcol.p, c->nF, L, dvec, t, c->scal.p, part, tickets, rif_ptr, rif_idx, hsend);
          if (p2p) e = halo_exchange(c, c->dim, nullptr, 0);
          else if (dist) e = interface_sum(c, t, c->dim, nullptr, 0, nullptr, 0);
          if (e) return e;
          const int lastj = j == deg - 1;
          if (c->dim == 3) k_cheb_step<3><<<gv, 256, 0, st>>>(c->scal.p, tickets + 1, 1, t, r, res, dvec, z, c->pcg_dinv.p, c->nF, cc1[j - 1], cc2[j - 1], lastj, part + kPcgBlocks, w, rif_ptr, rif_idx, hrecv);
          else k_cheb_step<2><<<gv, 256, 0, st>>>(c->scal.p, tickets + 1, 1, t, r, res, dvec, z, c->pcg_dinv.p, c->nF, cc1[j - 1], cc2[j - 1], lastj, part + kPcgBlocks, w, rif_ptr, rif_idx, hrecv);
        }
        c->launches += 2 * (deg - 1);
        if (dist && (e = allreduce_scalars(c, c->scal.p + 8, 2, false))) return e;
        k_pcg_dir_dist<<<gn, 256, 0, st>>>(c->scal.p, k, z, p, n);
        return MFB_OK;
      }
      const int dmode = dist ? (peer ? 2 : 1) : 0;
      if (c->dim == 3) k_pcg_update<3><<<gv, 256, 0, st>>>(c->scal.p, k, tickets + 1, p, t, x, r, z, c->pcg_dinv.p, c->nF, precond, part + kPcgBlocks, w, dmode, rif_ptr, rif_idx, hrecv, pvd, seqA);
      else k_pcg_update<2><<<gv, 256, 0, st>>>(c->scal.p, k, tickets + 1, p, t, x, r, z, c->pcg_dinv.p, c->nF, precond, part + kPcgBlocks, w, dmode, rif_ptr, rif_idx, hrecv, pvd, seqA);
      if (dist) {
        if (!peer && (e = allreduce_scalars(c, c->scal.p + 8, 2, false))) return e;
        k_pcg_dir_dist<<<gn, 256, 0, st>>>(c->scal.p, k, z, p, n, pvd, seqA + 1ull);
      } else {
        k_pcg_dir<<<gn, 256, 0, st>>>(c->scal.p, k, z, p, n);
      }
      return MFB_OK;
    };
    // the whole loop as one persistent cooperative kernel (k_pcg_persist) when nothing in it needs the host: one GPU, or
    // several with peer mailboxes
    // cg_persist 2 (default) = below 64 Ki block-rows, where the iteration is latency-bound (one GPU, per-rank sizes of C3 over
    // 8 / 4 / 2 / 1 GPUs: 24.4 / 28.9 / 43.2 / 73.2 us per iteration against 28.2 / 31.0 / 43.9 / 65.8 for three launches -
    // a kernel boundary lets the SpMV gather p through the L1, the persistent kernel must read it from the L2)
    bool persist = (c->cg_persist == 1 || (c->cg_persist == 2 && c->nF <= (1 << 16))) && !cheb && (!dist || peer) && c->nF > 0;
    if (persist) {
      static int coop_dev[64], gmax_dev[64][2];
      static bool coop_known[64] = {false};
      const int dv = c->device & 63;
      if (!coop_known[dv]) {
        int coop = 0, sms = 0, nb3 = 0, nb2 = 0;
        MFB_CUDA(cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, c->device));
        MFB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device));
        MFB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb3, k_pcg_persist<3>, 256, 0));
        MFB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb2, k_pcg_persist<2>, 256, 0));
        coop_dev[dv] = coop; gmax_dev[dv][0] = nb2 * sms; gmax_dev[dv][1] = nb3 * sms;
        coop_known[dv] = true;
      }
      const int gmax = std::min(gmax_dev[dv][c->dim == 3], kPcgBlocks);
      if (!coop_dev[dv] || gmax < 1) persist = false;
      else {
        if (!c->pcg_abort.p && (rc = c->pcg_abort.alloc(1))) return rc;
        MFB_CUDA(cudaMemsetAsync(c->pcg_abort.p, 0, sizeof(int), st));
        PcgPersist pa;
        pa.blockptr = c->blockptr.p; pa.bcol = c->bcol.p; pa.nF = c->nF; pa.L = L;
        pa.p = p; pa.t = t; pa.x = x; pa.r = r; pa.z = z;
        pa.dinv = c->pcg_dinv.p; pa.mode = precond; pa.w = w;
        pa.scal = c->scal.p; pa.part = part;
        pa.rowif_ptr = rif_ptr; pa.rowif_idx = rif_idx; pa.pv = pvd; pa.seq0 = seq0;
        pa.iter0 = 0; pa.niter = max_iters; pa.abort = c->pcg_abort.p;
        pa.prof = nullptr;
        if (c->profile) {
          if (!c->pcg_prof.p) { if ((rc = c->pcg_prof.alloc(16))) return rc; MFB_CUDA(cudaMemsetAsync(c->pcg_prof.p, 0, 16 * sizeof(long long), st)); }
          pa.prof = c->pcg_prof.p;
        }
        static const int gcap = getenv("MFB_PCG_G") ? atoi(getenv("MFB_PCG_G")) : 0;          // experiments: cap on the cooperative grid
        // blocks per SM by size: one per 9 472 block-rows (= 64 rows per block of 256 threads), 1 .. 4.  Measured on one GPU
        // (tools/run_cgsize.sh): 24 k rows 24.5 us per iteration at 4 blocks per SM, 21.6 at 2, 24.6 at 1; 48 k rows 29.3 / 28.6 /
        // 38.7 - fewer blocks make the barriers and the partial sums cheaper, more blocks hide the SpMV's latency
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device);
        const int per_sm = std::max(1, std::min(4, c->nF / (64 * sms)));
        const int gwant = gcap > 0 ? gcap : per_sm * sms;
        const unsigned G = (unsigned)std::max(1, std::min(std::min(gmax, gwant), (int)nblk((int64_t)c->nF * 16)));
        void* kargs[] = {(void*)&pa};
        if (c->dim == 3) MFB_CUDA(cudaLaunchCooperativeKernel((void*)k_pcg_persist<3>, dim3(G), dim3(256), kargs, 0, st));
        else MFB_CUDA(cudaLaunchCooperativeKernel((void*)k_pcg_persist<2>, dim3(G), dim3(256), kargs, 0, st));
        c->launches += 1;
        MFB_CUDA(cudaMemcpyAsync(c->h_scal, c->scal.p, 14 * sizeof(double), cudaMemcpyDeviceToHost, st));
        MFB_CUDA(cudaStreamSynchronize(st));
        if (c->h_scal[13] != 0.0) {
          g_last_error = "mfb_pcg_dev: a peer-mailbox exchange timed out (a rank of the communicator stopped, or peer memory is not coherent on this node); set option peer = 0 to use NCCL";
          return MFB_ERR_CUDA;
        }
      }
    }
    int launched = persist ? max_iters : 0, batch = 8;
    cudaGraphExec_t gexec = nullptr;
    const int gbatch = 32;                           // iterations per graph launch (even)
    bool try_graph = c->cg_graph != 0 && max_iters >= 3 * gbatch && !peer;      // the sequence numbers are launch parameters
    while (launched < max_iters) {
      if (try_graph && !gexec && launched >= 8 && max_iters - launched >= gbatch) {
        cudaGraph_t graph = nullptr;
        bool okc = cudaStreamBeginCapture(st, cudaStreamCaptureModeThreadLocal) == cudaSuccess;
        if (okc) {
          int e = MFB_OK;
          for (int k = launched; k < launched + gbatch && !e; ++k) e = enqueue_iteration(k);      // launched is even here
          const cudaError_t ce = cudaStreamEndCapture(st, &graph);
          okc = !e && ce == cudaSuccess && graph;
        }
        if (okc) okc = cudaGraphInstantiate(&gexec, graph, 0) == cudaSuccess;
        if (graph) cudaGraphDestroy(graph);
        if (!okc) { gexec = nullptr; try_graph = false; (void)cudaGetLastError(); }
      }
      int nb;
      if (gexec && max_iters - launched >= gbatch) {
        nb = gbatch;
        MFB_CUDA(cudaGraphLaunch(gexec, st));
      } else {
        nb = (max_iters - launched < batch) ? max_iters - launched : batch;
        if (gexec || try_graph) nb = std::min(nb, 8);                 // keep `launched` even and the eager prefix short
        for (int k = launched; k < launched + nb; ++k)
          if ((rc = enqueue_iteration(k))) { if (gexec) cudaGraphExecDestroy(gexec); return rc; }
      }
      launched += nb;
      c->launches += 3 * nb;
      MFB_CUDA(cudaMemcpyAsync(c->h_scal, c->scal.p, 14 * sizeof(double), cudaMemcpyDeviceToHost, st));
      MFB_CUDA(cudaStreamSynchronize(st));
      if (peer && c->h_scal[13] != 0.0) {
        if (gexec) cudaGraphExecDestroy(gexec);
        g_last_error = "mfb_pcg_dev: a peer-mailbox exchange timed out (a rank of the communicator stopped, or peer memory is not coherent on this node); set option peer = 0 to use NCCL";
        return MFB_ERR_CUDA;
      }
      if (c->h_scal[2] != 0.0) break;
      if (batch < 64) batch *= 2;
    }
    if (gexec) cudaGraphExecDestroy(gexec);
    it = (int)c->h_scal[3];
    resNorm2 = c->h_scal[5];
  }
  MFB_CUDA(cudaGetLastError());
  if (iters_out) *iters_out = it;
  if (relres_out) *relres_out = sqrt(resNorm2 / rhsNorm2);
  return MFB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// Hand-off to block solvers: the operator (A, or A + M when M values on the same pattern are
// given) as block-CSR with d x d row-major blocks - the layout of amgcl::backend::crs<amgcl::static_matrix<double,d,d>>
// that the reference's AMGCL solver builds from Eigen blocks (src/linear_solvers/amgcl_solver.h:40-75).
// ---------------------------------------------------------------------------------------------------------------
namespace {
template <int DIM>
__global__ void k_to_bsr(const int* __restrict__ blockptr, const int* __restrict__ brow, int nb, const double* __restrict__ A,
                         const double* __restrict__ Mv, double* __restrict__ out) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)nb * DIM * DIM) return;
  const int b = (int)(i / (DIM * DIM)), rc = (int)(i % (DIM * DIM)), r = rc / DIM, cc = rc % DIM;
  const int row = brow[b], b0 = blockptr[row], nbr = blockptr[row + 1] - b0;
  const int64_t src = (int64_t)DIM * DIM * b0 + (int64_t)r * DIM * nbr + DIM * (b - b0) + cc;
  out[i] = A[src] + (Mv ? Mv[src] : 0.0);
}
}  // namespace

extern "C" int mfb_lhs_bsr(mfb_ctx* c, const double* M_vals_dev, int32_t* ptr, int32_t* col, double* val) {
  CHECK_CTX(c);
  if (!ptr || !col || !val) return MFB_ERR_INVALID;
  const int dd = c->dim * c->dim;
  mfb::DBuf<double> tmp;
  int rc = tmp.alloc((size_t)c->nb * dd + 1);
  if (rc) return rc;
  cudaStream_t st = c->stream;
  if (c->nb > 0) {
    if (c->dim == 3) k_to_bsr<3><<<nblk((int64_t)c->nb * dd), 256, 0, st>>>(c->blockptr.p, c->brow.p, c->nb, c->Avals.p, M_vals_dev, tmp.p);
    else k_to_bsr<2><<<nblk((int64_t)c->nb * dd), 256, 0, st>>>(c->blockptr.p, c->brow.p, c->nb, c->Avals.p, M_vals_dev, tmp.p);
    c->launches += 1;
  }
  MFB_CUDA(cudaMemcpyAsync(ptr, c->blockptr.p, ((size_t)c->nF + 1) * 4, cudaMemcpyDeviceToHost, st));
  if (c->nb > 0) {
    MFB_CUDA(cudaMemcpyAsync(col, c->bcol.p, (size_t)c->nb * 4, cudaMemcpyDeviceToHost, st));
    MFB_CUDA(cudaMemcpyAsync(val, tmp.p, (size_t)c->nb * dd * 8, cudaMemcpyDeviceToHost, st));
  }
  MFB_CUDA(cudaStreamSynchronize(st));
  MFB_CUDA(cudaGetLastError());
  return MFB_OK;
}
