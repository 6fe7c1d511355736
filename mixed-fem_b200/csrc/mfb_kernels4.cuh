// k_update_tet_v4 : MixedStretch<3>::update() (mixed_stretch.cpp:158-243, 345-366) in ONE launch per step, plus a
// tail launch over the (normally empty or short) list of elements whose Hessian needs the eigen-decomposition.
//
// Per element (one thread):
//   1. F = Ds Dm^-1, warm-started Newton polar, R_, S_, grad_la_, dSdF_; B = (Js Jloc)^T goes to SHARED memory
//      (54 doubles per thread, thread-private column of a [54][128] tile: conflict-free, no barriers);
//   2. g = h^2 dpsi, H = h^2 d2psi, psd_fix_invert (psd_fix.h:10-34):
//        fast path  (psd_fast6): H has no eigenvalue below tol => fixed H is H, H^-1 by Cholesky, lambda_max by
//                   tridiagonal Newton.  ~1/6 of the FP64 work of the eigen-decomposition;
//        otherwise  the element index is appended to a list and the thread goes idle; k_update_tet_v4_tail redoes
//                   step 1 from the already converged R_ (no iteration), runs the Jacobi eigen-solver and continues;
//   3. u = Hfix (S - s) + g, gl_; vertex vectors grad_x_ / rhs_; B <- sqrt(vol) B L in place (Hfix = L L^T);
//      3x3 blocks B_j B_k^T of the element matrix for the 10 local vertex pairs j <= k.
//
// Assembly modes (template ASM):
//   0  one scattered FP64 RED per value into the reference-layout CSR values, upper triangle only, then k_mirror_lower;
//   1  WARP-COOPERATIVE reductions into an internal block-CSR copy of the upper triangle (9 contiguous doubles per
//      block): each lane parks its 3x3 block in a per-warp shared-memory buffer, then the 32 lanes re-read the 288
//      values so that 9 ADJACENT lanes cover the 9 CONTIGUOUS doubles of one block and issue 9 RED instructions for the
//      32 blocks.  Same number of RED instructions, but the LSU coalesces each into a few 72-byte runs.  Vertex
//      vectors the same way (6 contiguous doubles per vertex: grad_x | rhs interleaved).  k_finalize_csr then writes
//      every value of the reference-layout scalar CSR (both triangles), k_deinterleave_vec the two vertex vectors.
//      Measured on B200 (tools/ubench_red.cu, L2-resident targets): scattered RED.F64 179 G values/s; lane groups of
//      8 contiguous doubles 500 G values/s; 80-byte cp.reduce.async.bulk 57 G ops/s, but each bulk operation is a
//      warp-uniform instruction that ptxas serialises over the lanes (ELECT/R2UR/UBLKRED loop, ~9 issue slots per
//      operation) - tried, 6.3 ms per 10M tets against 7.3 ms for mode 0, replaced by this mode.
#pragma once
#include "mfb_kernels.cuh"

namespace mfb {

#ifndef MFB_V4_THREADS
#define MFB_V4_THREADS 128
#endif
#ifndef MFB_V4_MINB
#define MFB_V4_MINB 2      // CTAs per SM.  2 x 128 threads at 231 registers, no spills: 4.17 ms per 10M tets; 3 x 128 at 168 registers with
#endif                     // 96 B of spills: 4.36 ms (round 2, after the late loads moved to cp.async; round 1 measured them equal)
#ifndef MFB_V4_ROLL
#define MFB_V4_ROLL 1      // step-3 loops over vertices / block pairs stay rolled: 6984 -> 5944 SASS instructions, 5.31 -> 5.21 ms
#endif                      // per 10M tets (straight-line code of this size pays in instruction-cache misses)
#ifndef MFB_V4_ROLL_DSDF
#define MFB_V4_ROLL_DSDF 0
#endif
// register cap of the three element kernels: __launch_bounds__(threads, MINB) by default; ptxas only knows a few register
// levels for it (255 / 168 / 128 ...), so experiments with 9 or 10 warps per SM name the cap itself (MFB_V4_MAXNREG)
// Experiment (off): CTA barriers at the phase boundaries of the one-tile kernel keep its warps in the same region of the
// ~95 KB instruction stream (shared instruction-cache lines).
#ifndef MFB_V4_SYNC
#define MFB_V4_SYNC 0
#endif
#ifndef MFB_V4_TAIL_SYNC
#define MFB_V4_TAIL_SYNC 0
#endif
#define MFB_V4_SYNCPT() do { if (MFB_V4_SYNC && !TAIL && !PIPE) __syncthreads(); } while (0)
#ifdef MFB_V4_MAXNREG
#define MFB_V4_BOUNDS __maxnreg__(MFB_V4_MAXNREG)
#else
#define MFB_V4_BOUNDS __launch_bounds__(kV4Threads, MFB_V4_MINB)
#endif
constexpr int kV4Threads = MFB_V4_THREADS;
constexpr int kV4Warps = kV4Threads / 32;
// per warp: [0,288) reduction staging (one 3x3 block per lane; la[6][32] parks here until step 3),
// [288,384) the 6 element->block codes per lane (int[6][32]), [384,448) fidx of the 4 local vertices (int[4][32]).
// Codes and fidx stay in shared memory for the whole kernel: no registers, no spills, and their global loads are
// cp.async issued at the top (ncu r1k: the late fidx loads + their spill stores were 10 % of all stall samples).
constexpr int kV4SinkSlots = 128;          // x 32 doubles behind Ablk and behind gxrh
constexpr int kV4RedDoubles = 32 * 9;
constexpr int kV4StageDoubles = kV4RedDoubles + 96 + 64;
constexpr int kV4SmemB = 54 * kV4Threads * (int)sizeof(double);
constexpr int kV4SmemStage = kV4Warps * (kV4StageDoubles * (int)sizeof(double) + 32 * (int)sizeof(int));

// Pipelined (persistent) variant: per thread one staging slot for the NEXT tile's streamed inputs - dphi[9], vol, R[9],
// s[6] (25 doubles), T[4] (ints) - and one for its 12 gathered vertex coordinates, all thread-private columns
// (component-major, [c][thread]): no block-level synchronisation, cp.async completion is per thread.
constexpr int kV4PipeIn = 25;
constexpr int kV4SmemPipe = (kV4PipeIn + 12) * kV4Threads * (int)sizeof(double) + 4 * kV4Threads * (int)sizeof(int);

struct PipeStage {
  double* in;      // [25][kV4Threads]
  int* T;          // [4][kV4Threads]
  double* x;       // [12][kV4Threads]
  int e_next;      // element whose inputs are (being) staged, -1: none
};

struct DevAsm {
  double* Ablk;          // [nbu][9]  block-CSR copy of the upper triangle (row-major 3x3)
  double* gxrh;          // [nF][6]   grad_x | rhs interleaved
  const int* bmap;       // [6][ld]   pairs j < k: 2*upper_block + transposed; one vertex pinned: the free vertex's diagonal slot; both: -1
  int* slow_list;        // [nE]
  unsigned long long* slow_cnt;   // [0] entries appended, [1] entries an earlier tail launch has done (counters[5], [6], or a chunk's own pair)
  int sinkA, sinkV;      // first of 32 scratch doubles behind Ablk / gxrh (reductions without a target land there)
  double *Kelem, *Velem; // ASM 2: [nE][54] off-diagonal blocks K_e(j,k), j < k, row-major ; [nE][4][6] vertex values (grad_x | rhs)
};

__device__ __forceinline__ void cp_async8(void* sdst, const void* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"((unsigned)__cvta_generic_to_shared(sdst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async4(void* sdst, const void* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((unsigned)__cvta_generic_to_shared(sdst)), "l"(gsrc) : "memory");
}

// base[idx] += val when guard >= 0, as ONE predicated RED with a 32-bit index (9 * #blocks < 2^31 is implied by the
// int32 CSR of the reference).  Written in PTX because the C++ form `if (t >= 0) atomicAdd(...)` compiles to a
// branch with a BSSY/BSYNC pair and a 64-bit carry chain per reduction: 13 instructions instead of 5, 88 times per tet.
__device__ __forceinline__ void red_add_if(double* base, int idx, double val, int guard) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t.reg .u64 a;\n\t"
      "setp.ge.s32 p, %3, 0;\n\t"
      "mad.wide.s32 a, %2, 8, %0;\n\t"
      "@p red.global.add.f64 [a], %1;\n\t}"
      ::"l"(base), "d"(val), "r"(idx), "r"(guard));
}

__device__ __forceinline__ void red_add_idx(double* base, int idx, double val) {
  asm volatile(
      "{\n\t.reg .u64 a;\n\t"
      "mad.wide.s32 a, %1, 8, %0;\n\t"
      "red.global.add.f64 [a], %2;\n\t}"
      ::"l"(base), "r"(idx), "d"(val));
}

// Every lane holds NV values for NV contiguous doubles at base[target .. target+NV) (target < 0: nothing to add).
// The warp re-distributes them through its shared-memory buffer so that each RED instruction covers runs of NV
// contiguous doubles with NV adjacent lanes.  All 32 lanes must call this.
// `sink`: index of 32 scratch doubles behind the real targets of `base` (one of kV4SinkSlots groups, picked by the warp's
// element range: the pinned face of a mesh sends ~2M reductions there, and FP64 reductions on ONE address retire at
// about 16 ns each - with a single group of 32 the rank holding the pinned face of the 8-GPU run took 1.5 ms instead of
// 0.6); values without a target are added there, which keeps the reduction unconditional: ncu r1k showed a BSSY/BRA/BSYNC triple and a
// branch_resolving stall around every predicated RED, and every RED waiting on the shared-memory load issued right
// before it - here all values are fetched first, then the NV reductions go out back to back.
// `live` false (the lane's element went to the Jacobi tail list, or is grid padding): the lane adds 0.0 to its own
// targets - NOT to the sink, which would serialise a third of all reductions on 32 addresses when many Hessians are
// indefinite (measured: 50 ms instead of 9.5 ms per 10M tets at 30 % tail elements).
template <int NV>
__device__ __forceinline__ void coop_red(double* __restrict__ base, const double* v, int target, bool live, double* sW, int* sC, int lane, int sink) {
  __syncwarp();
#pragma unroll
  for (int i = 0; i < NV; ++i) sW[lane * NV + i] = live ? v[i] : 0.0;
  sC[lane] = target;
  __syncwarp();
  double val[NV];
  int idx[NV];
#pragma unroll
  for (int it = 0; it < NV; ++it) {
    const int q = it * 32 + lane;
    const int b = q / NV, i = q - b * NV;
    const int t = sC[b];
    val[it] = sW[q];
    idx[it] = (t >= 0) ? t + i : sink + lane;       // sink: already offset by the caller's warp slot (kV4SinkSlots x 32 doubles)
  }
#pragma unroll
  for (int it = 0; it < NV; ++it) red_add_idx(base, idx[it], val[it]);
}

// Per-element state is written once per update and not read again by this kernel: streaming (evict-first) stores keep
// the L2 for the reduction targets (the blocks of one z-plane of a cube-major mesh are touched again a plane later).
template <class T>
__device__ __forceinline__ void st_stream(T* p, T v) { __stcs(p, v); }

// local pair (j,k), j < k  ->  0..5
__device__ __forceinline__ constexpr int pair6(int j, int k) { return j * 3 - j * (j - 1) / 2 + (k - j - 1); }

// `inrange` false: the lane has no element (grid padding); it runs on a clamped element index so that the warp stays
// convergent for the cooperative reductions, and stores nothing.
// streamed inputs of element e -> the thread's staging slot (25 x 8-byte + 4 x 4-byte cp.async)
__device__ __forceinline__ void pipe_issue_inputs(const DevMesh& m, const DevState& st, const PipeStage& ps, int e) {
  const int64_t ld = m.ld;
  const int t = threadIdx.x;
#pragma unroll
  for (int k = 0; k < 4; ++k) cp_async4(&ps.T[k * kV4Threads + t], &m.T[k * ld + e]);
#pragma unroll
  for (int i = 0; i < 9; ++i) cp_async8(&ps.in[i * kV4Threads + t], &m.dphi[i * ld + e]);
  cp_async8(&ps.in[9 * kV4Threads + t], &m.vol[e]);
#pragma unroll
  for (int i = 0; i < 9; ++i) cp_async8(&ps.in[(10 + i) * kV4Threads + t], &st.R[i * ld + e]);
#pragma unroll
  for (int k = 0; k < 6; ++k) cp_async8(&ps.in[(19 + k) * kV4Threads + t], &st.s[k * ld + e]);
}
// the 12 vertex coordinates of the staged element (its T has landed) -> the thread's x slot
__device__ __forceinline__ void pipe_issue_gather(const PipeStage& ps, const double* __restrict__ x) {
  const int t = threadIdx.x;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int v = ps.T[k * kV4Threads + t];
#pragma unroll
    for (int l = 0; l < 3; ++l) cp_async8(&ps.x[(3 * k + l) * kV4Threads + t], &x[3 * (int64_t)v + l]);
  }
}

template <bool FUSE_RHS, bool TAIL, int ASM, bool PIPE = false>
__device__ __forceinline__ void update_tet_v4_body(const DevMesh& m, const DevState& st, const DevAsm& as, const double* __restrict__ x,
                                                   int e, bool inrange, double h2, double* sB, double* sW, int* sC,
                                                   PipeStage ps = PipeStage{nullptr, nullptr, nullptr, -1}) {
  double* myB = sB + threadIdx.x;                       // element (row, n) at myB[(6*row + n) * kV4Threads]
  const int lane = threadIdx.x & 31;
  const int64_t ld = m.ld;
  int status = 0, iters = 0, sweeps = 0;

  // ---- 1 : deformation gradient, polar, stretch, dS/dF
  double dphi[9], F[9], R[9], Y[9];
  double s[6];
  double vol;
  if constexpr (PIPE) {
    // this element's inputs were staged while the previous tile computed: everything outstanding has to land first
    asm volatile("cp.async.wait_all;" ::: "memory");
    const int t = threadIdx.x;
#pragma unroll
    for (int i = 0; i < 9; ++i) dphi[i] = ps.in[i * kV4Threads + t];
    vol = ps.in[9 * kV4Threads + t];
#pragma unroll
    for (int i = 0; i < 9; ++i) R[i] = ps.in[(10 + i) * kV4Threads + t];
#pragma unroll
    for (int k = 0; k < 6; ++k) s[k] = ps.in[(19 + k) * kV4Threads + t];
    double x0[3], dx[9];
#pragma unroll
    for (int l = 0; l < 3; ++l) x0[l] = ps.x[l * kV4Threads + t];
#pragma unroll
    for (int k = 0; k < 3; ++k)
#pragma unroll
      for (int l = 0; l < 3; ++l) dx[3 * k + l] = ps.x[(3 * (k + 1) + l) * kV4Threads + t] - x0[l];
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
      for (int r = 0; r < 3; ++r) F[3 * c + r] = dx[r] * dphi[c] + dx[3 + r] * dphi[3 + c] + dx[6 + r] * dphi[6 + c];
    asm volatile("" ::: "memory");        // the slot is read: refill it with the next tile's inputs (lands during the polar iteration)
    if (ps.e_next >= 0) pipe_issue_inputs(m, st, ps, ps.e_next);
  } else {
    tet_F(m, e, x, dphi, F);
#pragma unroll
    for (int i = 0; i < 9; ++i) R[i] = __ldcs(&st.R[i * ld + e]);
#pragma unroll
    for (int k = 0; k < 6; ++k) s[k] = __ldcs(&st.s[k * ld + e]);
    vol = __ldcs(&m.vol[e]);
  }
  // step-3 inputs (la, element->block codes) start moving now, asynchronously and without registers, into the warp's
  // reduction buffer, which is idle until step 3: layout la[n] at sW[n*32 + lane], code[q] at int(sW + 192)[q*32 + lane]
  if (ASM >= 1) {
    __syncwarp();     // tail kernel: the previous element's last reduction round has finished reading the buffer
#pragma unroll
    for (int n = 0; n < 6; ++n) cp_async8(&sW[n * 32 + lane], &st.la[n * ld + e]);
    int* sI = reinterpret_cast<int*>(sW + kV4RedDoubles);
#pragma unroll
    for (int q = 0; q < 6; ++q) cp_async4(&sI[q * 32 + lane], &as.bmap[q * ld + e]);
    int* sF = reinterpret_cast<int*>(sW + kV4RedDoubles + 96);
#pragma unroll
    for (int k = 0; k < 4; ++k) cp_async4(&sF[k * 32 + lane], &m.fidx[k * ld + e]);
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  if (!TAIL) {
    iters = procrustes3(R, F, Y, 1e-6, 100, &status);
    if (inrange) {
#pragma unroll
      for (int i = 0; i < 9; ++i) st_stream(&st.R[i * ld + e], R[i]);
      st_stream(&st.iters[e], iters);
    }
  } else {
#pragma unroll
    for (int c = 0; c < 3; ++c)
#pragma unroll
      for (int r = 0; r < 3; ++r) Y[3 * c + r] = R[r] * F[c] + R[3 + r] * F[3 + c] + R[6 + r] * F[6 + c];
  }
  MFB_V4_SYNCPT();
  // The block targets of step 3 are known (the codes have landed long ago): pull their lines into the L2 now.
  // An FP64 reduction whose line is not in the L2 waits for the DRAM fill and, behind it, so does everything else in
  // the SM's memory pipe: without this the reductions cost 1.6 ms per 10M tets, with it 0.7 ms.  (The vertex targets
  // are L2-resident already; prefetching them, a second line per block, or the next CTAs' inputs measured slower.)
  if (ASM == 2) asm volatile("cp.async.wait_all;" ::: "memory");
  if (ASM == 1) {
    asm volatile("cp.async.wait_all;" ::: "memory");
    if constexpr (PIPE) {              // the next tile's T has landed with the rest: start its vertex gather now
      if (ps.e_next >= 0) { pipe_issue_gather(ps, x); asm volatile("cp.async.commit_group;" ::: "memory"); }
    }
    const int* sI = reinterpret_cast<const int*>(sW + kV4RedDoubles);
#pragma unroll
    for (int q = 0; q < 6; ++q) {
      const int cd = sI[q * 32 + lane];
      if (cd >= 0) asm volatile("prefetch.global.L2 [%0];" ::"l"(as.Ablk + 9 * (int64_t)(cd >> 1)));
    }
  }
  double u[6];    // S - s first, then Hfix (S - s) + g
  {
    double S[6];
    stretch6(R, F, S);
#pragma unroll
    for (int k = 0; k < 6; ++k) u[k] = S[k] - s[k];
    if (!TAIL) {
      if (inrange) {
#pragma unroll
        for (int k = 0; k < 6; ++k) st_stream(&st.S[k * ld + e], S[k]);
      }
      // evaluate_constraint's own polar from a copy of R_ (mixed_stretch.cpp:358-360): zero iterations when the
      // iterated Y = R F^T already passes the tolerance test, which it does unless the iteration cap was hit
      const double q0 = Y[5] - Y[7], q1 = Y[6] - Y[2], q2 = Y[1] - Y[3];
      if (!(sqrt(q0 * q0 + q1 * q1 + q2 * q2) < 1e-6)) {
        M33 Rc, Fc;
#pragma unroll
        for (int i = 0; i < 9; ++i) { Rc.a[i] = R[i]; Fc.a[i] = F[i]; }
        const V6 r = polar_from_copy_slow(Rc, Fc);
#pragma unroll
        for (int k = 0; k < 6; ++k) S[k] = r.a[k];
      }
      if (inrange) {
#pragma unroll
        for (int k = 0; k < 6; ++k) st_stream(&st.grad_la[k * ld + e], ((k < 3) ? 1.0 : 2.0) * (S[k] - s[k]) * vol);
      }
    }
  }
  {
    double Hinv3[6];
    if (!procrustes_hinv(Y, Hinv3)) status |= ST_POLAR_HFAIL;
#if MFB_V4_ROLL_DSDF
#pragma unroll 1
#else
#pragma unroll
#endif
    for (int n = 0; n < 6; ++n) {
      double col[9];
#if MFB_V4_ROLL_DSDF
      dsdf3_column_rt(n, R, F, Hinv3, col);
#else
      dsdf3_column(n, R, F, Hinv3, col);
#endif
      if (!TAIL && inrange && !(m.exp & 1)) {
        const double sy = (n < 3) ? 1.0 : 2.0;
#pragma unroll
        for (int k = 0; k < 9; ++k) st_stream(&st.dSdF[(n * 9 + k) * ld + e], sy * col[k]);
      }
#pragma unroll
      for (int k = 0; k < 3; ++k)
#pragma unroll
        for (int r = 0; r < 3; ++r)
          myB[(6 * (3 * k + r) + n) * kV4Threads] = dphi[3 * k] * col[r] + dphi[3 * k + 1] * col[3 + r] + dphi[3 * k + 2] * col[6 + r];
      if (n == 0 && !(col[0] == col[0])) status |= ST_NONFINITE;
    }
  }

  MFB_V4_SYNCPT();
  // ---- 2 : material gradient / Hessian, PSD fix, inverse, lambda_max
  bool live = inrange;       // false: this lane stores / adds nothing from here on
  double L[21];
  {
    const MaterialDev mat = m.matid ? m.mats[m.matid[e]] : m.mat0;
    double g[6], Hinv[21], lmax;
    material_grad_hess6(mat.model, mat.mu, mat.la, s, h2, g, L);     // L holds H for now
    if (!TAIL) {
      // H_ and Hinv_ leave as soon as they exist: a lane that later turns out to need the eigen-decomposition has
      // them overwritten by the tail kernel, and the 42 values do not stay live across the lambda_max iteration
      double Hm[21];
#pragma unroll
      for (int i = 0; i < 21; ++i) Hm[i] = L[i];
      if (inrange) {
        const double iv = 1.0 / vol;
#pragma unroll
        for (int i = 0; i < 6; ++i)
#pragma unroll
          for (int j = 0; j <= i; ++j) {
            const double w = ((i < 3) ? 1.0 : 0.5) * ((j < 3) ? 1.0 : 0.5);
            st_stream(&st.H[symidx(i, j) * ld + e], iv * w * Hm[symidx(i, j)]);
          }
      }
      bool ok = psd_fast6_factor(Hm, L, Hinv, 1e-8, 1e-5);
      if (inrange && !(m.exp & 4)) {
#pragma unroll
        for (int i = 0; i < 21; ++i) st_stream(&st.Hinv[i * ld + e], Hinv[i]);
      }
      int nit;
      if (m.exp & 8) { lmax = 1.0; nit = 0; } else ok = psd_fast6_lmax(Hm, &lmax, &nit) && ok;
      // elements that need the eigen-decomposition: append to the list (one atomic per warp); the lane goes idle
      const bool fail = inrange && !ok;
      const unsigned failm = __ballot_sync(0xffffffffu, fail);
      if (failm) {
        const int leader = __ffs(failm) - 1;
        unsigned long long base = 0;
        if (lane == leader) base = atomicAdd(as.slow_cnt, (unsigned long long)__popc(failm));
        base = __shfl_sync(0xffffffffu, base, leader);
        if (fail) as.slow_list[base + __popc(failm & ((1u << lane) - 1u))] = e;
        live = live && ok;
        if (__ballot_sync(0xffffffffu, live) == 0u) {     // nothing left to do for this warp
          count_status(st.counters, inrange ? status : 0, inrange ? iters : 0, 0);
          return;
        }
      }
    } else {
      lmax = psd_fix_invert<6>(L, Hinv, 1e-8, &sweeps);
      if (live) {
        const double iv = 1.0 / vol;
#pragma unroll
        for (int i = 0; i < 6; ++i)
#pragma unroll
          for (int j = 0; j <= i; ++j) {
            const double w = ((i < 3) ? 1.0 : 0.5) * ((j < 3) ? 1.0 : 0.5);
            st_stream(&st.H[symidx(i, j) * ld + e], iv * w * L[symidx(i, j)]);
          }
      }
      cholesky6(L);
    }
    if (live && !(m.exp & 4)) {
      if (TAIL) {
#pragma unroll
        for (int i = 0; i < 21; ++i) st_stream(&st.Hinv[i * ld + e], Hinv[i]);
      }
#pragma unroll
      for (int k = 0; k < 6; ++k) st_stream(&st.g[k * ld + e], g[k]);
      st_stream(&st.evals[e], h2 * lmax);
      if (!(lmax == lmax) || lmax > 1e300) status |= ST_NONFINITE;
    }
    // u = Hfix (S - s) + g = L (L^T (S - s)) + g ; gl = Syminv u   (rhs(): gl = vol H_ Sym (S-s) + Syminv g)
    if (FUSE_RHS) {
      double t[6];
#pragma unroll
      for (int j = 0; j < 6; ++j) { double a = 0.0;
#pragma unroll
        for (int i = j; i < 6; ++i) a += L[symidx(i, j)] * u[i];
        t[j] = a; }
#pragma unroll
      for (int i = 0; i < 6; ++i) { double a = g[i];
#pragma unroll
        for (int j = 0; j <= i; ++j) a += L[symidx(i, j)] * t[j];
        u[i] = a; if (live) st_stream(&st.gl[i * ld + e], ((i < 3) ? 1.0 : 0.5) * a); }
    }
  }

  MFB_V4_SYNCPT();
  // ---- 3 : vertex vectors, and B <- sqrt(vol) B L in place (one row at a time through registers)
  int f[4];
  double tl[6];
  const int sink_ofs = ((e >> 5) & (kV4SinkSlots - 1)) * 32;
  const int* sI = reinterpret_cast<const int*>(sW + kV4RedDoubles);          // codes, fidx: read where they are used
  const int* sF = reinterpret_cast<const int*>(sW + kV4RedDoubles + 96);
  if (ASM >= 1) {
    asm volatile("cp.async.wait_all;" ::: "memory");
#pragma unroll
    for (int n = 0; n < 6; ++n) tl[n] = ((n < 3) ? 1.0 : 2.0) * sW[n * 32 + lane];
  } else {
#pragma unroll
    for (int k = 0; k < 4; ++k) f[k] = live ? m.fidx[k * ld + e] : -1;
#pragma unroll
    for (int n = 0; n < 6; ++n) tl[n] = ((n < 3) ? 1.0 : 2.0) * st.la[n * ld + e];
  }
  const double sv = sqrt(vol);
  double B0[3][6];
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int n = 0; n < 6; ++n) B0[r][n] = 0.0;
  {
    double v0[6] = {0, 0, 0, 0, 0, 0};
#if MFB_V4_ROLL
#pragma unroll 1
#else
#pragma unroll
#endif
    for (int k = 0; k < 3; ++k) {
      const int fk = (ASM >= 1) ? sF[(k + 1) * 32 + lane] : ((k == 0) ? f[1] : (k == 1) ? f[2] : f[3]);
      double vv[6];
#pragma unroll
      for (int r = 0; r < 3; ++r) {
        double row[6];
#pragma unroll
        for (int n = 0; n < 6; ++n) row[n] = myB[(6 * (3 * k + r) + n) * kV4Threads];
        double a = 0.0, b = 0.0;
#pragma unroll
        for (int n = 0; n < 6; ++n) { a += row[n] * tl[n]; if (FUSE_RHS) b += row[n] * u[n]; }
        a *= vol; b *= -vol;
        v0[r] -= a; v0[3 + r] -= b;
        vv[r] = a; vv[3 + r] = b;
#pragma unroll
        for (int j = 0; j < 6; ++j) {
          double t = 0.0;
#pragma unroll
          for (int i = j; i < 6; ++i) t += row[i] * L[symidx(i, j)];
          t *= sv;
          myB[(6 * (3 * k + r) + j) * kV4Threads] = t;
          B0[r][j] -= t;
        }
      }
      if (ASM == 0) {
        if (fk >= 0) {
#pragma unroll
          for (int r = 0; r < 3; ++r) {
            red_add(&st.grad_x[3 * (int64_t)fk + r], vv[r]);
            if (FUSE_RHS) red_add(&st.rhs[3 * (int64_t)fk + r], vv[3 + r]);
          }
        }
      } else if (ASM == 2) {          // deterministic mode: the element's own slot, summed per vertex by k_det_gather_vec
        if (live) {
#pragma unroll
          for (int i = 0; i < 6; ++i) as.Velem[(int64_t)e * 24 + (k + 1) * 6 + i] = vv[i];
        }
      } else {
        if (!(m.exp & 2)) coop_red<6>(as.gxrh, vv, (fk >= 0 && inrange) ? 6 * fk : -1, live, sW, sC, lane, as.sinkV + sink_ofs);
      }
    }
    if (ASM == 0) {
      if (f[0] >= 0) {
#pragma unroll
        for (int r = 0; r < 3; ++r) {
          red_add(&st.grad_x[3 * (int64_t)f[0] + r], v0[r]);
          if (FUSE_RHS) red_add(&st.rhs[3 * (int64_t)f[0] + r], v0[3 + r]);
        }
      }
    } else if (ASM == 2) {
      if (live) {
#pragma unroll
        for (int i = 0; i < 6; ++i) as.Velem[(int64_t)e * 24 + i] = v0[i];
      }
    } else {
      const int f0 = sF[lane];
      if (!(m.exp & 2)) coop_red<6>(as.gxrh, v0, (f0 >= 0 && inrange) ? 6 * f0 : -1, live, sW, sC, lane, as.sinkV + sink_ofs);
    }
  }

  MFB_V4_SYNCPT();
  // ---- 3 : matrix blocks for the local pairs j <= k
  if (ASM == 0) {
    int stride[4], sidx[16];
#pragma unroll
    for (int k = 0; k < 4; ++k) stride[k] = m.sstride[k * ld + e];
#pragma unroll
    for (int q = 0; q < 16; ++q) sidx[q] = live ? m.smap[q * ld + e] : -1;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      double bj[3][6];
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int n = 0; n < 6; ++n) bj[r][n] = (j == 0) ? B0[r][n] : myB[(6 * (3 * (j - 1) + r) + n) * kV4Threads];
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
          for (int n = 0; n < 6; ++n) bk[n] = (k == j) ? bj[c][n] : myB[(6 * (3 * (k - 1) + c) + n) * kV4Threads];
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
  } else {
    // off-diagonal pairs only: the diagonal blocks follow from the zero block-row sums in k_finalize_rows
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      double bj[3][6];
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int n = 0; n < 6; ++n) bj[r][n] = (j == 0) ? B0[r][n] : myB[(6 * (3 * (j - 1) + r) + n) * kV4Threads];
#if MFB_V4_ROLL
#pragma unroll 1
#else
#pragma unroll
#endif
      for (int k = j + 1; k < 4; ++k) {
        const int cd = sI[pair6(j, k) * 32 + lane];
        double a[3][3];      // a[r][c] = (B_j B_k^T)(r,c)
#pragma unroll
        for (int c = 0; c < 3; ++c) {
          double bk[6];
#pragma unroll
          for (int n = 0; n < 6; ++n) bk[n] = myB[(6 * (3 * (k - 1) + c) + n) * kV4Threads];
#pragma unroll
          for (int r = 0; r < 3; ++r) {
            double t = 0.0;
#pragma unroll
            for (int n = 0; n < 6; ++n) t += bj[r][n] * bk[n];
            a[r][c] = t;
          }
        }
        if (ASM == 2) {             // deterministic mode: K_e(j,k) as computed, row-major, gathered per CSR block later
          if (live) {
            double* dst = as.Kelem + (int64_t)e * 54 + pair6(j, k) * 9;
#pragma unroll
            for (int r = 0; r < 3; ++r)
#pragma unroll
              for (int c = 0; c < 3; ++c) dst[3 * r + c] = a[r][c];
          }
          continue;
        }
        const bool tr = (cd & 1) != 0;     // the stored block is (g_k, g_j): park the transpose
        double v[9];
#pragma unroll
        for (int r = 0; r < 3; ++r)
#pragma unroll
          for (int c = 0; c < 3; ++c) v[3 * r + c] = tr ? a[c][r] : a[r][c];
        if (!(m.exp & 2)) coop_red<9>(as.Ablk, v, (cd >= 0 && inrange) ? 9 * (cd >> 1) : -1, live, sW, sC, lane, as.sinkA + sink_ofs);
      }
    }
  }
  count_status(st.counters, inrange ? status : 0, inrange ? iters : 0, inrange ? sweeps : 0);
}

template <bool FUSE_RHS, int ASM>
__global__ void MFB_V4_BOUNDS k_update_tet_v4(DevMesh m, DevState st, DevAsm as, const double* __restrict__ x, double h2,
                                                                 int e_begin) {
  extern __shared__ __align__(128) double smem4[];
  const int e = e_begin + blockIdx.x * blockDim.x + threadIdx.x;      // e_begin: a multiple of the CTA size (chunked launches)
  const int w = threadIdx.x >> 5;
  double* sW = smem4 + 54 * kV4Threads + w * kV4StageDoubles;
  int* sC = reinterpret_cast<int*>(smem4 + 54 * kV4Threads + kV4Warps * kV4StageDoubles) + w * 32;
  if ((e & ~31) >= m.nE) return;                       // whole warp out of range
  update_tet_v4_body<FUSE_RHS, false, ASM>(m, st, as, x, e < m.nE ? e : m.nE - 1, e < m.nE, h2, smem4, sW, sC);
}

// Persistent, software-pipelined form of k_update_tet_v4 (ASM 1): each CTA walks the element tiles
// e_begin + (blockIdx.x + i gridDim.x) 128 < e_end and, while it computes tile i, cp.async moves tile i+1's streamed inputs
// (at the top) and its 12 gathered vertex coordinates (after the polar iteration, when tile i+1's vertex ids have landed)
// into thread-private shared-memory slots.  ncu (profiles/README.md, round 2): 12.8 % of the stall samples of the plain
// kernel sit on the two chained global latencies at its top (T, then x[T]), which 2 warps per scheduler cannot cover.
template <bool FUSE_RHS>
__global__ void MFB_V4_BOUNDS k_update_tet_v4_pipe(DevMesh m, DevState st, DevAsm as, const double* __restrict__ x, double h2,
                                                                      int e_begin, int e_end) {
  extern __shared__ __align__(128) double smem4[];
  const int w = threadIdx.x >> 5;
  double* sW = smem4 + 54 * kV4Threads + w * kV4StageDoubles;
  int* sC = reinterpret_cast<int*>(smem4 + 54 * kV4Threads + kV4Warps * kV4StageDoubles) + w * 32;
  PipeStage ps;
  ps.in = reinterpret_cast<double*>(reinterpret_cast<char*>(smem4) + kV4SmemB + kV4SmemStage);
  ps.x = ps.in + kV4PipeIn * kV4Threads;
  ps.T = reinterpret_cast<int*>(ps.x + 12 * kV4Threads);
  const int last = e_end - 1;
  const int ntiles = (e_end - e_begin + kV4Threads - 1) / kV4Threads;
  int tile = blockIdx.x;
  if (tile >= ntiles) return;
  // prologue: the first tile's inputs, then its gather (lanes past the end work on a clamped element and store nothing)
  {
    const int e0 = e_begin + tile * kV4Threads + threadIdx.x;
    ps.e_next = e0 <= last ? e0 : last;
    pipe_issue_inputs(m, st, ps, ps.e_next);
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_all;" ::: "memory");
    pipe_issue_gather(ps, x);
    asm volatile("cp.async.commit_group;" ::: "memory");
  }
  for (; tile < ntiles; tile += gridDim.x) {
    // re-align the CTA's warps once per tile: nothing else synchronises them, and warps that drift into different phases
    // of the ~100 KB instruction stream thrash the instruction cache (ncu: no_instruction 4.6 % -> 22 % of the samples
    // without this barrier; the one-tile kernel keeps its warps together simply because they start together)
    __syncthreads();
    const int e = e_begin + tile * kV4Threads + threadIdx.x;
    const int nxt = tile + gridDim.x;
    const int en = e_begin + nxt * kV4Threads + threadIdx.x;
    ps.e_next = nxt < ntiles ? (en <= last ? en : last) : -1;
    if ((e & ~31) > last) {           // whole warp past the end (last tile only): nothing staged for a next tile either
      asm volatile("cp.async.wait_all;" ::: "memory");
      continue;
    }
    update_tet_v4_body<FUSE_RHS, false, 1, true>(m, st, as, x, e <= last ? e : last, e <= last, h2, smem4, sW, sC, ps);
  }
}

// elements the fast PSD path refused: grid-stride over the list the main kernel built (count in counters[5])
template <bool FUSE_RHS, int ASM>
__global__ void MFB_V4_BOUNDS k_update_tet_v4_tail(DevMesh m, DevState st, DevAsm as, const double* __restrict__ x, double h2) {
  extern __shared__ __align__(128) double smem4[];
  const int w = threadIdx.x >> 5;
  double* sW = smem4 + 54 * kV4Threads + w * kV4StageDoubles;
  int* sC = reinterpret_cast<int*>(smem4 + 54 * kV4Threads + kV4Warps * kV4StageDoubles) + w * 32;
  // list entries [counters[6], counters[5]): counters[6] is 0 unless an earlier tail launch of the same update() has
  // already processed a prefix (interface elements first, see mfb_update_dev / k_mark_tail)
  const long long n = (long long)as.slow_cnt[0];
#if MFB_V4_TAIL_SYNC
  // experiment: a CTA barrier per trip keeps the four warps in the same region of the instruction stream (the persistent
  // pipelined kernel lost 22 % of its samples to instruction-cache misses once its warps drifted apart).  The trip count is
  // CTA-uniform; whole warps past the end skip the body but stay in the loop.
  for (long long b0 = (long long)as.slow_cnt[1] + (long long)blockIdx.x * blockDim.x; b0 < n; b0 += (long long)gridDim.x * blockDim.x) {
    __syncthreads();
    const long long i0 = b0 + (threadIdx.x & ~31);
    if (i0 < n) {
      const long long i = i0 + (threadIdx.x & 31);
      const bool in = i < n;
      update_tet_v4_body<FUSE_RHS, true, ASM>(m, st, as, x, as.slow_list[in ? i : n - 1], in, h2, smem4, sW, sC);
    }
  }
#else
  for (long long i0 = (long long)as.slow_cnt[1] + (long long)blockIdx.x * blockDim.x + (threadIdx.x & ~31); i0 < n; i0 += (long long)gridDim.x * blockDim.x) {
    const long long i = i0 + (threadIdx.x & 31);
    const bool in = i < n;
    update_tet_v4_body<FUSE_RHS, true, ASM>(m, st, as, x, as.slow_list[in ? i : n - 1], in, h2, smem4, sW, sC);
  }
#endif
}

// Block-CSR upper triangle -> reference-layout scalar CSR values (sparse_utils.cpp:155-163), every entry written once.
// Half a warp per block-row gi, 15 active lanes: lane l handles the scalar columns q = l, l+15, l+30, ... of the row
// (so its column-in-block c = l % 3 is fixed) for the three scalar rows r.  Off-diagonal block (gi,gj): read from the
// upper-triangle copy (transposed when gj < gi), write, and accumulate the row sum.  Diagonal block: -(row sum of the
// off-diagonal blocks + the K_e(free,pinned) blocks accumulated in its own slot), because every element matrix has
// zero block-row sums.
// `rows` (may be null: all rows 0 .. nF-1 in order) lists the nF block-rows of this launch - the staged finalize of
// mfb_update_dev sends one launch per element chunk, over the rows that are final with that chunk.
static __global__ void __launch_bounds__(256) k_finalize_rows(const int* __restrict__ blockptr, const int* __restrict__ diagblk, const int* __restrict__ fcode,
                                                              const int* __restrict__ rows, int nF, const double* __restrict__ Ablk, double* __restrict__ A) {
  const int64_t gt = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int idx = (int)(gt >> 4), l = threadIdx.x & 15;
  const bool rowok = idx < nF;
  const int gi = rowok ? (rows ? rows[idx] : idx) : 0;
  int p0 = 0, nbr = 0, dslot = -1;
  if (rowok) { p0 = blockptr[gi]; nbr = blockptr[gi + 1] - p0; dslot = diagblk[gi] - p0; }
  const int rowlen = 3 * nbr, c = l % 3;
  const int64_t D0 = 9 * (int64_t)p0;
  double acc[3] = {0.0, 0.0, 0.0}, dval[3] = {0.0, 0.0, 0.0};
  int qd = -1;
  if (l < 15) {
    for (int q = l; q < rowlen; q += 15) {
      const int slot = q / 3;
      const int code = __ldg(&fcode[p0 + slot]);
      const double* src = Ablk + 9 * (int64_t)(code >> 1);
      if (slot == dslot) {
        qd = q;
#pragma unroll
        for (int r = 0; r < 3; ++r) dval[r] = src[3 * r + c];
      } else {
        const bool tr = code & 1;
#pragma unroll
        for (int r = 0; r < 3; ++r) {
          const double v = src[tr ? 3 * c + r : 3 * r + c];
          __stcs(&A[D0 + (int64_t)r * rowlen + q], v);      // streamed out: the block-CSR lines are the ones worth caching
          acc[r] += v;
        }
      }
    }
  }
#pragma unroll
  for (int r = 0; r < 3; ++r) {
    double t = acc[r] + __shfl_down_sync(0xffffffffu, acc[r], 3, 16);
    t += __shfl_down_sync(0xffffffffu, t, 6, 16);
    t += __shfl_down_sync(0xffffffffu, acc[r], 12, 16);
    const double tot = __shfl_sync(0xffffffffu, t, c, 16);       // lanes 0..2 hold the totals of columns c = 0..2
    if (qd >= 0) __stcs(&A[D0 + (int64_t)r * rowlen + qd], -(dval[r] + tot));
  }
}

// staged finalize: the per-chunk counts of slow elements -> counters[5] (what mfb_status reports)
static __global__ void k_fold_chunk_counts(const unsigned long long* __restrict__ cnt, int n, unsigned long long* counters) {
  unsigned long long t = 0;
  for (int k = threadIdx.x; k < n; k += 32) t += cnt[2 * k];
  for (int o = 16; o; o >>= 1) t += __shfl_down_sync(0xffffffffu, t, o);
  if (threadIdx.x == 0) { counters[5] = t; counters[6] = t; }
}

static __global__ void k_mark_tail(unsigned long long* counters) { counters[6] = counters[5]; }

// ---- deterministic assembly (asm_mode 2): owner-computes gather, the reference's own structure (sparse_utils.cpp:124-170:
// one block-row per OpenMP iteration, contributions summed in a fixed order, no atomics).  One thread per CSR block: its
// contributions (element, local pair) in ascending element order; the diagonal pair (j,j) of an element is minus the sum
// of its three off-diagonal blocks of row j (zero block-row sums of the element matrix), taken in ascending m.
static __global__ void __launch_bounds__(256) k_det_gather_blocks(const int* __restrict__ blockptr, const int* __restrict__ brow,
                                                                  const int* __restrict__ bptr, const int* __restrict__ bent, int nb,
                                                                  const double* __restrict__ Kelem, double* __restrict__ A) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= nb) return;
  double acc[9];
#pragma unroll
  for (int i = 0; i < 9; ++i) acc[i] = 0.0;
  for (int i = bptr[b]; i < bptr[b + 1]; ++i) {
    const int p = bent[i];
    const int e = p >> 4, j = (p >> 2) & 3, k = p & 3;
    const double* K = Kelem + (int64_t)e * 54;
    if (j < k) {
      const double* q = K + pair6(j, k) * 9;
#pragma unroll
      for (int t = 0; t < 9; ++t) acc[t] += q[t];
    } else if (j > k) {
      const double* q = K + pair6(k, j) * 9;
#pragma unroll
      for (int r = 0; r < 3; ++r)
#pragma unroll
        for (int c = 0; c < 3; ++c) acc[3 * r + c] += q[3 * c + r];
    } else {
      for (int m2 = 0; m2 < 4; ++m2) {
        if (m2 == j) continue;
        if (m2 > j) {
          const double* q = K + pair6(j, m2) * 9;
#pragma unroll
          for (int t = 0; t < 9; ++t) acc[t] -= q[t];
        } else {
          const double* q = K + pair6(m2, j) * 9;
#pragma unroll
          for (int r = 0; r < 3; ++r)
#pragma unroll
            for (int c = 0; c < 3; ++c) acc[3 * r + c] -= q[3 * c + r];
        }
      }
    }
  }
  const int gi = brow[b], p0 = blockptr[gi], rowlen = 3 * (blockptr[gi + 1] - p0), rank = b - p0;
#pragma unroll
  for (int r = 0; r < 3; ++r)
#pragma unroll
    for (int c = 0; c < 3; ++c) A[9 * (int64_t)p0 + (int64_t)r * rowlen + 3 * rank + c] = acc[3 * r + c];
}
static __global__ void __launch_bounds__(256) k_det_gather_vec(const int* __restrict__ vptr, const int* __restrict__ vent, int nF,
                                                               const double* __restrict__ Velem, double* __restrict__ gx, double* __restrict__ rh, int with_rhs) {
  const int v = blockIdx.x * blockDim.x + threadIdx.x;
  if (v >= nF) return;
  double acc[6] = {0, 0, 0, 0, 0, 0};
  for (int i = vptr[v]; i < vptr[v + 1]; ++i) {
    const int q = vent[i];
    const double* s = Velem + (int64_t)(q >> 2) * 24 + (q & 3) * 6;
#pragma unroll
    for (int t = 0; t < 6; ++t) acc[t] += s[t];
  }
#pragma unroll
  for (int r = 0; r < 3; ++r) { gx[3 * (int64_t)v + r] = acc[r]; if (with_rhs) rh[3 * (int64_t)v + r] = acc[3 + r]; }
}

static __global__ void __launch_bounds__(256) k_deinterleave_vec(const double* __restrict__ gxrh, int nF, double* __restrict__ gx, double* __restrict__ rh, int with_rhs) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)nF * 3) return;
  const int64_t v = i / 3; const int r = (int)(i % 3);
  gx[i] = gxrh[6 * v + r];
  if (with_rhs) rh[i] = gxrh[6 * v + 3 + r];
}

}  // namespace mfb
