// C ABI entry points of the mixed-stretch path (see include/mfem_b200.h for the reference interface each replaces).
#include "mfb_ctx.cuh"
#include <cstdlib>
#include <algorithm>
#include <cstring>

using namespace mfb;

#define CHECK_CTX(c) do { if (!(c)) { g_last_error = "null context"; return MFB_ERR_INVALID; } cudaSetDevice((c)->device); } while (0)
#define LAUNCHED(c, n) ((c)->launches += (n))

static inline unsigned nblk(int64_t n, int t = 256) { return (unsigned)((n + t - 1) / t); }

extern "C" const char* mfb_last_error_string(void) { return g_last_error.c_str(); }
extern "C" const char* mfb_version(void) { return "mfem_b200 0.1 (sm_100a, FP64)"; }

extern "C" int mfb_set_stream(mfb_ctx* c, void* s) { CHECK_CTX(c); c->stream = (cudaStream_t)s; return MFB_OK; }

extern "C" int mfb_set_option(mfb_ctx* c, const char* name, int value) {
  CHECK_CTX(c);
  if (!name) return MFB_ERR_INVALID;
  if (!strcmp(name, "fuse_rhs")) { c->fuse_rhs = value ? 1 : 0; return MFB_OK; }
  if (!strcmp(name, "assembly")) { c->assembly = value; return MFB_OK; }
  if (!strcmp(name, "profile")) { c->profile = value ? 1 : 0; return MFB_OK; }
  if (!strcmp(name, "update_exchange")) { c->update_exchange = value ? 1 : 0; return MFB_OK; }
  if (!strcmp(name, "p2p")) { c->use_p2p = value ? 1 : 0; return MFB_OK; }
  if (!strcmp(name, "cg_graph")) { c->cg_graph = value ? 1 : 0; return MFB_OK; }
  if (!strcmp(name, "convergence_log")) { c->convergence_log = value ? 1 : 0; return MFB_OK; }
  if (!strcmp(name, "cg_persist")) { c->cg_persist = value; return MFB_OK; }      // persistent cooperative CG kernel: 2 (default) below 64 Ki block-rows, 1 always, 0 never (three launches per iteration)
  if (!strcmp(name, "peer")) { c->use_peer = value != 0; return MFB_OK; }      // 1 (default): CG halo + dot products through peer mailboxes (CUDA IPC) when every rank can open them; 0: NCCL
  if (!strcmp(name, "zero_overlap")) { c->zero_overlap = value != 0; return MFB_OK; }   // 1: double-buffered accumulator, the other half cleared on a side stream during the element kernel (default 0: no gain measured)
  if (!strcmp(name, "stage_finalize")) { c->stage_finalize = value != 0; return MFB_OK; }   // 1: k_finalize_rows per element chunk on a side stream under the next chunks (default 0: measured slower)
  if (!strcmp(name, "pipe")) { c->pipe = value; return MFB_OK; }       // 0 off, n > 0: persistent pipelined update kernel, n CTAs per SM slot
  if (!strcmp(name, "cheb_degree")) { if (value < 1 || value > 16) return MFB_ERR_INVALID; c->cheb_degree = value; return MFB_OK; }
  if (!strcmp(name, "cheb_ratio")) { if (value < 2) return MFB_ERR_INVALID; c->cheb_ratio = value; return MFB_OK; }
  if (!strcmp(name, "smem_b")) {      // 3 = k_update_tet_v4, the only tet update kernel left (0 / 1 / 2: the round-1 kernels, removed)
    if (value != 3) { g_last_error = "option smem_b: only 3 (k_update_tet_v4) exists; the round-1 kernels 0 / 1 / 2 were removed"; return MFB_ERR_INVALID; }
    c->smem_b = 3; return MFB_OK;
  }
  if (!strcmp(name, "exp")) { c->exp = value; return MFB_OK; }
  if (!strcmp(name, "cg_zero_guess")) { c->cg_zero_guess = value ? 1 : 0; return MFB_OK; }
  if (!strcmp(name, "asm_mode")) {   // v4: 1 bulk reductions into block-CSR, 0 RED into scalar CSR, 2 deterministic gather
    if (value < 0 || value > 2) return MFB_ERR_INVALID;
    c->asm_mode = value;
    return MFB_OK;
  }
  g_last_error = "unknown option";
  return MFB_ERR_INVALID;
}

extern "C" int mfb_dims(const mfb_ctx* c, int32_t* dim, int32_t* nV, int32_t* nE, int32_t* nF, int32_t* N, int64_t* nnz) {
  if (!c) return MFB_ERR_INVALID;
  if (dim) *dim = c->dim; if (nV) *nV = c->nV; if (nE) *nE = c->nE; if (nF) *nF = c->nF; if (N) *N = c->N; if (nnz) *nnz = c->nnz;
  return MFB_OK;
}

extern "C" int mfb_num_kernel_launches(mfb_ctx* c, int64_t* n, int reset) {
  if (!c || !n) return MFB_ERR_INVALID;
  *n = c->launches;
  if (reset) c->launches = 0;
  return MFB_OK;
}

extern "C" int mfb_csr_pattern(mfb_ctx* c, const int32_t** outer, const int32_t** inner, int64_t* nnz, int32_t* nrows) {
  CHECK_CTX(c);
  if (c->h_outer.empty()) {
    c->h_outer.resize((size_t)c->dim * c->nF + 1);
    c->h_inner.resize((size_t)c->nnz);
    MFB_CUDA(cudaMemcpy(c->h_outer.data(), c->outer.p, c->h_outer.size() * 4, cudaMemcpyDeviceToHost));
    if (c->nnz) MFB_CUDA(cudaMemcpy(c->h_inner.data(), c->inner.p, c->h_inner.size() * 4, cudaMemcpyDeviceToHost));
  }
  if (outer) *outer = c->h_outer.data();
  if (inner) *inner = c->h_inner.data();
  if (nnz) *nnz = c->nnz;
  if (nrows) *nrows = c->dim * c->nF;
  return MFB_OK;
}

// host layout [e][j][k] / [e][j] from the device component-major arrays
extern "C" int mfb_scatter_map(mfb_ctx* c, const int32_t** map, const int32_t** stride) {
  CHECK_CTX(c);
  if (c->h_smap.empty()) {
    const int nn = c->npe * c->npe;
    std::vector<int32_t> a((size_t)nn * c->ld), b((size_t)c->npe * c->ld);
    MFB_CUDA(cudaMemcpy(a.data(), c->smap.p, a.size() * 4, cudaMemcpyDeviceToHost));
    MFB_CUDA(cudaMemcpy(b.data(), c->sstride.p, b.size() * 4, cudaMemcpyDeviceToHost));
    c->h_smap.resize((size_t)c->nE * nn);
    c->h_sstride.resize((size_t)c->nE * c->npe);
    for (int e = 0; e < c->nE; ++e) {
      for (int q = 0; q < nn; ++q) c->h_smap[(size_t)e * nn + q] = a[(size_t)q * c->ld + e];
      for (int j = 0; j < c->npe; ++j) c->h_sstride[(size_t)e * c->npe + j] = b[(size_t)j * c->ld + e];
    }
  }
  if (map) *map = c->h_smap.data();
  if (stride) *stride = c->h_sstride.data();
  return MFB_OK;
}

static int upload_x(mfb_ctx* c, const double* x) {
  MFB_CUDA(cudaMemcpyAsync(c->xdev.p, x, (size_t)c->dim * c->nV * 8, cudaMemcpyHostToDevice, c->stream));
  return MFB_OK;
}

extern "C" int mfb_reset(mfb_ctx* c, const double* x_or_null) {
  CHECK_CTX(c);
  const double* xd = nullptr;
  if (x_or_null) { int rc = upload_x(c, x_or_null); if (rc) return rc; xd = c->xdev.p; }
  else xd = c->vref.p;   // x = vec(V_^T) (mixed_stretch.cpp:314-317); row-major Vref is the same interleaving
  MFB_CUDA(cudaMemsetAsync(c->counters.p, 0, 64, c->stream));
  if (c->kind == MFB_SHELL) k_reset<K_SHELL><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state(), xd);
  else if (c->dim == 3) k_reset<3><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state(), xd);
  else k_reset<2><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state(), xd);
  LAUNCHED(c, 1);
  MFB_CUDA(cudaGetLastError());
  MFB_CUDA(cudaMemsetAsync(c->grad_x.p, 0, ((size_t)c->dim * c->nF + 1) * 8, c->stream));
  MFB_CUDA(cudaMemsetAsync(c->rhs.p, 0, ((size_t)c->dim * c->nF + 1) * 8, c->stream));
  MFB_CUDA(cudaStreamSynchronize(c->stream));
  c->rhs_valid = false; c->updated = false;
  return MFB_OK;
}

// update(): one fused kernel (plus the zero-fill of the accumulation targets)
extern "C" int mfb_update_dev(mfb_ctx* c, const double* x_dev, double dt) {
  CHECK_CTX(c);
  if (!x_dev) return MFB_ERR_INVALID;
  cudaStream_t st = c->stream;
  const double h2 = dt * dt;
  if (c->kind == MFB_SHELL)   // rhs / solve / products re-form Jloc(x) from the positions of this update (TriMesh::update_jacobian)
    MFB_CUDA(cudaMemcpyAsync(c->xupd.p, x_dev, (size_t)c->dim * c->nV * 8, cudaMemcpyDeviceToDevice, st));
  const bool v4 = c->kind == MFB_TET && c->smem_b == 3;
  const bool bulk = v4 && c->asm_mode == 1;
  const bool det = v4 && c->asm_mode == 2;
  if (det) { int rcd = build_det_lists(c); if (rcd) return rcd; }
  const bool big = c->nE >= (1 << 18);
  const bool multi_x = bulk && c->world > 1 && c->update_exchange && c->use_p2p && c->ent_local.p && c->n_if_elems > 0 && c->n_if_elems < c->nE;
  const bool chunk_x = bulk && c->x_host_chunked && big;
  const bool staged = bulk && c->stage_finalize && c->fin_rows.p && big && !multi_x && c->pipe == 0 && c->nb > 0;
  MFB_CUDA(cudaMemsetAsync(c->counters.p, 0, 64, st));
  bool flip_ablk = false;
  if (det) {
    // every CSR value and every vertex entry is WRITTEN by the gather kernels: nothing to zero
  } else if (bulk) {
    MFB_CUDA(cudaMemsetAsync(c->gxrh.p, 0, (size_t)6 * c->nF * 8, st));
    const size_t ablk_bytes = (size_t)c->nbu * 9 * 8;
    const bool pingpong = staged || (c->zero_overlap && big && c->Ablk2.p);
    if (pingpong) {
      if (!c->fin_stream) {
        int lo = 0, hi = 0;
        MFB_CUDA(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        static const int fin_prio = getenv("MFB_FIN_PRIO") ? atoi(getenv("MFB_FIN_PRIO")) : 1;      // 1: highest priority
        MFB_CUDA(cudaStreamCreateWithPriority(&c->fin_stream, cudaStreamNonBlocking, fin_prio ? hi : lo));
        MFB_CUDA(cudaStreamCreateWithFlags(&c->zero_stream, cudaStreamNonBlocking));
        for (int k = 0; k < mfb_ctx::kChunks; ++k) MFB_CUDA(cudaEventCreateWithFlags(&c->ev_stage[k], cudaEventDisableTiming));
        MFB_CUDA(cudaEventCreateWithFlags(&c->ev_fin, cudaEventDisableTiming));
        MFB_CUDA(cudaEventCreateWithFlags(&c->ev_upd, cudaEventDisableTiming));
        for (int k = 0; k < 2; ++k) MFB_CUDA(cudaEventCreateWithFlags(&c->ev_zero[k], cudaEventDisableTiming));
      }
      const int cur = c->ablk_cur, oth = cur ^ 1;
      if (ablk_bytes) {
        if (c->ablk_clean[cur]) MFB_CUDA(cudaStreamWaitEvent(st, c->ev_zero[cur], 0));
        else MFB_CUDA(cudaMemsetAsync(c->ablk(cur), 0, ablk_bytes, st));
        // the other accumulator is cleared for the NEXT update while this one runs: its last reader (the finalize of the
        // previous update) was joined into this stream before anything queued here
        if (!c->zero_overlap) { c->ablk_clean[oth] = false; goto zero_done; }
        MFB_CUDA(cudaEventRecord(c->ev_upd, st));
        MFB_CUDA(cudaStreamWaitEvent(c->zero_stream, c->ev_upd, 0));
        MFB_CUDA(cudaMemsetAsync(c->ablk(oth), 0, ablk_bytes, c->zero_stream));
        MFB_CUDA(cudaEventRecord(c->ev_zero[oth], c->zero_stream));
        c->ablk_clean[oth] = true;
      }
      zero_done:
      c->ablk_clean[cur] = false;
      if (staged) MFB_CUDA(cudaMemsetAsync(c->chunk_cnt.p, 0, 2 * mfb_ctx::kChunks * sizeof(unsigned long long), st));
      flip_ablk = !staged;           // the staged loop flips it itself
    } else {
      // (a pending clear of this buffer on the zero stream, if any, is ordered before this one: same event)
      if (c->ablk_clean[c->ablk_cur] && c->ev_zero[c->ablk_cur]) MFB_CUDA(cudaStreamWaitEvent(st, c->ev_zero[c->ablk_cur], 0));
      else if (ablk_bytes) MFB_CUDA(cudaMemsetAsync(c->ablk(c->ablk_cur), 0, ablk_bytes, st));
      c->ablk_clean[c->ablk_cur] = false;
    }
  } else {
    MFB_CUDA(cudaMemsetAsync(c->grad_x.p, 0, (size_t)c->dim * c->nF * 8, st));
    if (c->fuse_rhs) MFB_CUDA(cudaMemsetAsync(c->rhs.p, 0, (size_t)c->dim * c->nF * 8, st));
    if (c->nnz) MFB_CUDA(cudaMemsetAsync(c->Avals.p, 0, (size_t)c->nnz * 8, st));
  }
  const DevMesh m = c->mesh();
  const DevState s = c->state();
  bool vec_recorded = false, exchanged = false;
  int slot = -1;
  if (c->profile && c->ev_count < 4096) {
    slot = c->ev_count++;
    if ((int)c->ev0.size() <= slot) {
      cudaEvent_t a, b;
      MFB_CUDA(cudaEventCreate(&a)); MFB_CUDA(cudaEventCreate(&b));
      c->ev0.push_back(a); c->ev1.push_back(b);
    }
    MFB_CUDA(cudaEventRecord(c->ev0[slot], st));
  }
  if (v4) {
    const DevAsm a = c->asmdesc();
    static const int extra_smem = getenv("MFB_V4_EXTRA_SMEM") ? atoi(getenv("MFB_V4_EXTRA_SMEM")) : 0;   // occupancy experiments
    const int smem = kV4SmemB + kV4SmemStage + extra_smem;
    // cudaFuncSetAttribute applies to the CURRENT device: one flag per device ordinal (a second context on another GPU
    // of the same process needs its own opt-in)
    static bool attr4_done_dev[64] = {false};
    bool& attr4_done = attr4_done_dev[c->device & 63];
    if (!attr4_done) {
      MFB_CUDA(cudaFuncSetAttribute(k_update_tet_v4<true, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      MFB_CUDA(cudaFuncSetAttribute(k_update_tet_v4<false, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      MFB_CUDA(cudaFuncSetAttribute(k_update_tet_v4<true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      MFB_CUDA(cudaFuncSetAttribute(k_update_tet_v4<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      MFB_CUDA(cudaFuncSetAttribute(k_update_tet_v4_tail<true, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      MFB_CUDA(cudaFuncSetAttribute(k_update_tet_v4_tail<false, 0>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      MFB_CUDA(cudaFuncSetAttribute(k_update_tet_v4_tail<true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      MFB_CUDA(cudaFuncSetAttribute(k_update_tet_v4_tail<false, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      MFB_CUDA(cudaFuncSetAttribute(k_update_tet_v4<true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      MFB_CUDA(cudaFuncSetAttribute(k_update_tet_v4<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      MFB_CUDA(cudaFuncSetAttribute(k_update_tet_v4_tail<true, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      MFB_CUDA(cudaFuncSetAttribute(k_update_tet_v4_tail<false, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
      MFB_CUDA(cudaFuncSetAttribute(k_update_tet_v4_pipe<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem + kV4SmemPipe));
      MFB_CUDA(cudaFuncSetAttribute(k_update_tet_v4_pipe<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem + kV4SmemPipe));
      // two CTAs of 110 KB only fit with the whole L1 / shared array configured as shared memory
      MFB_CUDA(cudaFuncSetAttribute(k_update_tet_v4_pipe<true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
      MFB_CUDA(cudaFuncSetAttribute(k_update_tet_v4_pipe<false>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared));
      attr4_done = true;
    }
    const int grid4 = (c->nE + kV4Threads - 1) / kV4Threads;
    const int tail_grid = grid4 < 148 * MFB_V4_MINB ? grid4 : 148 * MFB_V4_MINB;
    if (det) {
      // deterministic mode: the element kernel stores its blocks / vertex values, two gather kernels sum them in a fixed
      // order per CSR block and per vertex (no atomics anywhere: two runs give bit-identical results)
      if (c->fuse_rhs) { k_update_tet_v4<true, 2><<<grid4, kV4Threads, smem, st>>>(m, s, a, x_dev, h2, 0); k_update_tet_v4_tail<true, 2><<<tail_grid, kV4Threads, smem, st>>>(m, s, a, x_dev, h2); }
      else { k_update_tet_v4<false, 2><<<grid4, kV4Threads, smem, st>>>(m, s, a, x_dev, h2, 0); k_update_tet_v4_tail<false, 2><<<tail_grid, kV4Threads, smem, st>>>(m, s, a, x_dev, h2); }
      if (c->nF > 0) k_det_gather_vec<<<nblk(c->nF), 256, 0, st>>>(c->det_vptr.p, c->det_vent.p, c->nF, c->Velem.p, c->grad_x.p, c->rhs.p, c->fuse_rhs);
      if (c->ev_vec) { MFB_CUDA(cudaEventRecord(c->ev_vec, st)); vec_recorded = true; }
      if (c->nb > 0) k_det_gather_blocks<<<nblk(c->nb), 256, 0, st>>>(c->blockptr.p, c->brow.p, c->det_bptr.p, c->det_bent.p, c->nb, c->Kelem.p, c->Avals.p);
      LAUNCHED(c, 3);
    } else if (bulk) {
      if (chunk_x || staged) {
        // element chunks.  Host positions (mfb_update): x goes up in pieces on the copy stream and each chunk starts when
        // its vertices are there.  Staged finalize: behind each chunk, on a side stream, its Jacobi tail and the block-rows
        // that are final with it (see mfb_ctx::stage_finalize).
        const double* xh = chunk_x ? c->x_host_chunked : nullptr;
        if (xh) {
          MFB_CUDA(cudaEventRecord(c->ev_in, st));                     // everything queued so far (earlier readers of xdev)
          MFB_CUDA(cudaStreamWaitEvent(c->copy_stream, c->ev_in, 0));
        }
        static const int fin_threads = getenv("MFB_FIN_THREADS") ? atoi(getenv("MFB_FIN_THREADS")) : 256;
        int vdone = 0, nlaunch = 0;
        for (int k = 0; k < c->nchunks; ++k) {
          const int vhi = k + 1 == c->nchunks ? c->nV : c->chunk_vhi[k];
          if (xh && vhi > vdone) {
            MFB_CUDA(cudaMemcpyAsync(c->xdev.p + (size_t)3 * vdone, xh + (size_t)3 * vdone, (size_t)(vhi - vdone) * 24, cudaMemcpyHostToDevice, c->copy_stream));
            vdone = vhi;
            if (k + 1 == c->nchunks) MFB_CUDA(cudaEventRecord(c->ev_in, c->copy_stream));     // the caller's buffer is free after this
            MFB_CUDA(cudaEventRecord(c->ev_chunk[k], c->copy_stream));
            MFB_CUDA(cudaStreamWaitEvent(st, c->ev_chunk[k], 0));
          } else if (xh && k + 1 == c->nchunks) {
            MFB_CUDA(cudaEventRecord(c->ev_in, c->copy_stream));
          }
          const int e0 = c->chunk_e0[k], ne = c->chunk_e0[k + 1] - e0;
          if (ne <= 0) continue;
          const unsigned g = (unsigned)((ne + kV4Threads - 1) / kV4Threads);
          DevAsm ak = a;
          if (staged) { ak.slow_list = a.slow_list + e0; ak.slow_cnt = c->chunk_cnt.p + 2 * k; }
          if (c->fuse_rhs) k_update_tet_v4<true, 1><<<g, kV4Threads, smem, st>>>(m, s, ak, x_dev, h2, e0);
          else k_update_tet_v4<false, 1><<<g, kV4Threads, smem, st>>>(m, s, ak, x_dev, h2, e0);
          ++nlaunch;
          if (!staged) continue;
          MFB_CUDA(cudaEventRecord(c->ev_stage[k], st));
          MFB_CUDA(cudaStreamWaitEvent(c->fin_stream, c->ev_stage[k], 0));
          const unsigned gt = std::min(g, (unsigned)tail_grid);
          if (c->fuse_rhs) k_update_tet_v4_tail<true, 1><<<gt, kV4Threads, smem, c->fin_stream>>>(m, s, ak, x_dev, h2);
          else k_update_tet_v4_tail<false, 1><<<gt, kV4Threads, smem, c->fin_stream>>>(m, s, ak, x_dev, h2);
          const int nrows = c->fin_ptr[k + 1] - c->fin_ptr[k];
          if (nrows > 0)
            k_finalize_rows<<<(unsigned)(((int64_t)nrows * 16 + fin_threads - 1) / fin_threads), fin_threads, 0, c->fin_stream>>>(
                c->blockptr.p, c->diagblk.p, c->fcode.p, c->fin_rows.p + c->fin_ptr[k], nrows, a.Ablk, c->Avals.p);
          nlaunch += 2;
        }
        if (staged) {
          MFB_CUDA(cudaEventRecord(c->ev_fin, c->fin_stream));
          MFB_CUDA(cudaStreamWaitEvent(st, c->ev_fin, 0));
          k_fold_chunk_counts<<<1, 32, 0, st>>>(c->chunk_cnt.p, c->nchunks, c->counters.p);
          ++nlaunch;
          c->ablk_cur ^= 1;
        }
        LAUNCHED(c, nlaunch - 1);
        if (xh) c->x_host_chunked = nullptr;                           // consumed
      } else if (multi_x) {
        // multi-GPU with a partition: the leading INTERFACE elements first (and their share of the Jacobi tail), then the
        // point-to-point sum of their vertex vectors (grad_x | rhs, still interleaved) on a side stream while the
        // interior elements - which touch no shared vertex - run on this one
        if (!c->xstream) {
          MFB_CUDA(cudaStreamCreateWithFlags(&c->xstream, cudaStreamNonBlocking));
          MFB_CUDA(cudaEventCreateWithFlags(&c->ev_x0, cudaEventDisableTiming));
          MFB_CUDA(cudaEventCreateWithFlags(&c->ev_x1, cudaEventDisableTiming));
        }
        const int n1 = std::min(c->nE, ((c->n_if_elems + kV4Threads - 1) / kV4Threads) * kV4Threads);
        const int g1 = (n1 + kV4Threads - 1) / kV4Threads, g2 = (c->nE - n1 + kV4Threads - 1) / kV4Threads;
        if (c->fuse_rhs) { k_update_tet_v4<true, 1><<<g1, kV4Threads, smem, st>>>(m, s, a, x_dev, h2, 0); k_update_tet_v4_tail<true, 1><<<std::min(g1, tail_grid), kV4Threads, smem, st>>>(m, s, a, x_dev, h2); }
        else { k_update_tet_v4<false, 1><<<g1, kV4Threads, smem, st>>>(m, s, a, x_dev, h2, 0); k_update_tet_v4_tail<false, 1><<<std::min(g1, tail_grid), kV4Threads, smem, st>>>(m, s, a, x_dev, h2); }
        k_mark_tail<<<1, 1, 0, st>>>(c->counters.p);
        MFB_CUDA(cudaEventRecord(c->ev_x0, st));
        MFB_CUDA(cudaStreamWaitEvent(c->xstream, c->ev_x0, 0));
        { int rcx = p2p_interface_sum_on(c, c->gxrh.p, 6, nullptr, 0, nullptr, 0, c->xstream); if (rcx) return rcx; }
        MFB_CUDA(cudaEventRecord(c->ev_x1, c->xstream));
        if (g2 > 0) {
          if (c->fuse_rhs) k_update_tet_v4<true, 1><<<g2, kV4Threads, smem, st>>>(m, s, a, x_dev, h2, n1);
          else k_update_tet_v4<false, 1><<<g2, kV4Threads, smem, st>>>(m, s, a, x_dev, h2, n1);
        }
        if (c->fuse_rhs) k_update_tet_v4_tail<true, 1><<<tail_grid, kV4Threads, smem, st>>>(m, s, a, x_dev, h2);
        else k_update_tet_v4_tail<false, 1><<<tail_grid, kV4Threads, smem, st>>>(m, s, a, x_dev, h2);
        MFB_CUDA(cudaStreamWaitEvent(st, c->ev_x1, 0));
        LAUNCHED(c, 4);
        exchanged = true;
      } else if (c->pipe > 0) {
        // persistent pipelined variant: MFB_V4_MINB CTAs per SM, `pipe` waves of tiles interleaved over them
        const int gp = std::min(grid4, 148 * MFB_V4_MINB * c->pipe);
        if (c->fuse_rhs) k_update_tet_v4_pipe<true><<<gp, kV4Threads, smem + kV4SmemPipe, st>>>(m, s, a, x_dev, h2, 0, c->nE);
        else k_update_tet_v4_pipe<false><<<gp, kV4Threads, smem + kV4SmemPipe, st>>>(m, s, a, x_dev, h2, 0, c->nE);
      } else if (c->fuse_rhs) k_update_tet_v4<true, 1><<<grid4, kV4Threads, smem, st>>>(m, s, a, x_dev, h2, 0);
      else k_update_tet_v4<false, 1><<<grid4, kV4Threads, smem, st>>>(m, s, a, x_dev, h2, 0);
      if (!exchanged && !staged) {
        if (c->fuse_rhs) k_update_tet_v4_tail<true, 1><<<tail_grid, kV4Threads, smem, st>>>(m, s, a, x_dev, h2);
        else k_update_tet_v4_tail<false, 1><<<tail_grid, kV4Threads, smem, st>>>(m, s, a, x_dev, h2);
      }
      // vertex vectors first: grad_x_ / rhs_ are final after this launch (a host reader copies them out while the
      // CSR values are still being written, see mfb_rhs)
      if (c->nF > 0) k_deinterleave_vec<<<nblk((int64_t)c->nF * 3), 256, 0, st>>>(c->gxrh.p, c->nF, c->grad_x.p, c->rhs.p, c->fuse_rhs);
      if (c->world > 1 && c->update_exchange && !exchanged) {
        // no overlap with the interior elements on this path (host positions arriving in chunks): the interface sum still goes
        // BEFORE the CSR values, so that a host reader of rhs_ does not wait for k_finalize_rows
        int rcx = interface_sum(c, c->grad_x.p, c->dim, c->fuse_rhs ? c->rhs.p : nullptr, c->dim, nullptr, 0);
        if (rcx) return rcx;
        exchanged = true;
      }
      if (c->ev_vec) { MFB_CUDA(cudaEventRecord(c->ev_vec, st)); vec_recorded = true; }
      if (c->nb > 0 && !staged) k_finalize_rows<<<nblk((int64_t)c->nF * 16), 256, 0, st>>>(c->blockptr.p, c->diagblk.p, c->fcode.p, nullptr, c->nF, a.Ablk, c->Avals.p);
      if (flip_ablk) c->ablk_cur ^= 1;          // the next update reduces into the half that is being cleared beside this one
      LAUNCHED(c, staged ? 1 : 3);
    } else {
      if (c->fuse_rhs) { k_update_tet_v4<true, 0><<<grid4, kV4Threads, smem, st>>>(m, s, a, x_dev, h2, 0); k_update_tet_v4_tail<true, 0><<<tail_grid, kV4Threads, smem, st>>>(m, s, a, x_dev, h2); }
      else { k_update_tet_v4<false, 0><<<grid4, kV4Threads, smem, st>>>(m, s, a, x_dev, h2, 0); k_update_tet_v4_tail<false, 0><<<tail_grid, kV4Threads, smem, st>>>(m, s, a, x_dev, h2); }
      LAUNCHED(c, 1);
    }
  } else if (c->kind == MFB_SHELL) {
    if (c->fuse_rhs) k_update_shell<true><<<c->grid(), 128, 0, st>>>(m, s, x_dev, h2);
    else k_update_shell<false><<<c->grid(), 128, 0, st>>>(m, s, x_dev, h2);
  } else if (c->dim == 3) {
    // (the round-1 tet kernels - register B, shared-memory B, single fused launch: option smem_b 0 / 1 / 2 - were removed in
    // round 2; mfb_set_option rejects those values)
    g_last_error = "mfb_update_dev: no tet update kernel selected";
    return MFB_ERR_UNSUPPORTED;
  } else {
    if (c->fuse_rhs) k_update_tri2d<true, true><<<c->grid(), 128, 0, st>>>(m, s, x_dev, h2);
    else k_update_tri2d<false, true><<<c->grid(), 128, 0, st>>>(m, s, x_dev, h2);
  }
  if (c->kind == MFB_TET && c->nb > 0 && !bulk && !det) {   // RED variants accumulate the upper triangle only
    k_mirror_lower<<<nblk((int64_t)c->n_mirror * 9), 256, 0, st>>>(c->mirror.p, c->n_mirror, 3, c->Avals.p);
    LAUNCHED(c, 1);
  }
  if (c->world > 1 && c->update_exchange && !exchanged) {     // same semantics without the overlap (other element kinds, all-reduce path)
    int rcx = interface_sum(c, c->grad_x.p, c->dim, c->fuse_rhs ? c->rhs.p : nullptr, c->dim, nullptr, 0);
    if (rcx) return rcx;
    vec_recorded = false;
  }
  if (slot >= 0) MFB_CUDA(cudaEventRecord(c->ev1[slot], st));
  LAUNCHED(c, 1);
  MFB_CUDA(cudaGetLastError());
  c->rhs_valid = c->fuse_rhs != 0;
  c->ev_vec_valid = vec_recorded && c->rhs_valid;
  c->updated = true;
  return MFB_OK;
}

extern "C" int mfb_update(mfb_ctx* c, const double* x, double dt) {
  CHECK_CTX(c);
  if (!x) return MFB_ERR_INVALID;
  if (!c->copy_stream) {
    MFB_CUDA(cudaStreamCreateWithFlags(&c->copy_stream, cudaStreamNonBlocking));
    MFB_CUDA(cudaEventCreateWithFlags(&c->ev_in, cudaEventDisableTiming));
    MFB_CUDA(cudaEventCreateWithFlags(&c->ev_vec, cudaEventDisableTiming));
  }
  int rc;
  const bool chunked = c->kind == MFB_TET && c->smem_b == 3 && c->asm_mode == 1 && c->nE >= (1 << 18);
  if (chunked) {                     // mfb_update_dev uploads x piecewise, interleaved with the element chunks
    if (!c->ev_chunk[0]) for (int k = 0; k < mfb_ctx::kChunks; ++k) MFB_CUDA(cudaEventCreateWithFlags(&c->ev_chunk[k], cudaEventDisableTiming));
    c->x_host_chunked = x;
  } else {
    if ((rc = upload_x(c, x))) return rc;
    MFB_CUDA(cudaEventRecord(c->ev_in, c->stream));
  }
  rc = mfb_update_dev(c, c->xdev.p, dt);
  c->x_host_chunked = nullptr;
  if (rc) return rc;
  // The caller's buffer is free once the upload has finished.  The kernels are stream-ordered: every entry point that
  // hands results to the host synchronises on what it reads, so update() does not wait for the device here.
  MFB_CUDA(cudaEventSynchronize(c->ev_in));
  return MFB_OK;
}

static int ensure_rhs(mfb_ctx* c) {
  if (c->rhs_valid) return MFB_OK;
  MFB_CUDA(cudaMemsetAsync(c->rhs.p, 0, (size_t)c->dim * c->nF * 8, c->stream));
  if (c->kind == MFB_SHELL) k_rhs<K_SHELL><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state());
  else if (c->dim == 3) k_rhs<3><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state());
  else k_rhs<2><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state());
  LAUNCHED(c, 1);
  MFB_CUDA(cudaGetLastError());
  c->rhs_valid = true;
  c->ev_vec_valid = false;
  return MFB_OK;
}

extern "C" int mfb_rhs_dev(mfb_ctx* c, const double** out_dev) {
  CHECK_CTX(c);
  int rc = ensure_rhs(c);
  if (rc) return rc;
  if (out_dev) *out_dev = c->rhs.p;
  return MFB_OK;
}
extern "C" int mfb_rhs(mfb_ctx* c, double* out) {
  CHECK_CTX(c);
  if (!out) return MFB_ERR_INVALID;
  int rc = ensure_rhs(c);
  if (rc) return rc;
  if (c->ev_vec_valid && c->copy_stream) {       // rhs_ came out of the fused update: do not wait for the CSR values
    MFB_CUDA(cudaStreamWaitEvent(c->copy_stream, c->ev_vec, 0));
    MFB_CUDA(cudaMemcpyAsync(out, c->rhs.p, (size_t)c->dim * c->nF * 8, cudaMemcpyDeviceToHost, c->copy_stream));
    MFB_CUDA(cudaStreamSynchronize(c->copy_stream));
    return MFB_OK;
  }
  MFB_CUDA(cudaMemcpyAsync(out, c->rhs.p, (size_t)c->dim * c->nF * 8, cudaMemcpyDeviceToHost, c->stream));
  MFB_CUDA(cudaStreamSynchronize(c->stream));
  return MFB_OK;
}

extern "C" int mfb_solve_dev(mfb_ctx* c, const double* dx_dev) {
  CHECK_CTX(c);
  if (!dx_dev) return MFB_ERR_INVALID;
  int rc = ensure_rhs(c);   // solve() reads gl_, which rhs() produces (mixed_stretch.cpp:281)
  if (rc) return rc;
  if (c->kind == MFB_SHELL) k_solve<K_SHELL><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state(), dx_dev);
  else if (c->dim == 3) k_solve<3><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state(), dx_dev);
  else k_solve<2><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state(), dx_dev);
  LAUNCHED(c, 1);
  MFB_CUDA(cudaGetLastError());
  return MFB_OK;
}
extern "C" int mfb_solve(mfb_ctx* c, const double* dx) {
  CHECK_CTX(c);
  if (!dx) return MFB_ERR_INVALID;
  MFB_CUDA(cudaMemcpyAsync(c->vecF.p, dx, (size_t)c->dim * c->nF * 8, cudaMemcpyHostToDevice, c->stream));
  int rc = mfb_solve_dev(c, c->vecF.p);
  if (rc) return rc;
  MFB_CUDA(cudaStreamSynchronize(c->stream));
  return MFB_OK;
}

extern "C" int mfb_post_solve(mfb_ctx* c) {
  CHECK_CTX(c);
  MFB_CUDA(cudaMemsetAsync(c->la.p, 0, (size_t)c->N * c->ld * 8, c->stream));
  return MFB_OK;
}

static int finish_sum(mfb_ctx* c, double* e) {
  k_sum_partials<<<1, 256, 0, c->stream>>>(c->partials.p, c->grid(), c->scal.p);
  LAUNCHED(c, 1);
  MFB_CUDA(cudaMemcpyAsync(c->h_scal, c->scal.p, 8, cudaMemcpyDeviceToHost, c->stream));
  MFB_CUDA(cudaStreamSynchronize(c->stream));
  *e = c->h_scal[0];
  return MFB_OK;
}
static int aos_to_soa(mfb_ctx* c, const double* host_aos, mfb::DBuf<double>& soa, int ncomp) {
  int rc = c->aos.n < (size_t)ncomp * c->nE ? c->aos.alloc((size_t)c->M * c->N * c->nE) : MFB_OK;
  if (rc) return rc;
  MFB_CUDA(cudaMemcpyAsync(c->aos.p, host_aos, (size_t)ncomp * c->nE * 8, cudaMemcpyHostToDevice, c->stream));
  k_aos_to_soa<<<nblk((int64_t)ncomp * c->nE), 256, 0, c->stream>>>(c->aos.p, soa.p, ncomp, c->nE, c->ld);
  LAUNCHED(c, 1);
  MFB_CUDA(cudaGetLastError());
  return MFB_OK;
}

extern "C" int mfb_energy_dev(mfb_ctx* c, const double* s_soa, double* e) {
  CHECK_CTX(c);
  struct Span { mfb_ctx* c; Span(mfb_ctx* c_) : c(c_) { phase_begin(c, MFB_PH_ENERGY); } ~Span() { phase_end(c, MFB_PH_ENERGY); } } span(c);
  if (!e) return MFB_ERR_INVALID;
  const double* sp = s_soa ? s_soa : c->s.p;
  if (c->kind == MFB_SHELL) k_energy<K_SHELL><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), sp, c->partials.p);
  else if (c->dim == 3) k_energy<3><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), sp, c->partials.p);
  else k_energy<2><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), sp, c->partials.p);
  LAUNCHED(c, 1);
  MFB_CUDA(cudaGetLastError());
  return finish_sum(c, e);
}
extern "C" int mfb_energy(mfb_ctx* c, const double* /*x*/, const double* s, double* e) {
  CHECK_CTX(c);
  if (!s || !e) return MFB_ERR_INVALID;
  int rc = aos_to_soa(c, s, c->soa_tmp, c->N);
  if (rc) return rc;
  return mfb_energy_dev(c, c->soa_tmp.p, e);
}
extern "C" int mfb_constraint_value_dev(mfb_ctx* c, const double* x_dev, const double* s_soa, double* e) {
  CHECK_CTX(c);
  struct Span { mfb_ctx* c; Span(mfb_ctx* c_) : c(c_) { phase_begin(c, MFB_PH_CONSTRAINT); } ~Span() { phase_end(c, MFB_PH_CONSTRAINT); } } span(c);
  if (!x_dev || !e) return MFB_ERR_INVALID;
  const double* sp = s_soa ? s_soa : c->s.p;
  if (c->kind == MFB_SHELL) k_constraint<K_SHELL, false><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state(), x_dev, sp, c->partials.p, nullptr);
  else if (c->dim == 3) k_constraint<3, false><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state(), x_dev, sp, c->partials.p, nullptr);
  else k_constraint<2, false><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state(), x_dev, sp, c->partials.p, nullptr);
  LAUNCHED(c, 1);
  MFB_CUDA(cudaGetLastError());
  return finish_sum(c, e);
}
extern "C" int mfb_constraint_value(mfb_ctx* c, const double* x, const double* s, double* e) {
  CHECK_CTX(c);
  if (!x || !s || !e) return MFB_ERR_INVALID;
  int rc = upload_x(c, x);
  if (rc) return rc;
  rc = aos_to_soa(c, s, c->soa_tmp, c->N);
  if (rc) return rc;
  return mfb_constraint_value_dev(c, c->xdev.p, c->soa_tmp.p, e);
}
extern "C" int mfb_evaluate_constraint(mfb_ctx* c, const double* x, double* out) {
  CHECK_CTX(c);
  if (!x || !out) return MFB_ERR_INVALID;
  int rc = upload_x(c, x);
  if (rc) return rc;
  if (c->kind == MFB_SHELL) k_constraint<K_SHELL, true><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state(), c->xdev.p, c->s.p, nullptr, c->soa_tmp.p);
  else if (c->dim == 3) k_constraint<3, true><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state(), c->xdev.p, c->s.p, nullptr, c->soa_tmp.p);
  else k_constraint<2, true><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state(), c->xdev.p, c->s.p, nullptr, c->soa_tmp.p);
  LAUNCHED(c, 1);
  MFB_CUDA(cudaGetLastError());
  if (c->aos.n < (size_t)c->N * c->nE) { rc = c->aos.alloc((size_t)c->M * c->N * c->nE); if (rc) return rc; }
  k_soa_to_aos<<<nblk((int64_t)c->N * c->nE), 256, 0, c->stream>>>(c->soa_tmp.p, c->aos.p, c->N, c->nE, c->ld);
  LAUNCHED(c, 1);
  MFB_CUDA(cudaMemcpyAsync(out, c->aos.p, (size_t)c->N * c->nE * 8, cudaMemcpyDeviceToHost, c->stream));
  MFB_CUDA(cudaStreamSynchronize(c->stream));
  return MFB_OK;
}

extern "C" int mfb_lhs_values_dev(mfb_ctx* c, const double** v) { CHECK_CTX(c); if (v) *v = c->Avals.p; return MFB_OK; }
extern "C" int mfb_lhs_values(mfb_ctx* c, double* values) {
  CHECK_CTX(c);
  if (!values) return MFB_ERR_INVALID;
  if (c->nnz) MFB_CUDA(cudaMemcpyAsync(values, c->Avals.p, (size_t)c->nnz * 8, cudaMemcpyDeviceToHost, c->stream));
  MFB_CUDA(cudaStreamSynchronize(c->stream));
  return MFB_OK;
}

extern "C" int mfb_product(mfb_ctx* c, int which, const double* x, double* out, int transposed) {
  CHECK_CTX(c);
  if (!x || !out || which < 0 || which > 3) return MFB_ERR_INVALID;
  const size_t nEl = (size_t)c->N * c->nE, nVx = (size_t)c->dim * c->nF;
  const bool x_is_elem = (which != 3) || transposed;
  const bool out_is_elem = (which != 3) || !transposed;
  int rc;
  const double* xin;
  double* o;
  if (x_is_elem) { rc = aos_to_soa(c, x, c->soa_tmp, c->N); if (rc) return rc; xin = c->soa_tmp.p; }
  else { MFB_CUDA(cudaMemcpyAsync(c->vecF.p, x, nVx * 8, cudaMemcpyHostToDevice, c->stream)); xin = c->vecF.p; }
  if (out_is_elem) {
    // second staging buffer for out
    mfb::DBuf<double> stage;
    rc = stage.alloc(nEl); if (rc) return rc;
    MFB_CUDA(cudaMemcpyAsync(stage.p, out, nEl * 8, cudaMemcpyHostToDevice, c->stream));
    k_aos_to_soa<<<nblk((int64_t)nEl), 256, 0, c->stream>>>(stage.p, c->soa_tmp2.p, c->N, c->nE, c->ld);
    o = c->soa_tmp2.p;
    if (c->kind == MFB_SHELL) k_product<K_SHELL><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state(), which, transposed, xin, o);
    else if (c->dim == 3) k_product<3><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state(), which, transposed, xin, o);
    else k_product<2><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state(), which, transposed, xin, o);
    k_soa_to_aos<<<nblk((int64_t)nEl), 256, 0, c->stream>>>(o, stage.p, c->N, c->nE, c->ld);
    LAUNCHED(c, 3);
    MFB_CUDA(cudaMemcpyAsync(out, stage.p, nEl * 8, cudaMemcpyDeviceToHost, c->stream));
    MFB_CUDA(cudaStreamSynchronize(c->stream));
  } else {
    mfb::DBuf<double> stage;
    rc = stage.alloc(nVx + 1); if (rc) return rc;
    MFB_CUDA(cudaMemcpyAsync(stage.p, out, nVx * 8, cudaMemcpyHostToDevice, c->stream));
    if (c->kind == MFB_SHELL) k_product<K_SHELL><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state(), which, transposed, xin, stage.p);
    else if (c->dim == 3) k_product<3><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state(), which, transposed, xin, stage.p);
    else k_product<2><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state(), which, transposed, xin, stage.p);
    LAUNCHED(c, 1);
    MFB_CUDA(cudaMemcpyAsync(out, stage.p, nVx * 8, cudaMemcpyDeviceToHost, c->stream));
    MFB_CUDA(cudaStreamSynchronize(c->stream));
  }
  MFB_CUDA(cudaGetLastError());
  return MFB_OK;
}

extern "C" int64_t mfb_jacobian_x_nnz(mfb_ctx* c) {
  if (!c) return 0;
  return (int64_t)c->nE * c->N * c->npe * c->dim;
}

extern "C" int mfb_jacobian_x_coo(mfb_ctx* c, int32_t* rows, int32_t* cols, double* vals) {
  CHECK_CTX(c);
  if (!rows || !cols || !vals) return MFB_ERR_INVALID;
  const int64_t n = mfb_jacobian_x_nnz(c);
  mfb::DBuf<int> dr, dc;
  mfb::DBuf<double> dv;
  int rc;
  if ((rc = dr.alloc((size_t)n + 1)) || (rc = dc.alloc((size_t)n + 1)) || (rc = dv.alloc((size_t)n + 1))) return rc;
  if (c->kind == MFB_SHELL) k_jacobian_x_coo<K_SHELL><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state(), dr.p, dc.p, dv.p);
  else if (c->dim == 3) k_jacobian_x_coo<3><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state(), dr.p, dc.p, dv.p);
  else k_jacobian_x_coo<2><<<c->grid(), 128, 0, c->stream>>>(c->mesh(), c->state(), dr.p, dc.p, dv.p);
  LAUNCHED(c, 1);
  MFB_CUDA(cudaMemcpyAsync(rows, dr.p, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream));
  MFB_CUDA(cudaMemcpyAsync(cols, dc.p, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream));
  MFB_CUDA(cudaMemcpyAsync(vals, dv.p, (size_t)n * 8, cudaMemcpyDeviceToHost, c->stream));
  MFB_CUDA(cudaStreamSynchronize(c->stream));
  MFB_CUDA(cudaGetLastError());
  return MFB_OK;
}

extern "C" int mfb_last_status(mfb_ctx* c, mfb_status* st) {
  CHECK_CTX(c);
  if (!st) return MFB_ERR_INVALID;
  unsigned long long h[6];
  MFB_CUDA(cudaMemcpyAsync(h, c->counters.p, sizeof(h), cudaMemcpyDeviceToHost, c->stream));
  MFB_CUDA(cudaStreamSynchronize(c->stream));
  st->polar_maxiter = (int64_t)h[0]; st->polar_hfail = (int64_t)h[1]; st->nonfinite = (int64_t)h[2];
  st->polar_iters = (int64_t)h[3]; st->eig_sweeps = (int64_t)h[4]; st->psd_slow = (int64_t)h[5];
  return MFB_OK;
}

// sum / count of the update-kernel durations recorded since the last call (option "profile" = 1); resets the ring
#if __has_include(<nvtx3/nvToolsExt.h>)
#include <nvtx3/nvToolsExt.h>
#define MFB_NVTX 1
#endif
static const char* const kPhaseNames[MFB_PHASE_COUNT] = {"update", "rhs", "linsolve", "solve", "LS", "energy", "constraint"};
extern "C" const char* mfb_phase_name(int phase) { return (phase >= 0 && phase < MFB_PHASE_COUNT) ? kPhaseNames[phase] : ""; }

void mfb::phase_begin(mfb_ctx* c, int key) {
  if (!c->profile || key < 0 || key >= MFB_PHASE_COUNT || c->ph_open[key]) return;
#ifdef MFB_NVTX
  nvtxRangePushA(kPhaseNames[key]);
#endif
  cudaEvent_t e;
  if (!c->ph_free.empty()) { e = c->ph_free.back(); c->ph_free.pop_back(); }
  else if (cudaEventCreate(&e) != cudaSuccess) return;
  cudaEventRecord(e, c->stream);
  c->ph_open[key] = e;
}
void mfb::phase_end(mfb_ctx* c, int key) {
  if (key < 0 || key >= MFB_PHASE_COUNT || !c->ph_open[key]) return;
#ifdef MFB_NVTX
  nvtxRangePop();
#endif
  cudaEvent_t e;
  if (!c->ph_free.empty()) { e = c->ph_free.back(); c->ph_free.pop_back(); }
  else if (cudaEventCreate(&e) != cudaSuccess) { c->ph_free.push_back(c->ph_open[key]); c->ph_open[key] = 0; return; }
  cudaEventRecord(e, c->stream);
  c->ph_spans.push_back({key, c->ph_open[key], e});
  c->ph_open[key] = 0;
}
extern "C" int mfb_phase_times(mfb_ctx* c, double* ms, int32_t* counts, int reset) {
  CHECK_CTX(c);
  static thread_local double acc[MFB_PHASE_COUNT];
  static thread_local int32_t cnt[MFB_PHASE_COUNT];
  MFB_CUDA(cudaStreamSynchronize(c->stream));
  for (int k = 0; k < MFB_PHASE_COUNT; ++k) { acc[k] = 0; cnt[k] = 0; }
  for (auto& sp : c->ph_spans) {
    float t = 0.f;
    if (cudaEventElapsedTime(&t, sp.a, sp.b) == cudaSuccess) { acc[sp.key] += t; cnt[sp.key] += 1; }
  }
  for (int k = 0; k < MFB_PHASE_COUNT; ++k) { if (ms) ms[k] = acc[k]; if (counts) counts[k] = cnt[k]; }
  if (reset) {
    for (auto& sp : c->ph_spans) { c->ph_free.push_back(sp.a); c->ph_free.push_back(sp.b); }
    c->ph_spans.clear();
  }
  return MFB_OK;
}

// resident CTAs per SM of the tet update kernels as the runtime computes them (which: 0 plain, 1 pipelined) - tools only
extern "C" int mfb_debug_occupancy(int which) {
  int n = -1;
  const int smem = kV4SmemB + kV4SmemStage;
  cudaFuncSetAttribute(k_update_tet_v4<true, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
  cudaFuncSetAttribute(k_update_tet_v4_pipe<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem + kV4SmemPipe);
  cudaFuncSetAttribute(k_update_tet_v4_pipe<true>, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
  if (which == 0) cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_update_tet_v4<true, 1>, kV4Threads, smem);
  else cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, k_update_tet_v4_pipe<true>, kV4Threads, smem + kV4SmemPipe);
  return n;
}

extern "C" int mfb_kernel_time_ms(mfb_ctx* c, double* total_ms, int* count) {
  CHECK_CTX(c);
  MFB_CUDA(cudaStreamSynchronize(c->stream));
  double tot = 0.0;
  for (int i = 0; i < c->ev_count; ++i) { float ms = 0.f; MFB_CUDA(cudaEventElapsedTime(&ms, c->ev0[i], c->ev1[i])); tot += ms; }
  if (total_ms) *total_ms = tot;
  if (count) *count = c->ev_count;
  c->ev_count = 0;
  return MFB_OK;
}
