// Context object behind the C ABI: owns every device allocation of one MixedStretch instance.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <string>
#include <vector>
#include "../../include/mfem_b200.h"
#include "mfb_kernels2.cuh"
#include "mfb_kernels4.cuh"
#include "mfb_kernels5.cuh"

namespace mfb {

extern thread_local std::string g_last_error;

#define MFB_CUDA(call)                                                                                   \
  do {                                                                                                   \
    cudaError_t err__ = (call);                                                                          \
    if (err__ != cudaSuccess) {                                                                          \
      mfb::g_last_error = std::string(#call) + ": " + cudaGetErrorString(err__);                         \
      return MFB_ERR_CUDA;                                                                               \
    }                                                                                                    \
  } while (0)

template <typename T>
struct DBuf {   // plain device buffer
  T* p = nullptr;
  size_t n = 0;
  int alloc(size_t count) {
    free();
    n = count;
    if (count == 0) return MFB_OK;
    cudaError_t e = cudaMalloc((void**)&p, count * sizeof(T));
    if (e != cudaSuccess) { g_last_error = std::string("cudaMalloc: ") + cudaGetErrorString(e); p = nullptr; return MFB_ERR_CUDA; }
    return MFB_OK;
  }
  void free() { if (p) cudaFree(p); p = nullptr; n = 0; }
  ~DBuf() { free(); }
};

}  // namespace mfb

struct mfb_ctx;
struct mfb_partition;
namespace mfb {
// ---- peer mailboxes (mfb_dist.cu: peer_setup; used inside the CG kernels of mfb_pcg.cu) -----------------------------
// One cudaMalloc'ed mailbox per rank, opened by every other rank of the node through CUDA IPC (NVLink / NVSwitch peer
// memory).  A rank STORES its interface rows of t = L p straight into its neighbours' mailboxes from inside the SpMV
// kernel, and the last block of a kernel stores the rank's partial dot products into EVERY rank's mailbox followed by a
// sequence number; the consuming kernel of the same CG iteration polls its OWN mailbox (local L2) until all sequence
// numbers have arrived and adds the partial sums in rank order - the same bits on every rank.  No NCCL call and no host
// round trip inside the iteration.  Layout in 8-byte words:
//   [0, 16)   flag[stage][src rank]         sequence number of the last exchange src has posted for this stage
//   [16, 80)  sc[stage][parity][src rank][2] partial sums (stage 0: p.Lp ; stage 1: r.r, r.z)
//   [80, 96)  bc[stage][parity][4]           local: the sums block 0 has collected (sum0, sum1, ok) for this GPU's other blocks
//   [96, 98)  go[stage]                      local: sequence number block 0 has published
//   [128, ..) halo[parity][n_ent * 3]       this rank's receive segments, one per neighbour, in its own entry order
// Spins are bounded (kPeerSpinLimit polls of >= 128 ns): on a timeout the kernel raises scal[13] and the stop flag.
constexpr int kPeerMax = 8;
constexpr int kPeerFlagWords = 2 * kPeerMax, kPeerScWords = 2 * 2 * kPeerMax * 2, kPeerHaloWord0 = 128;
constexpr long long kPeerSpinLimit = 1LL << 24;
struct PeerView {
  int world, rank, n_ent;
  unsigned long long* flag_at[2][kPeerMax];     // [stage][r]: &flag[stage][me] in rank r's mailbox (r == me: the local one)
  double* sc_at[2][kPeerMax];                   // [stage][r]: &sc[stage][0][me][0] in rank r's mailbox
  unsigned long long* my_flag;                  // local mailbox
  double* my_sc;
  double* my_halo;                              // halo[0]; parity 1 follows at 3 n_ent doubles
  double* const* dst;                           // [parity][n_ent]: where entry s of this rank's send list lands (peer memory)
};
// mfb_dist.cu (no-ops when world == 1)
int interface_sum(mfb_ctx* c, double* a, int na, double* b, int nb, double* scal, int ns);
int allreduce_scalars(mfb_ctx* c, double* v, int n, bool maxop);
void comm_free(mfb_ctx* c);
// mfb_partition.cu -> mfb_dist.cu
struct PartitionView {
  int nparts, rank, nF, nE, n_if_elems, n_shared;
  const std::vector<int32_t>*if_local, *if_pos, *nbr_rank, *nbr_count, *nbr_local;
  const std::vector<double>* dot_weight;
};
// mfb_newton.cu: what the saddle-point systems (mfb_saddle.cu) read from the Displacement / BDF1 state
struct NewtonView {
  double h, f[3];
  const double* emass;    // per element: density x vol
  double *Mvals, *x, *xt, *diff, *dx;
  int* vfree;
};
// CUDA-event spans under the reference's Timer keys (optimizer_data.cpp:127-256; option "profile" = 1), also NVTX ranges
void phase_begin(mfb_ctx* c, int key);                            // mfb_api.cu
void phase_end(mfb_ctx* c, int key);
int partition_view(const mfb_partition* p, PartitionView* v);      // mfb_partition.cu
int build_det_lists(mfb_ctx* c);                                  // mfb_create.cu (asm_mode 2)
int peer_setup(mfb_ctx* c);                                       // mfb_dist.cu: collective; sets c->peer_state
// point-to-point halo of the CG (mfb_dist.cu): group{ send/recv of the packed interface rows ; all-reduce of `ns` scalars }
int halo_exchange(mfb_ctx* c, int ncomp, double* scal, int ns);
int halo_buffers(mfb_ctx* c, int ncomp, double** send, double** recv);
int p2p_interface_sum_on(mfb_ctx* c, double* a, int na, double* b, int nb, double* scal, int ns, cudaStream_t st);
void saddle_free(mfb_ctx* c);                                     // mfb_saddle.cu
void stretch_free(mfb_ctx* c);                                    // mfb_stretch.cu
int newton_view(mfb_ctx* c, NewtonView* v);                       // MFB_ERR_INVALID before mfb_newton_init
int newton_disp_gradient(mfb_ctx* c, double* out_free_dev);       // Displacement::gradient() = PM (x - x_tilde - h^2 f), d*nF
}  // namespace mfb

struct mfb_ctx {
  int device = 0;
  cudaStream_t stream = 0;
  int kind = 0;               // mfb_elem
  int dim = 3, npe = 4, N = 6, M = 9, NP = 21;
  int nV = 0, nE = 0, nF = 0, nb = 0;
  int64_t ld = 0, nnz = 0;
  // options
  int fuse_rhs = 1;
  int assembly = 0;
  int smem_b = 3;       // tet update kernel: 3 = k_update_tet_v4, the only one (the round-1 variants 0 / 1 / 2 were removed in round 2)
  int exp = 0;
  int asm_mode = 1;     // v4 assembly: 1 = warp-cooperative coalesced RED into the block-CSR upper triangle, 0 = FP64 RED into the scalar CSR,
                        // 2 = deterministic owner-computes gather (no atomics; bit-identical from run to run)
  // mesh (device)
  mfb::DBuf<int> T, fidx, matid, smap, sstride;
  mfb::DBuf<double> dphi, vol, vref, nrest, xupd;   // xupd: shells, positions of the last update()
  mfb::DBuf<mfb::MaterialDev> mats;
  mfb::MaterialDev mat0_host{0, 0.0, 0.0};
  // CSR structure (device)
  mfb::DBuf<int> blockptr, brow, bcol, bT, diagblk, outer, inner;
  mfb::DBuf<int4> mirror;     // (dst0, dst stride, src0, src stride) per lower-triangle block, see k_mirror_lower
  int n_mirror = 0;
  // v4 assembly: block-CSR upper triangle (9 doubles per block), element->block map (6 pairs j<k), per-block source code, interleaved vectors
  mfb::DBuf<double> Ablk, gxrh;
  mfb::DBuf<int> bmap, slow_list;
  mfb::DBuf<int> fcode;       // per block: 2 * upper block + transposed
  // deterministic assembly (asm_mode 2, built on first use): per CSR block the (element, local pair) contributions in ascending
  // element order - the reference Assembler's own structure (sparse_utils.cpp:58-99) - per free vertex its (element, local
  // vertex) incidences, and the per-element scratch the gather kernels read (6 off-diagonal 3x3 blocks, 4 x 6 vertex values)
  mfb::DBuf<int> det_bptr, det_bent, det_vptr, det_vent;
  mfb::DBuf<double> Kelem, Velem;
  bool det_ready = false;
  int nbu = 0;
  // state (device)
  mfb::DBuf<double> R, s, la, ds, S, dSdF, g, H, Hinv, evals, gl, grad_la, grad_x, rhs, Avals;
  mfb::DBuf<int> iters;
  mfb::DBuf<unsigned long long> counters;
  // scratch
  mfb::DBuf<double> xdev, vecF, partials, scal, aos, soa_tmp, soa_tmp2;
  // pcg work vectors
  mfb::DBuf<double> pcg_r, pcg_p, pcg_z, pcg_t, pcg_dinv, pcg_part, pcg_lhs;
  mfb::DBuf<int> pcg_abort;
  mfb::DBuf<long long> pcg_prof;
  int cg_persist = 2;        // option: the CG loop as one persistent cooperative kernel (mfb_pcg.cu: k_pcg_persist); 2 = for small systems
  mfb::DBuf<double> pcg_res, pcg_d;          // Chebyshev preconditioner (precond 3): inner residual and correction
  int cheb_degree = 6, cheb_ratio = 120;       // options "cheb_degree", "cheb_ratio" (lambda_max / lambda_min of the target interval)
  double* h_scal = nullptr;   // pinned
  // host copies (lazy)
  std::vector<int32_t> h_outer, h_inner, h_smap, h_sstride;
  std::vector<int32_t> h_matid;          // kept for mfb_newton_init's per-material densities
  std::vector<double> h_mat_density;
  std::vector<int32_t> h_mat_model;      // model of every material (the Stretch variable checks that each has an F-form)
  // flags
  bool rhs_valid = false;     // rhs_/gl_ match the current S_, s_, g_, H_, dSdF_
  bool updated = false;
  int64_t launches = 0;
  // optional CUDA-event timing of the update kernel alone (bench.py roofline): ring of event pairs on the stream
  int profile = 0;
  int cg_zero_guess = 0;   // mfb_newton_step: start CG from 0 instead of the reference's guess (parity runs, see mfb_newton.cu)
  std::vector<cudaEvent_t> ev0, ev1;
  int ev_count = 0;
  struct PhaseSpan { int key; cudaEvent_t a, b; };
  std::vector<PhaseSpan> ph_spans;         // closed spans since the last mfb_phase_times
  std::vector<cudaEvent_t> ph_free;        // event pool
  cudaEvent_t ph_open[MFB_PHASE_COUNT] = {0};
  void* newton = nullptr;     // mfb_newton_state (mfb_newton.cu)
  void* saddle = nullptr;     // scratch of the saddle-point systems (mfb_saddle.cu)
  void* stretch = nullptr;    // the non-mixed Stretch variable's lhs values / gradient (mfb_stretch.cu)
  // host-buffer entry points (mfb_update / mfb_rhs): the device->host copy of rhs_ runs on its own stream as soon as
  // the vertex vectors are final, beside the kernel that writes the CSR values
  cudaStream_t copy_stream = 0;
  cudaEvent_t ev_in = 0, ev_vec = 0;
  bool ev_vec_valid = false;
  // chunked upload of x in mfb_update: element chunk k needs the vertices below chunk_vhi[k] (prefix maximum over the
  // chunk's elements, from mfb_create), so its kernel starts as soon as that part of x has arrived
  static constexpr int kChunks = 32;      // at most
  int nchunks = 8;
  int chunk_e0[kChunks + 1] = {0};
  int chunk_vhi[kChunks] = {0};
  cudaEvent_t ev_chunk[kChunks] = {0};
  // staged finalize (option stage_finalize; OFF by default - measured on B200 at 10M tets: 5.15 .. 5.5 ms per update against
  // 5.07 ms for the single pass, for 2 .. 16 chunks, either stream priority, finalize CTAs of 64 .. 256 threads, with and
  // without the concurrent memset: the element kernel is latency-bound and loses more to the extra memory traffic beside it
  // than the 0.8 ms of finalize + memset it hides; profiles/README.md): block-row gi of the CSR is final once
  // the last element that holds vertex gi has run, so k_finalize_rows goes out per element chunk, on a side stream, over
  // the rows whose last element lies in that chunk (fin_rows sorted by chunk, fin_ptr[k] .. fin_ptr[k+1]), while the next
  // chunks run; each chunk keeps its own list of slow elements (slow_list + chunk_e0[k], counter chunk_cnt[2k]) so that its
  // Jacobi tail can run right behind it.  The upper-triangle accumulator is double-buffered: while update n reduces into
  // Ablk_buf[n % 2], a memset on a third stream clears the other one for update n + 1.
  int stage_finalize = 0;
  int zero_overlap = 0;      // 1: with >= 2^18 tets the accumulator is double-buffered and the half for the NEXT update is cleared on a side stream
                             // while this one runs.  Measured at 10M tets: the step stays at 5.04 ms - the 0.155 ms memset leaves the
                             // critical path and the element kernel gives 0.13 ms back to the traffic beside it (4.89 -> 5.02 ms) - off
  mfb::DBuf<int> fin_rows;
  int fin_ptr[kChunks + 1] = {0};
  mfb::DBuf<unsigned long long> chunk_cnt;       // [2 kChunks]
  mfb::DBuf<double> Ablk2;
  int ablk_cur = 0;                              // the buffer asmdesc() hands out
  bool ablk_clean[2] = {false, false};           // zero (or being zeroed, ev_zero pending) since its last finalize
  cudaStream_t fin_stream = nullptr, zero_stream = nullptr;
  cudaEvent_t ev_stage[kChunks] = {0}, ev_fin = nullptr, ev_zero[2] = {nullptr, nullptr}, ev_upd = nullptr;
  double* ablk(int i) const { return i ? Ablk2.p : Ablk.p; }
  const double* x_host_chunked = nullptr;   // set by mfb_update around its call of mfb_update_dev
  // multi-GPU (mfb_dist.cu): NCCL communicator, this rank's interface vertices (local free id, slot in the packed
  // buffer over all shared vertices), dot-product weights (1 interior / owned interface vertex, 0 copy owned elsewhere)
  void* comm = nullptr;
  int rank = 0, world = 1;
  mfb::DBuf<int> if_local, if_pos, if_inv;   // if_inv[slot] = local free id or -1
  int n_if_local = 0, n_shared = 0;
  mfb::DBuf<double> if_buf, dot_w;
  // point-to-point exchange (mfb_comm_set_partition): neighbour ranks, entry offsets, local free id per entry, and per
  // local free row the list of its entries (rowif_ptr / rowif_idx: the CG kernels write / read their halo through it)
  int use_p2p = 1, update_exchange = 0;
  int convergence_log = 0;                    // mfb_newton_step records the non-mixed gradient norm per iteration (apps/convergence.cpp)
  int pipe = 0;                               // update kernel: persistent software-pipelined variant (option "pipe")
  int cg_graph = 0;                           // mfb_pcg_dev: replay batches of 32 iterations as a CUDA graph (option "cg_graph"; measured: no gain, the
                                              // iteration is bound by its kernels on 1 GPU and by the latency of its 3 NCCL kernels on several)
  int n_if_elems = 0, n_nbr = 0, n_ent = 0;
  std::vector<int> nbr_rank, nbr_off;        // nbr_off[n_nbr + 1]
  mfb::DBuf<int> ent_local, rowif_ptr, rowif_idx;
  mfb::DBuf<double> xsend, xrecv;
  int xbuf_comp = 0;
  // peer mailboxes (option "peer", default 1; state: 0 not tried, 1 in use, -1 unavailable on this node -> NCCL path)
  int use_peer = 1, peer_state = 0;
  void* peer_box = nullptr;                      // this rank's mailbox (cudaMalloc, exported)
  void* peer_base[mfb::kPeerMax] = {nullptr};    // the others', opened with cudaIpcOpenMemHandle
  mfb::DBuf<double*> peer_dst;
  mfb::DBuf<mfb::PeerView> peer_view;
  unsigned long long peer_seq = 0;               // sequence numbers handed out so far (same on every rank)
  cudaStream_t xstream = 0;                   // side stream of the overlapped exchange in mfb_update_dev
  cudaEvent_t ev_x0 = 0, ev_x1 = 0;

  mfb::DevMesh mesh() const {
    mfb::DevMesh m;
    m.dim = dim; m.npe = npe; m.nV = nV; m.nE = nE; m.nF = nF; m.ld = ld;
    m.T = T.p; m.fidx = fidx.p; m.dphi = dphi.p; m.vol = vol.p; m.matid = matid.p; m.mats = mats.p; m.mat0 = mat0_host; m.exp = exp;
    m.smap = smap.p; m.sstride = sstride.p;
    m.nrest = nrest.p; m.xlast = xupd.p;
    return m;
  }
  mfb::DevState state() const {
    mfb::DevState st;
    st.R = R.p; st.s = s.p; st.la = la.p; st.ds = ds.p; st.S = S.p; st.dSdF = dSdF.p; st.g = g.p; st.H = H.p;
    st.Hinv = Hinv.p; st.evals = evals.p; st.gl = gl.p; st.grad_la = grad_la.p; st.iters = iters.p;
    st.grad_x = grad_x.p; st.rhs = rhs.p; st.Avals = Avals.p; st.counters = counters.p;
    return st;
  }
  mfb::DevAsm asmdesc() const {
    mfb::DevAsm a;
    a.Ablk = ablk(ablk_cur); a.gxrh = gxrh.p; a.bmap = bmap.p; a.slow_list = slow_list.p; a.slow_cnt = counters.p + 5;
    a.sinkA = nbu * 9; a.sinkV = 6 * nF;
    a.Kelem = Kelem.p; a.Velem = Velem.p;
    return a;
  }
  int grid() const { return (nE + 127) / 128; }
};
