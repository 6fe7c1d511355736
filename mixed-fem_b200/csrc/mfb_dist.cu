// Multi-GPU exchange step of the path (the reference is a single-address-space OpenMP code, so this
// layer has no reference counterpart).  One process per GPU, each with a LOCAL context built on its slab of elements;
// the assembled matrix stays sub-assembled (A = sum_r P_r^T A_r P_r), and the only coupling is at free vertices shared
// between slabs:
//   * per-vertex vectors (grad_x_, rhs_, matvec results, diagonal blocks): pack the interface entries into ONE buffer
//     laid out over ALL shared vertices (zero where this rank has no copy), ncclAllReduce(sum), unpack;
//   * scalars (dot products, energies, norms): ncclAllReduce on the context's scalar block.
// Everything is enqueued on the context stream: no host synchronisation is added by this file.
// NCCL is bound at run time (dlopen "libnccl.so.2": the copy the host process already holds - e.g. torch's - else the
// system one), so the library has no NCCL dependency for single-GPU users.
//
// Two exchange paths.  (1) all-reduce path (mfb_comm_set_interface): ONE buffer over all shared vertices of the partition.
// (2) point-to-point path (mfb_comm_set_partition, default when set): every rank packs only ITS cuts, one grouped
// ncclSend / ncclRecv per neighbour (<= 2 for slabs), arrivals are added to the local values - a + b on one side, b + a
// on the other, bit-identical copies for vertices shared by two parts; scalar sums ride in the same NCCL group.
// (3) peer mailboxes (peer_setup below; default for the CG when a partition is set and all ranks sit on one node): one
// cudaMalloc'ed buffer per rank, opened by the others through CUDA IPC; the CG kernels of mfb_pcg.cu store halo rows and
// partial dot products into each other's memory over NVLink and synchronise on sequence numbers - no NCCL call inside
// the iteration.  NCCL is still what sets it up (handle exchange) and what everything else uses.
#include "mfb_ctx.cuh"
#include <dlfcn.h>
#if __has_include(<nccl.h>)
#include <nccl.h>
#else      // the few NCCL declarations this file needs, so that the single-GPU library builds without NCCL headers
typedef struct ncclComm* ncclComm_t;
typedef struct { char internal[128]; } ncclUniqueId;
typedef enum { ncclSuccess = 0 } ncclResult_t;
typedef enum { ncclSum = 0, ncclProd = 1, ncclMax = 2, ncclMin = 3 } ncclRedOp_t;
typedef enum { ncclInt8 = 0, ncclUint8 = 1, ncclInt32 = 2, ncclUint32 = 3, ncclInt64 = 4, ncclUint64 = 5, ncclFloat16 = 6, ncclFloat32 = 7, ncclFloat64 = 8, ncclDouble = 8 } ncclDataType_t;
#endif
#include <cstring>

using namespace mfb;

#define CHECK_CTX(c) do { if (!(c)) { g_last_error = "null context"; return MFB_ERR_INVALID; } cudaSetDevice((c)->device); } while (0)

namespace {

struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  bool ok = false;
};

NcclApi* nccl() {
  static NcclApi api;
  static bool tried = false;
  if (!tried) {
    tried = true;
    void* h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_LOCAL);
    if (!h) h = dlopen("libnccl.so", RTLD_NOW | RTLD_LOCAL);
    if (h) {
      api.GetUniqueId = (decltype(api.GetUniqueId))dlsym(h, "ncclGetUniqueId");
      api.CommInitRank = (decltype(api.CommInitRank))dlsym(h, "ncclCommInitRank");
      api.AllReduce = (decltype(api.AllReduce))dlsym(h, "ncclAllReduce");
      api.CommDestroy = (decltype(api.CommDestroy))dlsym(h, "ncclCommDestroy");
      api.GetErrorString = (decltype(api.GetErrorString))dlsym(h, "ncclGetErrorString");
      api.Send = (decltype(api.Send))dlsym(h, "ncclSend");
      api.Recv = (decltype(api.Recv))dlsym(h, "ncclRecv");
      api.GroupStart = (decltype(api.GroupStart))dlsym(h, "ncclGroupStart");
      api.GroupEnd = (decltype(api.GroupEnd))dlsym(h, "ncclGroupEnd");
      api.ok = api.GetUniqueId && api.CommInitRank && api.AllReduce && api.CommDestroy && api.GetErrorString && api.Send && api.Recv &&
               api.GroupStart && api.GroupEnd;
    }
  }
  return &api;
}

int nccl_fail(const char* what, ncclResult_t r) {
  g_last_error = std::string(what) + ": " + (nccl()->GetErrorString ? nccl()->GetErrorString(r) : "NCCL error");
  return MFB_ERR_CUDA;
}

inline unsigned nblk(int64_t n, int t = 256) { return (unsigned)((n + t - 1) / t); }

// buf over ALL shared vertices: (a | b) where this rank has a copy (inv[pos] = local free id), zero elsewhere; the
// ns scalars ride at the end (one collective instead of two).  One launch packs, one unpacks.
__global__ void k_if_pack(const int* __restrict__ inv, int n_shared, const double* __restrict__ a, int na, const double* __restrict__ b, int nb,
                          const double* __restrict__ scal, int ns, double* __restrict__ buf) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int w = na + nb, nv = n_shared * w;
  if (i >= nv + ns) return;
  if (i >= nv) { buf[i] = scal[i - nv]; return; }
  const int v = i / w, k = i - v * w;
  const int64_t l = inv[v];
  buf[i] = l < 0 ? 0.0 : ((k < na) ? a[l * na + k] : b[l * nb + (k - na)]);
}
__global__ void k_if_unpack(const int* __restrict__ inv, int n_shared, double* __restrict__ a, int na, double* __restrict__ b, int nb,
                            double* __restrict__ scal, int ns, const double* __restrict__ buf) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int w = na + nb, nv = n_shared * w;
  if (i >= nv + ns) return;
  if (i >= nv) { scal[i - nv] = buf[i]; return; }
  const int v = i / w, k = i - v * w;
  const int64_t l = inv[v];
  if (l < 0) return;
  if (k < na) a[l * na + k] = buf[i]; else b[l * nb + (k - na)] = buf[i];
}
__global__ void k_if_invert(const int* __restrict__ if_local, const int* __restrict__ if_pos, int n, int* __restrict__ inv) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) inv[if_pos[i]] = if_local[i];
}

// point-to-point path: entry i (concatenated over the neighbours) <-> local free vertex ent[i]
__global__ void k_p2p_pack(const int* __restrict__ ent, int n_ent, const double* __restrict__ a, int na, const double* __restrict__ b, int nb,
                           double* __restrict__ send) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int w = na + nb;
  if (i >= n_ent * w) return;
  const int v = i / w, k = i - v * w;
  const int64_t l = ent[v];
  send[i] = (k < na) ? a[l * na + k] : b[l * nb + (k - na)];
}
// a vertex shared with more than one neighbour has several entries: atomic adds (order-dependent only in that case)
__global__ void k_p2p_unpack_add(const int* __restrict__ ent, int n_ent, double* __restrict__ a, int na, double* __restrict__ b, int nb,
                                 const double* __restrict__ recv) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int w = na + nb;
  if (i >= n_ent * w) return;
  const int v = i / w, k = i - v * w;
  const int64_t l = ent[v];
  if (k < na) atomicAdd(&a[l * na + k], recv[i]); else atomicAdd(&b[l * nb + (k - na)], recv[i]);
}

int ensure_xbuf(mfb_ctx* c, int ncomp) {
  if (c->xbuf_comp >= ncomp) return MFB_OK;
  int rc;
  const size_t n = (size_t)c->n_ent * ncomp + 1;
  if ((rc = c->xsend.alloc(n)) || (rc = c->xrecv.alloc(n))) return rc;
  c->xbuf_comp = ncomp;
  return MFB_OK;
}

}  // namespace

namespace mfb {

// group{ send / recv of xsend / xrecv (ncomp doubles per entry) with every neighbour ; all-reduce(sum) of ns scalars }
// on `st`.  The send buffer must have been filled on `st` (k_p2p_pack, or the CG's SpMV kernel).
static int halo_exchange_on(mfb_ctx* c, int ncomp, double* scal, int ns, cudaStream_t st) {
  NcclApi* api = nccl();
  ncclResult_t r = api->GroupStart();
  if (r != ncclSuccess) return nccl_fail("ncclGroupStart", r);
  for (int q = 0; q < c->n_nbr; ++q) {
    const size_t off = (size_t)c->nbr_off[q] * ncomp, cnt = (size_t)(c->nbr_off[q + 1] - c->nbr_off[q]) * ncomp;
    if ((r = api->Send(c->xsend.p + off, cnt, ncclDouble, c->nbr_rank[q], (ncclComm_t)c->comm, st)) != ncclSuccess) return nccl_fail("ncclSend", r);
    if ((r = api->Recv(c->xrecv.p + off, cnt, ncclDouble, c->nbr_rank[q], (ncclComm_t)c->comm, st)) != ncclSuccess) return nccl_fail("ncclRecv", r);
  }
  if (ns > 0 && (r = api->AllReduce(scal, scal, (size_t)ns, ncclDouble, ncclSum, (ncclComm_t)c->comm, st)) != ncclSuccess) return nccl_fail("ncclAllReduce", r);
  if ((r = api->GroupEnd()) != ncclSuccess) return nccl_fail("ncclGroupEnd", r);
  return MFB_OK;
}
int halo_buffers(mfb_ctx* c, int ncomp, double** send, double** recv) {
  int rc = ensure_xbuf(c, ncomp);
  if (rc) return rc;
  *send = c->xsend.p; *recv = c->xrecv.p;
  return MFB_OK;
}
int halo_exchange(mfb_ctx* c, int ncomp, double* scal, int ns) {
  if (c->world <= 1) return MFB_OK;
  int rc = ensure_xbuf(c, ncomp);
  if (rc) return rc;
  return halo_exchange_on(c, ncomp, scal, ns, c->stream);
}

// the whole point-to-point interface sum of (a | b) on stream st
int p2p_interface_sum_on(mfb_ctx* c, double* a, int na, double* b, int nb, double* scal, int ns, cudaStream_t st) {
  const int w = na + (b ? nb : 0);
  int rc = ensure_xbuf(c, w);
  if (rc) return rc;
  const int64_t n = (int64_t)c->n_ent * w;
  if (n > 0) k_p2p_pack<<<nblk(n), 256, 0, st>>>(c->ent_local.p, c->n_ent, a, na, b, b ? nb : 0, c->xsend.p);
  if ((rc = halo_exchange_on(c, w, scal, ns, st))) return rc;
  if (n > 0) k_p2p_unpack_add<<<nblk(n), 256, 0, st>>>(c->ent_local.p, c->n_ent, a, na, b, b ? nb : 0, c->xrecv.p);
  c->launches += 2;
  MFB_CUDA(cudaGetLastError());
  return MFB_OK;
}

// a (na per free vertex) and optionally b (nb per free vertex) <- sums over ranks at interface vertices; `scal` (ns
// doubles, device) is summed over ranks in the same collective.
int interface_sum(mfb_ctx* c, double* a, int na, double* b, int nb, double* scal, int ns) {
  if (c->world <= 1) return MFB_OK;
  if (c->use_p2p && c->ent_local.p) return p2p_interface_sum_on(c, a, na, b, nb, scal, ns, c->stream);
  if (!c->if_inv.p) { g_last_error = "interface_sum: no interface description (mfb_comm_set_interface / mfb_comm_set_partition)"; return MFB_ERR_INVALID; }
  NcclApi* api = nccl();
  const int w = na + (b ? nb : 0);
  const size_t need = (size_t)c->n_shared * w + ns;
  if (c->if_buf.n < need + 1) { int rc = c->if_buf.alloc(need + 1); if (rc) return rc; }
  cudaStream_t st = c->stream;
  if (need == 0) return MFB_OK;
  k_if_pack<<<nblk((int64_t)need), 256, 0, st>>>(c->if_inv.p, c->n_shared, a, na, b, b ? nb : 0, scal, ns, c->if_buf.p);
  ncclResult_t r = api->AllReduce(c->if_buf.p, c->if_buf.p, need, ncclDouble, ncclSum, (ncclComm_t)c->comm, st);
  if (r != ncclSuccess) return nccl_fail("ncclAllReduce", r);
  k_if_unpack<<<nblk((int64_t)need), 256, 0, st>>>(c->if_inv.p, c->n_shared, a, na, b, b ? nb : 0, scal, ns, c->if_buf.p);
  c->launches += 2;
  MFB_CUDA(cudaGetLastError());
  return MFB_OK;
}

int allreduce_scalars(mfb_ctx* c, double* v, int n, bool maxop) {
  if (c->world <= 1) return MFB_OK;
  ncclResult_t r = nccl()->AllReduce(v, v, (size_t)n, ncclDouble, maxop ? ncclMax : ncclSum, (ncclComm_t)c->comm, c->stream);
  if (r != ncclSuccess) return nccl_fail("ncclAllReduce", r);
  return MFB_OK;
}

// ---- peer mailboxes: see PeerView in mfb_ctx.cuh --------------------------------------------------------------------
static void peer_free(mfb_ctx* c) {
  for (int r = 0; r < kPeerMax; ++r) if (c->peer_base[r]) { cudaIpcCloseMemHandle(c->peer_base[r]); c->peer_base[r] = nullptr; }
  if (c->peer_box) { cudaFree(c->peer_box); c->peer_box = nullptr; }
  c->peer_state = 0;
  (void)cudaGetLastError();
}

// Collective over the communicator (every rank calls it at the same point: the first distributed mfb_pcg_dev after
// mfb_comm_set_partition).  Allocates and exports this rank's mailbox, gathers everybody's IPC handle and receive offsets
// (grouped ncclSend / ncclRecv), opens the others' mailboxes and builds the destination table of the halo stores.  Any
// failure on any rank (no IPC in this container, no peer access between two devices) leaves peer_state = -1 on EVERY
// rank - the decision is taken from gathered data - and the CG stays on the NCCL exchange.
int peer_setup(mfb_ctx* c) {
  if (c->peer_state != 0 || !c->use_peer) return MFB_OK;        // option peer = 0: not tried (state stays 0)
  c->peer_state = -1;
  if (c->world <= 1 || c->world > kPeerMax || !c->ent_local.p) return MFB_OK;
  NcclApi* api = nccl();
  cudaStream_t st = c->stream;
  const int W = c->world, me = c->rank;
  const size_t words = (size_t)kPeerHaloWord0 + 2 * (size_t)c->n_ent * 3 + 8;
  long long ok = 1;
  void* box = nullptr;
  cudaIpcMemHandle_t h;
  memset(&h, 0, sizeof(h));
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
  if (cudaMalloc(&box, words * 8) != cudaSuccess) { ok = 0; box = nullptr; }
  if (ok && cudaMemset(box, 0, words * 8) != cudaSuccess) ok = 0;
  if (ok && cudaIpcGetMemHandle(&h, box) != cudaSuccess) ok = 0;
  (void)cudaGetLastError();
  // blob per rank: [0,8) handle, [8] ok, [9] n_ent, [10 + r] first entry of the segment this rank receives from rank r (-1: none)
  const int bw = 10 + W;
  std::vector<long long> mine((size_t)bw, -1), all((size_t)bw * W, 0);
  memcpy(mine.data(), &h, 64);
  mine[8] = ok; mine[9] = c->n_ent;
  for (int q = 0; q < c->n_nbr; ++q) mine[10 + c->nbr_rank[q]] = c->nbr_off[q];
  DBuf<long long> dm, da;
  int rc;
  if ((rc = dm.alloc((size_t)bw)) || (rc = da.alloc((size_t)bw * W))) { if (box) cudaFree(box); return rc; }
  MFB_CUDA(cudaMemcpyAsync(dm.p, mine.data(), (size_t)bw * 8, cudaMemcpyHostToDevice, st));
  ncclResult_t r = api->GroupStart();
  if (r != ncclSuccess) return nccl_fail("ncclGroupStart", r);
  for (int q = 0; q < W; ++q) {
    if (q == me) continue;
    if ((r = api->Send(dm.p, (size_t)bw, ncclInt64, q, (ncclComm_t)c->comm, st)) != ncclSuccess) return nccl_fail("ncclSend", r);
    if ((r = api->Recv(da.p + (size_t)q * bw, (size_t)bw, ncclInt64, q, (ncclComm_t)c->comm, st)) != ncclSuccess) return nccl_fail("ncclRecv", r);
  }
  if ((r = api->GroupEnd()) != ncclSuccess) return nccl_fail("ncclGroupEnd", r);
  MFB_CUDA(cudaMemcpyAsync(all.data(), da.p, (size_t)bw * W * 8, cudaMemcpyDeviceToHost, st));
  MFB_CUDA(cudaStreamSynchronize(st));
  memcpy(&all[(size_t)me * bw], mine.data(), (size_t)bw * 8);
  bool all_ok = true;
  for (int q = 0; q < W; ++q) all_ok = all_ok && all[(size_t)q * bw + 8] == 1;
  // the two sides of every cut must describe the same segment
  for (int q = 0; q < c->n_nbr && all_ok; ++q) all_ok = all[(size_t)c->nbr_rank[q] * bw + 10 + me] >= 0;
  if (!all_ok) { if (box) cudaFree(box); (void)cudaGetLastError(); return MFB_OK; }
  c->peer_box = box;
  double opened = 1.0;
  for (int q = 0; q < W; ++q) {
    if (q == me) continue;
    cudaIpcMemHandle_t hq;
    memcpy(&hq, &all[(size_t)q * bw], 64);
    if (cudaIpcOpenMemHandle(&c->peer_base[q], hq, cudaIpcMemLazyEnablePeerAccess) != cudaSuccess) { c->peer_base[q] = nullptr; opened = 0.0; (void)cudaGetLastError(); }
  }
  {   // did everybody open everything?  min over ranks = -max(-x)
    DBuf<double> flag;
    if ((rc = flag.alloc(1))) return rc;
    const double neg = -opened;
    MFB_CUDA(cudaMemcpyAsync(flag.p, &neg, 8, cudaMemcpyHostToDevice, st));
    if ((rc = allreduce_scalars(c, flag.p, 1, true))) return rc;
    double res = 0.0;
    MFB_CUDA(cudaMemcpyAsync(&res, flag.p, 8, cudaMemcpyDeviceToHost, st));
    MFB_CUDA(cudaStreamSynchronize(st));
    if (res != -1.0) { peer_free(c); c->peer_state = -1; return MFB_OK; }
  }
  PeerView pv;
  memset(&pv, 0, sizeof(pv));
  pv.world = W; pv.rank = me; pv.n_ent = c->n_ent;
  for (int q = 0; q < W; ++q) {
    unsigned long long* b = (unsigned long long*)(q == me ? c->peer_box : c->peer_base[q]);
    for (int stg = 0; stg < 2; ++stg) {
      pv.flag_at[stg][q] = b + stg * kPeerMax + me;
      pv.sc_at[stg][q] = (double*)b + kPeerFlagWords + (size_t)stg * 2 * kPeerMax * 2 + (size_t)me * 2;
    }
  }
  pv.my_flag = (unsigned long long*)c->peer_box;
  pv.my_sc = (double*)c->peer_box + kPeerFlagWords;
  pv.my_halo = (double*)c->peer_box + kPeerHaloWord0;
  std::vector<double*> dst((size_t)2 * c->n_ent + 1, nullptr);
  for (int q = 0; q < c->n_nbr; ++q) {
    const int rr = c->nbr_rank[q];
    const long long off = all[(size_t)rr * bw + 10 + me], n_ent_r = all[(size_t)rr * bw + 9];
    double* halo_r = (double*)c->peer_base[rr] + kPeerHaloWord0;
    for (int sidx = c->nbr_off[q]; sidx < c->nbr_off[q + 1]; ++sidx)
      for (int par = 0; par < 2; ++par)
        dst[(size_t)par * c->n_ent + sidx] = halo_r + (size_t)par * 3 * n_ent_r + (size_t)(off + (sidx - c->nbr_off[q])) * c->dim;
  }
  if ((rc = c->peer_dst.alloc(dst.size())) || (rc = c->peer_view.alloc(1))) return rc;
  MFB_CUDA(cudaMemcpyAsync(c->peer_dst.p, dst.data(), dst.size() * sizeof(double*), cudaMemcpyHostToDevice, st));
  pv.dst = c->peer_dst.p;
  MFB_CUDA(cudaMemcpyAsync(c->peer_view.p, &pv, sizeof(pv), cudaMemcpyHostToDevice, st));
  MFB_CUDA(cudaStreamSynchronize(st));
  c->peer_seq = 0;
  c->peer_state = 1;
  return MFB_OK;
}

void comm_free(mfb_ctx* c) {
  peer_free(c);
  if (c->xstream) { cudaStreamDestroy(c->xstream); c->xstream = 0; }
  if (c->ev_x0) { cudaEventDestroy(c->ev_x0); cudaEventDestroy(c->ev_x1); c->ev_x0 = c->ev_x1 = 0; }
  if (c->comm && nccl()->ok) nccl()->CommDestroy((ncclComm_t)c->comm);
  c->comm = nullptr; c->world = 1; c->rank = 0;
}

}  // namespace mfb

extern "C" int mfb_comm_unique_id(char* id128) {
  if (!id128) return MFB_ERR_INVALID;
  NcclApi* api = nccl();
  if (!api->ok) { g_last_error = "NCCL (libnccl.so.2) could not be loaded"; return MFB_ERR_UNSUPPORTED; }
  ncclUniqueId id;
  ncclResult_t r = api->GetUniqueId(&id);
  if (r != ncclSuccess) return nccl_fail("ncclGetUniqueId", r);
  static_assert(sizeof(ncclUniqueId) == 128, "ncclUniqueId is 128 bytes");
  memcpy(id128, &id, 128);
  return MFB_OK;
}

extern "C" int mfb_comm_init(mfb_ctx* c, const char* id128, int rank, int world) {
  CHECK_CTX(c);
  if (!id128 || world < 1 || rank < 0 || rank >= world) return MFB_ERR_INVALID;
  NcclApi* api = nccl();
  if (!api->ok) { g_last_error = "NCCL (libnccl.so.2) could not be loaded"; return MFB_ERR_UNSUPPORTED; }
  comm_free(c);
  ncclUniqueId id;
  memcpy(&id, id128, 128);
  ncclComm_t comm;
  ncclResult_t r = api->CommInitRank(&comm, world, id, rank);
  if (r != ncclSuccess) return nccl_fail("ncclCommInitRank", r);
  c->comm = comm; c->rank = rank; c->world = world;
  return MFB_OK;
}

extern "C" int mfb_comm_set_interface(mfb_ctx* c, const int32_t* if_local, const int32_t* if_pos, int32_t n_if_local, int32_t n_shared,
                                      const double* dot_weight) {
  CHECK_CTX(c);
  if (n_if_local < 0 || n_shared < 0 || (n_if_local > 0 && (!if_local || !if_pos)) || !dot_weight) return MFB_ERR_INVALID;
  {   // a bad description would make the pack / unpack kernels read and write out of bounds: check it on the host
    std::vector<char> seen((size_t)n_shared, 0);
    for (int i = 0; i < n_if_local; ++i) {
      if (if_local[i] < 0 || if_local[i] >= c->nF || if_pos[i] < 0 || if_pos[i] >= n_shared || seen[if_pos[i]]) {
        g_last_error = "mfb_comm_set_interface: if_local / if_pos out of range or slots not unique";
        return MFB_ERR_INVALID;
      }
      seen[if_pos[i]] = 1;
    }
  }
  int rc;
  if ((rc = c->if_local.alloc((size_t)n_if_local + 1)) || (rc = c->if_pos.alloc((size_t)n_if_local + 1)) || (rc = c->dot_w.alloc((size_t)c->nF + 1))) return rc;
  if (n_if_local > 0) {
    MFB_CUDA(cudaMemcpyAsync(c->if_local.p, if_local, (size_t)n_if_local * 4, cudaMemcpyHostToDevice, c->stream));
    MFB_CUDA(cudaMemcpyAsync(c->if_pos.p, if_pos, (size_t)n_if_local * 4, cudaMemcpyHostToDevice, c->stream));
  }
  MFB_CUDA(cudaMemcpyAsync(c->dot_w.p, dot_weight, (size_t)c->nF * 8, cudaMemcpyHostToDevice, c->stream));
  if ((rc = c->if_inv.alloc((size_t)n_shared + 1))) return rc;
  MFB_CUDA(cudaMemsetAsync(c->if_inv.p, 0xff, ((size_t)n_shared + 1) * 4, c->stream));       // -1
  if (n_if_local > 0) k_if_invert<<<nblk(n_if_local), 256, 0, c->stream>>>(c->if_local.p, c->if_pos.p, n_if_local, c->if_inv.p);
  MFB_CUDA(cudaStreamSynchronize(c->stream));
  c->n_if_local = n_if_local; c->n_shared = n_shared;
  return MFB_OK;
}

extern "C" int mfb_comm_set_partition(mfb_ctx* c, const mfb_partition* p) {
  CHECK_CTX(c);
  PartitionView v;
  if (partition_view(p, &v)) { g_last_error = "mfb_comm_set_partition: null partition"; return MFB_ERR_INVALID; }
  if (v.nF != c->nF || v.nE != c->nE) { g_last_error = "mfb_comm_set_partition: the partition is not the one this context was created from"; return MFB_ERR_INVALID; }
  if (c->world != v.nparts || c->rank != v.rank) { g_last_error = "mfb_comm_set_partition: call mfb_comm_init with the partition's rank / world first"; return MFB_ERR_INVALID; }
  int rc = mfb_comm_set_interface(c, v.if_local->data(), v.if_pos->data(), (int32_t)v.if_local->size(), v.n_shared, v.dot_weight->data());
  if (rc) return rc;
  c->n_nbr = (int)v.nbr_rank->size();
  c->nbr_rank.assign(v.nbr_rank->begin(), v.nbr_rank->end());
  c->nbr_off.assign((size_t)c->n_nbr + 1, 0);
  for (int q = 0; q < c->n_nbr; ++q) c->nbr_off[q + 1] = c->nbr_off[q] + (*v.nbr_count)[q];
  c->n_ent = c->nbr_off[c->n_nbr];
  c->n_if_elems = v.n_if_elems;
  c->xbuf_comp = 0;
  peer_free(c);                  // sized and addressed for the previous partition, if any
  // per local free row: its entries
  std::vector<int> ptr((size_t)c->nF + 1, 0), idx((size_t)c->n_ent);
  for (int i = 0; i < c->n_ent; ++i) {
    const int l = (*v.nbr_local)[i];
    if (l < 0 || l >= c->nF) { g_last_error = "mfb_comm_set_partition: entry out of range"; return MFB_ERR_INVALID; }
    ptr[l + 1] += 1;
  }
  for (int l = 0; l < c->nF; ++l) ptr[l + 1] += ptr[l];
  { std::vector<int> fill(ptr.begin(), ptr.end() - 1);
    for (int i = 0; i < c->n_ent; ++i) idx[fill[(*v.nbr_local)[i]]++] = i; }
  if ((rc = c->ent_local.alloc((size_t)c->n_ent + 1)) || (rc = c->rowif_ptr.alloc((size_t)c->nF + 1)) || (rc = c->rowif_idx.alloc((size_t)c->n_ent + 1))) return rc;
  if (c->n_ent) {
    MFB_CUDA(cudaMemcpyAsync(c->ent_local.p, v.nbr_local->data(), (size_t)c->n_ent * 4, cudaMemcpyHostToDevice, c->stream));
    MFB_CUDA(cudaMemcpyAsync(c->rowif_idx.p, idx.data(), (size_t)c->n_ent * 4, cudaMemcpyHostToDevice, c->stream));
  }
  MFB_CUDA(cudaMemcpyAsync(c->rowif_ptr.p, ptr.data(), ((size_t)c->nF + 1) * 4, cudaMemcpyHostToDevice, c->stream));
  MFB_CUDA(cudaStreamSynchronize(c->stream));
  return MFB_OK;
}

extern "C" int mfb_comm_peer_state(mfb_ctx* c) { return c ? c->peer_state : 0; }

extern "C" int mfb_interface_sum_dev(mfb_ctx* c, double* a_dev, int ncomp_a, double* b_dev, int ncomp_b) {
  CHECK_CTX(c);
  if (!a_dev || ncomp_a <= 0 || (b_dev && ncomp_b <= 0)) return MFB_ERR_INVALID;
  if (c->world > 1 && !c->dot_w.p) { g_last_error = "mfb_interface_sum_dev: mfb_comm_set_interface has not been called"; return MFB_ERR_INVALID; }
  c->ev_vec_valid = false;     // rhs_ / grad_x_ may be among the summed vectors: host readers must wait for this stream
  return interface_sum(c, a_dev, ncomp_a, b_dev, ncomp_b, nullptr, 0);
}

extern "C" int mfb_allreduce_dev(mfb_ctx* c, double* v_dev, int n, int op) {
  CHECK_CTX(c);
  if (!v_dev || n <= 0) return MFB_ERR_INVALID;
  return allreduce_scalars(c, v_dev, n, op == 1);
}
