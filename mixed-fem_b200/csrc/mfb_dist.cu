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
#include "mfb_ctx.cuh"
#include <dlfcn.h>
#include <nccl.h>

using namespace mfb;

#define CHECK_CTX(c) do { if (!(c)) { g_last_error = "null context"; return MFB_ERR_INVALID; } cudaSetDevice((c)->device); } while (0)

namespace {

struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
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
      api.ok = api.GetUniqueId && api.CommInitRank && api.AllReduce && api.CommDestroy && api.GetErrorString;
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

}  // namespace

namespace mfb {

// a (na per free vertex) and optionally b (nb per free vertex) <- sums over ranks at interface vertices; `scal` (ns
// doubles, device) is summed over ranks in the same collective.
int interface_sum(mfb_ctx* c, double* a, int na, double* b, int nb, double* scal, int ns) {
  if (c->world <= 1) return MFB_OK;
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

void comm_free(mfb_ctx* c) {
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
