// mfb_energy_noalloc_dev: MixedStretch::energy() (mixed_stretch.cpp:93-107) for callers that must not allocate - code under
// CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  NO COUNTERPART IN THE REFERENCE.
// Same kernels as mfb_energy_dev (k_energy + k_sum_partials: per-block partial sums, then
// one block adds them in index order), but
//   * the scratch of the two-stage sum - one double per block of 128 elements - lives in a file-scope static __device__
//     array of MFB_NOALLOC_MAX_PARTIALS doubles (1 MiB; every GPU that loads the library holds it until exit) instead of
//     a buffer owned by the context;
//   * the result goes to a DEVICE pointer of the caller, stream-ordered on the context stream: no host copy, no
//     synchronisation, no event creation (the phase timers of mfb_energy_dev are not touched);
//   * nothing is allocated or freed: no cudaMalloc / cudaFree / cudaMallocHost / cudaEventCreate, no library routine that
//     allocates.  The one-time cudaGetSymbolAddress per device happens on the FIRST call on that device (call once before
//     starting a capture).
// Every call on a device shares the array: calls that overlap on the GPU (two contexts on different streams) race; order
// them on one stream or with events.  The array keeps the last call's contents; nothing in it has to start at zero -
// k_energy writes partials[0, grid) and k_sum_partials reads exactly that range.
#include "mfb_ctx.cuh"

using namespace mfb;

#define CHECK_CTX(c) do { if (!(c)) { g_last_error = "null context"; return MFB_ERR_INVALID; } cudaSetDevice((c)->device); } while (0)

static __device__ __align__(256) double g_noalloc_partials[MFB_NOALLOC_MAX_PARTIALS];

// In host code `g_noalloc_partials` is a host-side stand-in, not a device address: the kernels get the address that
// cudaGetSymbolAddress returns for the CURRENT device, looked up once per device ordinal.
static int noalloc_scratch(int device, double** out) {
  static double* addr[64] = {nullptr};
  if (device < 0 || device >= 64) { g_last_error = "mfb_energy_noalloc_dev: device ordinal out of range"; return MFB_ERR_INVALID; }
  if (!addr[device]) {
    double* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress((void**)&p, g_noalloc_partials);
    if (e != cudaSuccess || !p) {
      g_last_error = std::string("cudaGetSymbolAddress(g_noalloc_partials): ") + cudaGetErrorString(e);
      return MFB_ERR_CUDA;
    }
    addr[device] = p;
  }
  *out = addr[device];
  return MFB_OK;
}

extern "C" int64_t mfb_noalloc_max_elements(void) { return (int64_t)MFB_NOALLOC_MAX_PARTIALS * 128; }

extern "C" int mfb_energy_noalloc_dev(mfb_ctx* c, const double* s_dev_soa_or_null, double* e_dev) {
  CHECK_CTX(c);
  if (!e_dev) { g_last_error = "mfb_energy_noalloc_dev: e_dev is NULL"; return MFB_ERR_INVALID; }
  const int grid = c->grid();
  if (grid > MFB_NOALLOC_MAX_PARTIALS) {
    g_last_error = "mfb_energy_noalloc_dev: " + std::to_string(c->nE) + " elements exceed the static scratch (" +
                   std::to_string((long long)MFB_NOALLOC_MAX_PARTIALS * 128) + " elements); use mfb_energy_dev";
    return MFB_ERR_TOO_LARGE;
  }
  double* scratch = nullptr;
  int rc = noalloc_scratch(c->device, &scratch);
  if (rc) return rc;
  const double* sp = s_dev_soa_or_null ? s_dev_soa_or_null : c->s.p;
  cudaStream_t st = c->stream;
  if (c->kind == MFB_SHELL) k_energy<K_SHELL><<<grid, 128, 0, st>>>(c->mesh(), sp, scratch);
  else if (c->dim == 3) k_energy<3><<<grid, 128, 0, st>>>(c->mesh(), sp, scratch);
  else k_energy<2><<<grid, 128, 0, st>>>(c->mesh(), sp, scratch);
  k_sum_partials<<<1, 256, 0, st>>>(scratch, grid, e_dev);
  c->launches += 2;
  MFB_CUDA(cudaGetLastError());
  return MFB_OK;
}
