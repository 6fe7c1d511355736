// State access: host copies in the reference's array-of-structs layouts, setters for the members the reference's
// callers mutate in place (value(), lambda(), delta(): src/linesearch.h:120-125, src/linear_solvers/linear_system.h:97),
// and raw device pointers of the structure-of-arrays storage.
#include "mfb_ctx.cuh"

using namespace mfb;

#define CHECK_CTX(c) do { if (!(c)) { g_last_error = "null context"; return MFB_ERR_INVALID; } cudaSetDevice((c)->device); } while (0)
static inline unsigned nblk(int64_t n, int t = 256) { return (unsigned)((n + t - 1) / t); }

namespace {
struct FieldInfo { double* soa; int ncomp; int kind; };  // kind: 0 soa per-element, 1 packed symmetric, 2 vertex vector, 3 special
FieldInfo info(mfb_ctx* c, int f) {
  switch (f) {
    case MFB_F_S: return {c->s.p, c->N, 0};
    case MFB_F_DS: return {c->ds.p, c->N, 0};
    case MFB_F_LA: return {c->la.p, c->N, 0};
    case MFB_F_R: return {c->R.p, c->M, 0};
    case MFB_F_SROT: return {c->S.p, c->N, 0};
    case MFB_F_G: return {c->g.p, c->N, 0};
    case MFB_F_H: return {c->H.p, c->N * c->N, 1};
    case MFB_F_HINV: return {c->Hinv.p, c->N * c->N, 1};
    case MFB_F_HLOC: return {c->H.p, c->N * c->N, 1};
    case MFB_F_DSDF: return {c->dSdF.p, c->M * c->N, 0};
    case MFB_F_EVALS: return {c->evals.p, 1, 0};
    case MFB_F_GL: return {c->gl.p, c->N, 0};
    case MFB_F_GRAD: return {c->g.p, c->N, 3};
    case MFB_F_GRAD_LA: return {c->grad_la.p, c->N, 0};
    case MFB_F_GRAD_X: return {c->grad_x.p, c->dim, 2};
    case MFB_F_RHS: return {c->rhs.p, c->dim, 2};
    case MFB_F_VOLS: return {c->vol.p, 1, 0};
    case MFB_F_ITERS: return {nullptr, 1, 3};
  }
  return {nullptr, 0, -1};
}
}  // namespace

extern "C" int64_t mfb_field_size(const mfb_ctx* c, int f) {
  if (!c) return -1;
  FieldInfo fi = info(const_cast<mfb_ctx*>(c), f);
  if (fi.kind < 0) return -1;
  if (fi.kind == 2) return (int64_t)c->dim * c->nF;
  return (int64_t)fi.ncomp * c->nE;
}

extern "C" int mfb_get(mfb_ctx* c, int f, double* out) {
  CHECK_CTX(c);
  if (!out) return MFB_ERR_INVALID;
  FieldInfo fi = info(c, f);
  if (fi.kind < 0) { g_last_error = "unknown field"; return MFB_ERR_INVALID; }
  cudaStream_t st = c->stream;
  const int64_t n = mfb_field_size(c, f);
  if (fi.kind == 2) {
    if (f == MFB_F_RHS && !c->rhs_valid) { const double* p; int rc = mfb_rhs_dev(c, &p); if (rc) return rc; }
    MFB_CUDA(cudaMemcpyAsync(out, fi.soa, n * 8, cudaMemcpyDeviceToHost, st));
    MFB_CUDA(cudaStreamSynchronize(st));
    return MFB_OK;
  }
  if (c->aos.n < (size_t)n) { int rc = c->aos.alloc((size_t)c->M * c->N * c->nE); if (rc) return rc; }
  if (fi.kind == 0) k_soa_to_aos<<<nblk(n), 256, 0, st>>>(fi.soa, c->aos.p, fi.ncomp, c->nE, c->ld);
  else if (fi.kind == 1) k_sym_to_full<<<nblk(n), 256, 0, st>>>(fi.soa, c->vol.p, c->aos.p, c->N, c->dim, c->nE, c->ld, f == MFB_F_HLOC ? 1 : 0);
  else if (f == MFB_F_GRAD) k_grad_mixed<<<nblk(n), 256, 0, st>>>(c->g.p, c->vol.p, c->aos.p, c->N, c->nE, c->ld);
  else k_int_to_double<<<nblk(n), 256, 0, st>>>(c->iters.p, c->aos.p, c->nE);
  c->launches += 1;
  MFB_CUDA(cudaGetLastError());
  MFB_CUDA(cudaMemcpyAsync(out, c->aos.p, n * 8, cudaMemcpyDeviceToHost, st));
  MFB_CUDA(cudaStreamSynchronize(st));
  return MFB_OK;
}

extern "C" int mfb_set(mfb_ctx* c, int f, const double* in) {
  CHECK_CTX(c);
  if (!in) return MFB_ERR_INVALID;
  if (f != MFB_F_S && f != MFB_F_DS && f != MFB_F_LA && f != MFB_F_R) { g_last_error = "field is read-only"; return MFB_ERR_INVALID; }
  FieldInfo fi = info(c, f);
  const int64_t n = mfb_field_size(c, f);
  cudaStream_t st = c->stream;
  if (c->aos.n < (size_t)n) { int rc = c->aos.alloc((size_t)c->M * c->N * c->nE); if (rc) return rc; }
  MFB_CUDA(cudaMemcpyAsync(c->aos.p, in, n * 8, cudaMemcpyHostToDevice, st));
  k_aos_to_soa<<<nblk(n), 256, 0, st>>>(c->aos.p, fi.soa, fi.ncomp, c->nE, c->ld);
  c->launches += 1;
  MFB_CUDA(cudaGetLastError());
  MFB_CUDA(cudaStreamSynchronize(st));
  if (f != MFB_F_DS) c->rhs_valid = false;   // the fused rhs_/gl_ no longer match the state
  return MFB_OK;
}

extern "C" int mfb_field_dev(mfb_ctx* c, int f, double** ptr, int64_t* ld, int32_t* ncomp) {
  CHECK_CTX(c);
  FieldInfo fi = info(c, f);
  if (fi.kind < 0 || fi.kind == 3 || f == MFB_F_HLOC) { g_last_error = "field has no device storage"; return MFB_ERR_INVALID; }
  if (ptr) *ptr = fi.soa;
  if (ld) *ld = (fi.kind == 2) ? 0 : c->ld;
  if (ncomp) *ncomp = (fi.kind == 1) ? c->NP : fi.ncomp;
  if (f == MFB_F_S || f == MFB_F_LA || f == MFB_F_R) c->rhs_valid = false;   // caller may write through the pointer
  return MFB_OK;
}
