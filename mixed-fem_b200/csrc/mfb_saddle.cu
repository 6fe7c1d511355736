// The two saddle-point formulations of the mixed system, on the device and matrix-free:
//
//   SystemMatrixIndefinite<Scalar,DIM> ... src/linear_solvers/linear_system.h:64-118   (rhs, post_solve)
//   BlockMatrix<DIM> product ............. src/linear_solvers/block_matrix.h:74-125    (the KKT operator)
//   DualCondensedSystem<DIM> ............. src/linear_solvers/linear_system.h:120-239  (lhs, rhs, post_solve)
//   MixedStretch<DIM>::product_* ......... src/variables/mixed_stretch.cpp:406-468
//   Displacement<DIM>::product_hessian ... src/variables/displacement.h:63-67 (lhs_ = PMP^T, displacement.cpp:92)
//
// for a state with ONE mesh, the Displacement variable and ONE mixed-stretch variable (what mfb_newton_init set up).
//
//   KKT unknown  v = [dx (d nF) ; ds (N nE) ; la (N nE)]:
//        [ PMP^T        0        B   ] [dx]   [-grad_x(Displacement)]         B  = jacobian_x()  = -PJW blockdiag(dSdF)
//        [   0      blockdiag(Hloc) C ] [ds] = [-gradient_mixed()     ]       C  = jacobian_mixed() = blockdiag(vol Sym)
//        [  B^T         C        0   ] [la]   [-gradient_dual()      ]
//   The reference evaluates the product as eight separate passes (one per product_* call); here the element part is
//   ONE kernel (Hloc ds + C la, -dSdF^T J^T dx + C ds, and the scatter of -PJW dSdF la) plus one mass pass.
//
//   Dual-condensed:  (G^T G + C C Hinv) la = B^T Ml^-1 bx - C Hinv bs + bla,  G = Ml^-1/2 B,  Ml = lumped PMP^T,
//   Hinv = blockdiag(vol Hinv_).  The reference forms G^T G explicitly (a sparse matrix over element pairs sharing a
//   vertex: ~70 6x6 blocks per tet row); here it is applied as scatter -> diagonal scale -> gather, two element passes.
//
// Solvers: conjugate gradients (the Eigen ConjugateGradient recurrence the reference instantiates for the condensed
// system, identity preconditioner) and MINRES in the Lanczos form Eigen's MINRES uses for SystemMatrixIndefinite
// (identity preconditioner: the reference's BlockDiagonal / ADMM preconditioners need sparse direct factorisations,
// out of scope).  Eigen's MINRES lives in Eigen/unsupported (3.4.0), not under the reference tree: the recurrence is
// the published Paige-Saunders algorithm.
//
// Device vectors use the library's element layout: [x part d*nF, padded to a multiple of 32 | ds component-major
// N x ld | la component-major N x ld]; host entry points take the reference's layout (N*e + n) and convert.
#include "mfb_ctx.cuh"
#include <cmath>
#include <functional>

using namespace mfb;

#define CHECK_CTX(c) do { if (!(c)) { g_last_error = "null context"; return MFB_ERR_INVALID; } cudaSetDevice((c)->device); } while (0)

namespace {

inline unsigned nblk(int64_t n, int t = 256) { return (unsigned)((n + t - 1) / t); }
constexpr int kRedBlocks = 592;   // 4 x 148

template <int K> struct MassC { static constexpr double c0 = Dims<K>::NPE == 4 ? 0.1 : 1.0 / 6.0, c1 = Dims<K>::NPE == 4 ? 0.05 : 1.0 / 12.0; };

// out(free) += PMP^T v(free): element mass blocks mass*(c0 | c1) I_d applied to the free vertices of each element
template <int K>
__global__ void __launch_bounds__(128) k_mass_free_apply(DevMesh m, const double* __restrict__ emass, const double* __restrict__ v, double* __restrict__ out) {
  constexpr int DIM = Dims<K>::SD, NPE = Dims<K>::NPE;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m.nE) return;
  const int64_t ld = m.ld;
  const double mass = emass[e];
  double d[NPE][DIM], sum[DIM];
  int f[NPE];
#pragma unroll
  for (int l = 0; l < DIM; ++l) sum[l] = 0.0;
#pragma unroll
  for (int k = 0; k < NPE; ++k) {
    f[k] = m.fidx[k * ld + e];
#pragma unroll
    for (int l = 0; l < DIM; ++l) { d[k][l] = f[k] >= 0 ? v[DIM * (int64_t)f[k] + l] : 0.0; sum[l] += d[k][l]; }
  }
#pragma unroll
  for (int j = 0; j < NPE; ++j) {
    if (f[j] < 0) continue;
#pragma unroll
    for (int l = 0; l < DIM; ++l) red_add(&out[DIM * (int64_t)f[j] + l], mass * (MassC<K>::c1 * sum[l] + (MassC<K>::c0 - MassC<K>::c1) * d[j][l]));
  }
}

// Element part of the BlockMatrix product (block_matrix.h:84-124), all three block rows in one pass:
//   out_x  -= PJW dSdF la                      (product_jacobian_x(la, ., true))
//   out_s   = Hloc ds + vol Sym la             (product_hessian + product_jacobian_mixed)
//   out_la  = -dSdF^T (PJW^T dx) + vol Sym ds  (product_jacobian_x(dx, ., false) + product_jacobian_mixed)
template <int K>
__global__ void __launch_bounds__(128) k_kkt_elem(DevMesh m, DevState st, const double* __restrict__ dx, const double* __restrict__ ds,
                                                  const double* __restrict__ la, double* __restrict__ out_x, double* __restrict__ out_s,
                                                  double* __restrict__ out_la) {
  using D = Dims<K>;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m.nE) return;
  const int64_t ld = m.ld;
  const double vol = m.vol[e];
  double dsv[D::N], lav[D::N];
#pragma unroll
  for (int n = 0; n < D::N; ++n) { dsv[n] = ds[n * ld + e]; lav[n] = la[n * ld + e]; }
#pragma unroll
  for (int r = 0; r < D::N; ++r) {
    double a = 0.0;
#pragma unroll
    for (int k = 0; k < D::N; ++k) a += vol * vol * symw<K>(r) * symw<K>(k) * st.H[symidx(r, k) * ld + e] * dsv[k];
    out_s[r * ld + e] = a + vol * symw<K>(r) * lav[r];
  }
  double Jdx[D::M], t[D::M], dphi[D::M];
  gather_Jdx<K>(m, e, dx, vol, Jdx);
#pragma unroll
  for (int q = 0; q < D::M; ++q) { t[q] = 0.0; dphi[q] = (K == K_SHELL) ? 0.0 : m.dphi[q * ld + e]; }
#pragma unroll
  for (int n = 0; n < D::N; ++n) {
    double a = 0.0;
#pragma unroll
    for (int q = 0; q < D::M; ++q) {
      const double j = st.dSdF[(n * D::M + q) * ld + e];
      a += j * Jdx[q];
      t[q] += j * lav[n];
    }
    out_la[n * ld + e] = -a + vol * symw<K>(n) * dsv[n];
  }
  scatter_vertex<K>(m, e, dphi, t, -vol, out_x);
}

// t(free) += B la = -PJW dSdF la
template <int K>
__global__ void __launch_bounds__(128) k_dual_scatter(DevMesh m, DevState st, const double* __restrict__ la, double* __restrict__ tvec) {
  using D = Dims<K>;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m.nE) return;
  const int64_t ld = m.ld;
  double t[D::M], dphi[D::M];
#pragma unroll
  for (int q = 0; q < D::M; ++q) {
    double a = 0.0;
#pragma unroll
    for (int n = 0; n < D::N; ++n) a += st.dSdF[(n * D::M + q) * ld + e] * la[n * ld + e];
    t[q] = a; dphi[q] = (K == K_SHELL) ? 0.0 : m.dphi[q * ld + e];
  }
  scatter_vertex<K>(m, e, dphi, t, -m.vol[e], tvec);
}

// out_e = B_e^T w + cH * (C C Hinv y)_e + cg * (C Hinv gradient_mixed)_e + cl * grad_la_e,  C = vol Sym, Hinv = vol Hinv_:
//   lhs product : w = Ml^-1 B la, y = la, cH = 1, cg = 0,  cl = 0    (linear_system.h:155: G^T G + C*C*H, in THAT order:
//                 the product Sym^2 Hinv_ is not symmetric; kept as the reference forms it)
//   rhs         : w = -Ml^-1 gx,  cH = 0, cg = -1, cl = 1            (linear_system.h:158-165)
template <int K>
__global__ void __launch_bounds__(128) k_dual_gather(DevMesh m, DevState st, const double* __restrict__ w, const double* __restrict__ y, double cH,
                                                     double cg, double cl, double* __restrict__ out) {
  using D = Dims<K>;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m.nE) return;
  const int64_t ld = m.ld;
  const double vol = m.vol[e];
  double Jw[D::M];
  gather_Jdx<K>(m, e, w, vol, Jw);
  double z[D::N];
#pragma unroll
  for (int k = 0; k < D::N; ++k) z[k] = (cH != 0.0) ? y[k * ld + e] : vol * st.g[k * ld + e];
#pragma unroll
  for (int n = 0; n < D::N; ++n) {
    double a = 0.0;
#pragma unroll
    for (int q = 0; q < D::M; ++q) a += st.dSdF[(n * D::M + q) * ld + e] * Jw[q];
    double hz = 0.0;
#pragma unroll
    for (int k = 0; k < D::N; ++k) hz += st.Hinv[symidx(n, k) * ld + e] * z[k];
    const double cn = vol * symw<K>(n);
    const double coef = (cH != 0.0) ? cH * cn * cn * vol : cg * cn * vol;
    out[n * ld + e] = -a + coef * hz + (cl != 0.0 ? cl * st.grad_la[n * ld + e] : 0.0);
  }
}

// ds_e = vol Hinv_e (-vol g_e - vol Sym la_e)   (DualCondensedSystem::post_solve, linear_system.h:199-205)
template <int K>
__global__ void __launch_bounds__(128) k_dual_ds(DevMesh m, DevState st, const double* __restrict__ la, double* __restrict__ ds) {
  using D = Dims<K>;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m.nE) return;
  const int64_t ld = m.ld;
  const double vol = m.vol[e];
  double z[D::N];
#pragma unroll
  for (int k = 0; k < D::N; ++k) z[k] = -vol * st.g[k * ld + e] - vol * symw<K>(k) * la[k * ld + e];
#pragma unroll
  for (int n = 0; n < D::N; ++n) {
    double hz = 0.0;
#pragma unroll
    for (int k = 0; k < D::N; ++k) hz += st.Hinv[symidx(n, k) * ld + e] * z[k];
    ds[n * ld + e] = vol * hz;
  }
}

// rows e < nE of a component-major element array: o = a * x   (gradient_mixed = vol g ; padding stays untouched)
__global__ void k_elem_scale_vol(const double* __restrict__ vol, const double* __restrict__ x, double a, int ncomp, int nE, int64_t ld, double* __restrict__ o) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)ncomp * nE) return;
  const int n = (int)(i / nE), e = (int)(i - (int64_t)n * nE);
  o[n * ld + e] = a * vol[e] * x[n * ld + e];
}
__global__ void k_elem_scale(const double* __restrict__ x, double a, int ncomp, int nE, int64_t ld, double* __restrict__ o) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (int64_t)ncomp * nE) return;
  const int n = (int)(i / nE), e = (int)(i - (int64_t)n * nE);
  o[n * ld + e] = a * x[n * ld + e];
}
__global__ void k_vec_mul(const double* __restrict__ a, const double* __restrict__ b, double s, int64_t n, double* __restrict__ o) {   // o = s a .* b
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) o[i] = s * a[i] * b[i];
}
__global__ void k_vec_inv(double* __restrict__ a, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) a[i] = 1.0 / a[i];
}
__global__ void k_fill(double* __restrict__ a, double v, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) a[i] = v;
}
// y = a x + b y
__global__ void k_axpby(double a, const double* x, double b, double* y, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) y[i] = a * x[i] + b * y[i];
}
// o = a x + b y + c z
__global__ void k_lin3(double a, const double* __restrict__ x, double b, const double* __restrict__ y, double c, const double* __restrict__ z, double* __restrict__ o, int64_t n) {
  const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) o[i] = a * x[i] + b * y[i] + c * z[i];
}
// deterministic dot product: fixed block partials, summed in index order by one block
__global__ void __launch_bounds__(256) k_dot_part(const double* __restrict__ a, const double* __restrict__ b, int64_t n, double* __restrict__ part) {
  __shared__ double sh[256];
  double acc = 0.0;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) acc += a[i] * b[i];
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) { if ((int)threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s]; __syncthreads(); }
  if (threadIdx.x == 0) part[blockIdx.x] = sh[0];
}
__global__ void __launch_bounds__(256) k_dot_finish(const double* __restrict__ part, int np, double* out) {
  __shared__ double sh[256];
  double acc = 0.0;
  for (int i = threadIdx.x; i < np; i += 256) acc += part[i];
  sh[threadIdx.x] = acc;
  __syncthreads();
  for (int s = 128; s > 0; s >>= 1) { if ((int)threadIdx.x < s) sh[threadIdx.x] += sh[threadIdx.x + s]; __syncthreads(); }
  if (threadIdx.x == 0) *out = sh[0];
}

struct Saddle {
  DBuf<double> part, scal;           // reduction scratch
  DBuf<double> Mlinv, gx, tvec, wvec; // lumped PMP^T inverse (d nF), Displacement gradient, scatter targets
  DBuf<double> work[8];              // solver vectors
  int64_t ox = 0, os = 0, ola = 0, nk = 0, ne = 0;   // offsets of the three KKT parts, total KKT length, N*ld
  bool lumped = false;
};

#define KDISPATCH(c, KERNEL, ...)                                                                     \
  do {                                                                                                 \
    if ((c)->kind == MFB_SHELL) KERNEL<K_SHELL><<<(c)->grid(), 128, 0, (c)->stream>>>(__VA_ARGS__);    \
    else if ((c)->dim == 3) KERNEL<3><<<(c)->grid(), 128, 0, (c)->stream>>>(__VA_ARGS__);              \
    else KERNEL<2><<<(c)->grid(), 128, 0, (c)->stream>>>(__VA_ARGS__);                                 \
    (c)->launches += 1;                                                                                \
  } while (0)

int saddle_get(mfb_ctx* c, Saddle** out) {
  if (!c->saddle) {
    Saddle* s = new Saddle();
    c->saddle = s;
    const int64_t nf = (int64_t)c->dim * c->nF;
    s->ne = (int64_t)c->N * c->ld;
    s->ox = 0; s->os = (nf + 31) / 32 * 32; s->ola = s->os + s->ne; s->nk = s->ola + s->ne;
    int rc;
    if ((rc = s->part.alloc(kRedBlocks)) || (rc = s->scal.alloc(8)) || (rc = s->Mlinv.alloc(nf + 1)) || (rc = s->gx.alloc(nf + 1)) ||
        (rc = s->tvec.alloc(nf + 1)) || (rc = s->wvec.alloc(nf + 1)))
      return rc;
  }
  if (c->aos.n < (size_t)c->N * c->nE) { int rc = c->aos.alloc((size_t)c->M * c->N * c->nE); if (rc) return rc; }   // layout staging
  *out = (Saddle*)c->saddle;
  return MFB_OK;
}

int need_updated(mfb_ctx* c) {
  if (!c->updated) { g_last_error = "saddle systems: call mfb_update first"; return MFB_ERR_STATE; }
  return MFB_OK;
}

int dot(mfb_ctx* c, Saddle* s, const double* a, const double* b, int64_t n, double* out) {
  k_dot_part<<<kRedBlocks, 256, 0, c->stream>>>(a, b, n, s->part.p);
  k_dot_finish<<<1, 256, 0, c->stream>>>(s->part.p, kRedBlocks, s->scal.p);
  c->launches += 2;
  MFB_CUDA(cudaMemcpyAsync(c->h_scal, s->scal.p, 8, cudaMemcpyDeviceToHost, c->stream));
  MFB_CUDA(cudaStreamSynchronize(c->stream));
  *out = c->h_scal[0];
  return MFB_OK;
}

// lumped = PMP^T 1 ; Minv = 1 / lumped   (linear_system.h:130-134)
int ensure_lumped(mfb_ctx* c, Saddle* s, const NewtonView& nv) {
  if (s->lumped) return MFB_OK;
  const int64_t nf = (int64_t)c->dim * c->nF;
  k_fill<<<nblk(nf), 256, 0, c->stream>>>(s->wvec.p, 1.0, nf);
  MFB_CUDA(cudaMemsetAsync(s->Mlinv.p, 0, (size_t)nf * 8, c->stream));
  KDISPATCH(c, k_mass_free_apply, c->mesh(), nv.emass, s->wvec.p, s->Mlinv.p);
  k_vec_inv<<<nblk(nf), 256, 0, c->stream>>>(s->Mlinv.p, nf);
  c->launches += 2;
  s->lumped = true;
  return MFB_OK;
}

int kkt_apply(mfb_ctx* c, Saddle* s, const NewtonView& nv, const double* v, double* out) {
  const int64_t nf = (int64_t)c->dim * c->nF;
  MFB_CUDA(cudaMemsetAsync(out, 0, (size_t)s->nk * 8, c->stream));     // x part accumulates; the padding of ld must stay 0 (dot products)
  KDISPATCH(c, k_mass_free_apply, c->mesh(), nv.emass, v + s->ox, out + s->ox);
  KDISPATCH(c, k_kkt_elem, c->mesh(), c->state(), v + s->ox, v + s->os, v + s->ola, out + s->ox, out + s->os, out + s->ola);
  (void)nf;
  return MFB_OK;
}

int dual_apply(mfb_ctx* c, Saddle* s, const double* la, double* out) {
  const int64_t nf = (int64_t)c->dim * c->nF;
  MFB_CUDA(cudaMemsetAsync(s->tvec.p, 0, (size_t)nf * 8, c->stream));
  KDISPATCH(c, k_dual_scatter, c->mesh(), c->state(), la, s->tvec.p);
  k_vec_mul<<<nblk(nf), 256, 0, c->stream>>>(s->Mlinv.p, s->tvec.p, 1.0, nf, s->wvec.p);
  c->launches += 1;
  MFB_CUDA(cudaMemsetAsync(out, 0, (size_t)s->ne * 8, c->stream));
  KDISPATCH(c, k_dual_gather, c->mesh(), c->state(), s->wvec.p, la, 1.0, 0.0, 0.0, out);
  return MFB_OK;
}

int alloc_work(Saddle* s, int count, int64_t n) {
  for (int i = 0; i < count; ++i)
    if (s->work[i].n < (size_t)n) { int rc = s->work[i].alloc((size_t)n); if (rc) return rc; }
  return MFB_OK;
}

using ApplyFn = std::function<int(const double*, double*)>;

// Eigen::internal::conjugate_gradient with the identity preconditioner (the recurrence SystemMatrixPD's solver uses,
// restated in oracle/pcg.py): stops on |r|^2 < tol^2 |b|^2
int cg_identity(mfb_ctx* c, Saddle* s, const ApplyFn& A, const double* b, double* x, int64_t n, double tol, int max_iters, int* iters, double* err) {
  int rc;
  if ((rc = alloc_work(s, 3, n))) return rc;
  double *r = s->work[0].p, *p = s->work[1].p, *t = s->work[2].p;
  cudaStream_t st = c->stream;
  double rhs2, r2;
  if ((rc = dot(c, s, b, b, n, &rhs2))) return rc;
  if (rhs2 == 0.0) { MFB_CUDA(cudaMemsetAsync(x, 0, (size_t)n * 8, st)); *iters = 0; *err = 0.0; return MFB_OK; }
  const double thr = std::fmax(tol * tol * rhs2, 2.2250738585072014e-308);
  if ((rc = A(x, t))) return rc;
  k_lin3<<<nblk(n), 256, 0, st>>>(1.0, b, -1.0, t, 0.0, t, r, n);
  c->launches += 1;
  if ((rc = dot(c, s, r, r, n, &r2))) return rc;
  if (r2 < thr) { *iters = 0; *err = std::sqrt(r2 / rhs2); return MFB_OK; }
  MFB_CUDA(cudaMemcpyAsync(p, r, (size_t)n * 8, cudaMemcpyDeviceToDevice, st));
  double absNew = r2;
  int i = 0;
  while (i < max_iters) {
    if ((rc = A(p, t))) return rc;
    double pt;
    if ((rc = dot(c, s, p, t, n, &pt))) return rc;
    const double alpha = absNew / pt;
    k_axpby<<<nblk(n), 256, 0, st>>>(alpha, p, 1.0, x, n);
    k_axpby<<<nblk(n), 256, 0, st>>>(-alpha, t, 1.0, r, n);
    c->launches += 2;
    if ((rc = dot(c, s, r, r, n, &r2))) return rc;
    if (r2 < thr) break;
    const double absOld = absNew;
    absNew = r2;
    k_axpby<<<nblk(n), 256, 0, st>>>(1.0, r, absNew / absOld, p, n);
    c->launches += 1;
    ++i;
  }
  *iters = i; *err = std::sqrt(r2 / rhs2);
  return MFB_OK;
}

// MINRES (Paige & Saunders), identity preconditioner, in the Lanczos form of Eigen's MINRES: the residual norm is
// tracked by the recurrence |r|^2 *= s^2 and the stop test is |r|^2 < tol^2 |b|^2.
int minres_identity(mfb_ctx* c, Saddle* s, const ApplyFn& A, const double* b, double* x, int64_t n, double tol, int max_iters, int* iters, double* err) {
  int rc;
  if ((rc = alloc_work(s, 7, n))) return rc;
  double *v_old = s->work[0].p, *v = s->work[1].p, *v_new = s->work[2].p, *p_oold = s->work[3].p, *p_old = s->work[4].p, *p = s->work[5].p, *t = s->work[6].p;
  cudaStream_t st = c->stream;
  double rhs2;
  if ((rc = dot(c, s, b, b, n, &rhs2))) return rc;
  if (rhs2 == 0.0) { MFB_CUDA(cudaMemsetAsync(x, 0, (size_t)n * 8, st)); *iters = 0; *err = 0.0; return MFB_OK; }
  const double thr2 = tol * tol * rhs2;
  MFB_CUDA(cudaMemsetAsync(v, 0, (size_t)n * 8, st));
  if ((rc = A(x, t))) return rc;
  k_lin3<<<nblk(n), 256, 0, st>>>(1.0, b, -1.0, t, 0.0, t, v_new, n);
  c->launches += 1;
  double r2;
  if ((rc = dot(c, s, v_new, v_new, n, &r2))) return rc;
  double beta_new = std::sqrt(r2);
  const double beta_one = beta_new;
  double cs = 1.0, cs_old = 1.0, sn = 0.0, sn_old = 0.0, eta = 1.0;
  MFB_CUDA(cudaMemsetAsync(p_old, 0, (size_t)n * 8, st));
  MFB_CUDA(cudaMemsetAsync(p_oold, 0, (size_t)n * 8, st));
  MFB_CUDA(cudaMemsetAsync(p, 0, (size_t)n * 8, st));
  int it = 0;
  if (r2 < thr2) { *iters = 0; *err = std::sqrt(r2 / rhs2); return MFB_OK; }
  while (it < max_iters) {
    const double beta = beta_new;
    // v_old = v ; v = v_new / beta   (w == v with the identity preconditioner)
    { double* tmp = v_old; v_old = v; v = v_new; v_new = tmp; }
    k_axpby<<<nblk(n), 256, 0, st>>>(0.0, v, 1.0 / beta, v, n);
    c->launches += 1;
    if ((rc = A(v, t))) return rc;
    double alpha;
    if ((rc = dot(c, s, t, v, n, &alpha))) return rc;          // (A v - beta v_old).v = (A v).v since v_old is orthogonal to v
    k_lin3<<<nblk(n), 256, 0, st>>>(1.0, t, -beta, v_old, -alpha, v, v_new, n);
    c->launches += 1;
    double bn2;
    if ((rc = dot(c, s, v_new, v_new, n, &bn2))) return rc;
    beta_new = std::sqrt(bn2);
    const double r2h = sn * alpha + cs * cs_old * beta;
    const double r3 = sn_old * beta;
    const double r1_hat = cs * alpha - cs_old * sn * beta;
    const double r1 = std::sqrt(r1_hat * r1_hat + bn2);
    cs_old = cs; sn_old = sn;
    cs = r1_hat / r1; sn = beta_new / r1;
    { double* tmp = p_oold; p_oold = p_old; p_old = p; p = tmp; }
    k_lin3<<<nblk(n), 256, 0, st>>>(1.0 / r1, v, -r2h / r1, p_old, -r3 / r1, p_oold, p, n);
    k_axpby<<<nblk(n), 256, 0, st>>>(beta_one * cs * eta, p, 1.0, x, n);
    c->launches += 2;
    r2 *= sn * sn;
    if (r2 < thr2) break;
    eta = -sn * eta;
    ++it;
  }
  *iters = it; *err = std::sqrt(r2 / rhs2);
  return MFB_OK;
}

// host (reference layout) <-> device KKT vector
int kkt_upload(mfb_ctx* c, Saddle* s, const double* host, double* dev) {
  const size_t nf = (size_t)c->dim * c->nF, nel = (size_t)c->N * c->nE;
  cudaStream_t st = c->stream;
  MFB_CUDA(cudaMemsetAsync(dev, 0, (size_t)s->nk * 8, st));
  MFB_CUDA(cudaMemcpyAsync(dev, host, nf * 8, cudaMemcpyHostToDevice, st));
  for (int part = 0; part < 2; ++part) {
    MFB_CUDA(cudaMemcpyAsync(c->aos.p, host + nf + part * nel, nel * 8, cudaMemcpyHostToDevice, st));
    k_aos_to_soa<<<nblk((int64_t)nel), 256, 0, st>>>(c->aos.p, dev + (part ? s->ola : s->os), c->N, c->nE, c->ld);
    c->launches += 1;
  }
  return MFB_OK;
}
int kkt_download(mfb_ctx* c, Saddle* s, const double* dev, double* host) {
  const size_t nf = (size_t)c->dim * c->nF, nel = (size_t)c->N * c->nE;
  cudaStream_t st = c->stream;
  MFB_CUDA(cudaMemcpyAsync(host, dev, nf * 8, cudaMemcpyDeviceToHost, st));
  for (int part = 0; part < 2; ++part) {
    k_soa_to_aos<<<nblk((int64_t)nel), 256, 0, st>>>(dev + (part ? s->ola : s->os), c->aos.p, c->N, c->nE, c->ld);
    c->launches += 1;
    MFB_CUDA(cudaMemcpyAsync(host + nf + part * nel, c->aos.p, nel * 8, cudaMemcpyDeviceToHost, st));
  }
  MFB_CUDA(cudaStreamSynchronize(st));
  return MFB_OK;
}
int elem_upload(mfb_ctx* c, const double* host, double* dev_soa) {
  const size_t nel = (size_t)c->N * c->nE;
  MFB_CUDA(cudaMemsetAsync(dev_soa, 0, (size_t)c->N * c->ld * 8, c->stream));
  MFB_CUDA(cudaMemcpyAsync(c->aos.p, host, nel * 8, cudaMemcpyHostToDevice, c->stream));
  k_aos_to_soa<<<nblk((int64_t)nel), 256, 0, c->stream>>>(c->aos.p, dev_soa, c->N, c->nE, c->ld);
  c->launches += 1;
  return MFB_OK;
}
int elem_download(mfb_ctx* c, const double* dev_soa, double* host) {
  const size_t nel = (size_t)c->N * c->nE;
  k_soa_to_aos<<<nblk((int64_t)nel), 256, 0, c->stream>>>(dev_soa, c->aos.p, c->N, c->nE, c->ld);
  c->launches += 1;
  MFB_CUDA(cudaMemcpyAsync(host, c->aos.p, nel * 8, cudaMemcpyDeviceToHost, c->stream));
  MFB_CUDA(cudaStreamSynchronize(c->stream));
  return MFB_OK;
}

// b = [-Displacement::gradient ; -gradient_mixed ; -gradient_dual]   (linear_system.h:71-87)
int kkt_rhs_dev(mfb_ctx* c, Saddle* s, double* b) {
  int rc;
  MFB_CUDA(cudaMemsetAsync(b, 0, (size_t)s->nk * 8, c->stream));
  if ((rc = newton_disp_gradient(c, s->gx.p))) return rc;
  const int64_t nf = (int64_t)c->dim * c->nF, nel = (int64_t)c->N * c->nE;
  k_axpby<<<nblk(nf), 256, 0, c->stream>>>(-1.0, s->gx.p, 0.0, b + s->ox, nf);
  k_elem_scale_vol<<<nblk(nel), 256, 0, c->stream>>>(c->vol.p, c->g.p, -1.0, c->N, c->nE, c->ld, b + s->os);
  k_elem_scale<<<nblk(nel), 256, 0, c->stream>>>(c->grad_la.p, -1.0, c->N, c->nE, c->ld, b + s->ola);
  c->launches += 3;
  return MFB_OK;
}

int dual_rhs_dev(mfb_ctx* c, Saddle* s, double* b) {
  int rc;
  const int64_t nf = (int64_t)c->dim * c->nF;
  if ((rc = newton_disp_gradient(c, s->gx.p))) return rc;
  k_vec_mul<<<nblk(nf), 256, 0, c->stream>>>(s->Mlinv.p, s->gx.p, -1.0, nf, s->wvec.p);
  c->launches += 1;
  MFB_CUDA(cudaMemsetAsync(b, 0, (size_t)s->ne * 8, c->stream));
  KDISPATCH(c, k_dual_gather, c->mesh(), c->state(), s->wvec.p, (const double*)nullptr, 0.0, -1.0, 1.0, b);
  return MFB_OK;
}

// la_ = la ; dx = Ml^-1 (-gx - B la) ; ds = Hinv (-gradient_mixed - C la)   (linear_system.h:178-209)
int dual_post_solve_dev(mfb_ctx* c, Saddle* s, const NewtonView& nv, const double* la) {
  int rc;
  const int64_t nf = (int64_t)c->dim * c->nF;
  MFB_CUDA(cudaMemcpyAsync(c->la.p, la, (size_t)s->ne * 8, cudaMemcpyDeviceToDevice, c->stream));
  if ((rc = newton_disp_gradient(c, s->gx.p))) return rc;
  MFB_CUDA(cudaMemsetAsync(s->tvec.p, 0, (size_t)nf * 8, c->stream));
  KDISPATCH(c, k_dual_scatter, c->mesh(), c->state(), la, s->tvec.p);
  k_axpby<<<nblk(nf), 256, 0, c->stream>>>(-1.0, s->gx.p, -1.0, s->tvec.p, nf);        // -gx - B la
  k_vec_mul<<<nblk(nf), 256, 0, c->stream>>>(s->Mlinv.p, s->tvec.p, 1.0, nf, nv.dx);
  c->launches += 2;
  KDISPATCH(c, k_dual_ds, c->mesh(), c->state(), la, c->ds.p);
  return MFB_OK;
}

}  // namespace

namespace mfb {
void saddle_free(mfb_ctx* c) {
  if (c && c->saddle) { delete (Saddle*)c->saddle; c->saddle = nullptr; }
}
}  // namespace mfb

extern "C" int64_t mfb_kkt_size(mfb_ctx* c) { return c ? (int64_t)c->dim * c->nF + 2 * (int64_t)c->N * c->nE : 0; }

extern "C" int mfb_kkt_layout_dev(mfb_ctx* c, int64_t* off_x, int64_t* off_s, int64_t* off_la, int64_t* total) {
  CHECK_CTX(c);
  Saddle* s; int rc;
  if ((rc = saddle_get(c, &s))) return rc;
  if (off_x) *off_x = s->ox; if (off_s) *off_s = s->os; if (off_la) *off_la = s->ola; if (total) *total = s->nk;
  return MFB_OK;
}

extern "C" int mfb_kkt_product_dev(mfb_ctx* c, const double* v_dev, double* out_dev) {
  CHECK_CTX(c);
  Saddle* s; NewtonView nv; int rc;
  if (!v_dev || !out_dev) return MFB_ERR_INVALID;
  if ((rc = need_updated(c)) || (rc = saddle_get(c, &s)) || (rc = newton_view(c, &nv))) return rc;
  if ((rc = kkt_apply(c, s, nv, v_dev, out_dev))) return rc;
  MFB_CUDA(cudaGetLastError());
  return MFB_OK;
}

extern "C" int mfb_kkt_product(mfb_ctx* c, const double* v, double* out) {
  CHECK_CTX(c);
  Saddle* s; NewtonView nv; int rc;
  if (!v || !out) return MFB_ERR_INVALID;
  if ((rc = need_updated(c)) || (rc = saddle_get(c, &s)) || (rc = newton_view(c, &nv)) || (rc = alloc_work(s, 8, s->nk))) return rc;
  if ((rc = kkt_upload(c, s, v, s->work[6].p)) || (rc = kkt_apply(c, s, nv, s->work[6].p, s->work[7].p))) return rc;
  MFB_CUDA(cudaGetLastError());
  return kkt_download(c, s, s->work[7].p, out);
}

extern "C" int mfb_kkt_rhs(mfb_ctx* c, double* out) {
  CHECK_CTX(c);
  Saddle* s; int rc;
  if (!out) return MFB_ERR_INVALID;
  if ((rc = need_updated(c)) || (rc = saddle_get(c, &s)) || (rc = alloc_work(s, 8, s->nk))) return rc;
  if ((rc = kkt_rhs_dev(c, s, s->work[7].p))) return rc;
  MFB_CUDA(cudaGetLastError());
  return kkt_download(c, s, s->work[7].p, out);
}

// pre_solve + MINRES from the given guess + post_solve (x.delta() = dx, var.delta() = ds, var.lambda() = la).
// sol: host, reference layout, in = initial guess, out = solution.
extern "C" int mfb_kkt_solve_minres(mfb_ctx* c, double* sol, double tol, int max_iters, int* iters_out, double* relres_out) {
  CHECK_CTX(c);
  Saddle* s; NewtonView nv; int rc;
  if (!sol) return MFB_ERR_INVALID;
  if ((rc = need_updated(c)) || (rc = saddle_get(c, &s)) || (rc = newton_view(c, &nv))) return rc;
  DBuf<double> b, x;
  if ((rc = b.alloc((size_t)s->nk)) || (rc = x.alloc((size_t)s->nk))) return rc;
  if ((rc = kkt_rhs_dev(c, s, b.p)) || (rc = kkt_upload(c, s, sol, x.p))) return rc;
  int it = 0; double err = 0;
  ApplyFn A = [&](const double* in, double* out) { return kkt_apply(c, s, nv, in, out); };
  if ((rc = minres_identity(c, s, A, b.p, x.p, s->nk, tol, max_iters, &it, &err))) return rc;
  // SystemMatrixIndefinite::post_solve (linear_system.h:89-100)
  const size_t nf = (size_t)c->dim * c->nF;
  MFB_CUDA(cudaMemcpyAsync(nv.dx, x.p + s->ox, nf * 8, cudaMemcpyDeviceToDevice, c->stream));
  MFB_CUDA(cudaMemcpyAsync(c->ds.p, x.p + s->os, (size_t)s->ne * 8, cudaMemcpyDeviceToDevice, c->stream));
  MFB_CUDA(cudaMemcpyAsync(c->la.p, x.p + s->ola, (size_t)s->ne * 8, cudaMemcpyDeviceToDevice, c->stream));
  if (iters_out) *iters_out = it;
  if (relres_out) *relres_out = err;
  MFB_CUDA(cudaGetLastError());
  return kkt_download(c, s, x.p, sol);
}

extern "C" int mfb_dual_product(mfb_ctx* c, const double* la, double* out) {
  CHECK_CTX(c);
  Saddle* s; NewtonView nv; int rc;
  if (!la || !out) return MFB_ERR_INVALID;
  if ((rc = need_updated(c)) || (rc = saddle_get(c, &s)) || (rc = newton_view(c, &nv)) || (rc = ensure_lumped(c, s, nv)) || (rc = alloc_work(s, 8, s->ne))) return rc;
  if ((rc = elem_upload(c, la, s->work[6].p))) return rc;
  MFB_CUDA(cudaMemsetAsync(s->work[7].p, 0, (size_t)s->ne * 8, c->stream));
  if ((rc = dual_apply(c, s, s->work[6].p, s->work[7].p))) return rc;
  MFB_CUDA(cudaGetLastError());
  return elem_download(c, s->work[7].p, out);
}

extern "C" int mfb_dual_product_dev(mfb_ctx* c, const double* la_dev_soa, double* out_dev_soa) {
  CHECK_CTX(c);
  Saddle* s; NewtonView nv; int rc;
  if (!la_dev_soa || !out_dev_soa) return MFB_ERR_INVALID;
  if ((rc = need_updated(c)) || (rc = saddle_get(c, &s)) || (rc = newton_view(c, &nv)) || (rc = ensure_lumped(c, s, nv))) return rc;
  if ((rc = dual_apply(c, s, la_dev_soa, out_dev_soa))) return rc;
  MFB_CUDA(cudaGetLastError());
  return MFB_OK;
}

extern "C" int mfb_dual_rhs(mfb_ctx* c, double* out) {
  CHECK_CTX(c);
  Saddle* s; NewtonView nv; int rc;
  if (!out) return MFB_ERR_INVALID;
  if ((rc = need_updated(c)) || (rc = saddle_get(c, &s)) || (rc = newton_view(c, &nv)) || (rc = ensure_lumped(c, s, nv)) || (rc = alloc_work(s, 8, s->ne))) return rc;
  if ((rc = dual_rhs_dev(c, s, s->work[7].p))) return rc;
  MFB_CUDA(cudaGetLastError());
  return elem_download(c, s->work[7].p, out);
}

// DualCondensedSystem: pre_solve, CG on (G^T G + C C Hinv) la = rhs from the guess in la (host, reference layout),
// post_solve (lambda() = la, x.delta() = Ml^-1(-gx - B la), delta() = Hinv(-gradient_mixed - C la)).
// dx_out (d nF) / ds_out (N nE): host or NULL.
extern "C" int mfb_dual_solve_cg(mfb_ctx* c, double* la, double tol, int max_iters, int* iters_out, double* relres_out, double* dx_out, double* ds_out) {
  CHECK_CTX(c);
  Saddle* s; NewtonView nv; int rc;
  if (!la) return MFB_ERR_INVALID;
  if ((rc = need_updated(c)) || (rc = saddle_get(c, &s)) || (rc = newton_view(c, &nv)) || (rc = ensure_lumped(c, s, nv))) return rc;
  DBuf<double> b, x;
  if ((rc = b.alloc((size_t)s->ne)) || (rc = x.alloc((size_t)s->ne))) return rc;
  if ((rc = dual_rhs_dev(c, s, b.p)) || (rc = elem_upload(c, la, x.p))) return rc;
  int it = 0; double err = 0;
  ApplyFn A = [&](const double* in, double* out) { return dual_apply(c, s, in, out); };
  if ((rc = cg_identity(c, s, A, b.p, x.p, s->ne, tol, max_iters, &it, &err))) return rc;
  if ((rc = dual_post_solve_dev(c, s, nv, x.p))) return rc;
  if (iters_out) *iters_out = it;
  if (relres_out) *relres_out = err;
  MFB_CUDA(cudaGetLastError());
  if ((rc = elem_download(c, x.p, la))) return rc;
  if (dx_out) {
    MFB_CUDA(cudaMemcpyAsync(dx_out, nv.dx, (size_t)c->dim * c->nF * 8, cudaMemcpyDeviceToHost, c->stream));
    MFB_CUDA(cudaStreamSynchronize(c->stream));
  }
  if (ds_out && (rc = elem_download(c, c->ds.p, ds_out))) return rc;
  return MFB_OK;
}
