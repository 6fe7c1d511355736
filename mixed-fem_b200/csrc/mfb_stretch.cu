// The NON-mixed variable of the reference, for the FEM-vs-MFEM comparisons of apps/convergence.cpp:
//
//   Stretch<DIM>::energy / update / rhs / gradient ... src/variables/stretch.cpp:16-68
//   F-form material virtuals ......................... src/energies/material_model.h:43-62
//   StableNeohookean energy/gradient/hessian(Vector9d)  src/energies/stable_neohookean.cpp:93-232
//   sim::simple_psd_fix(H, 0.0) ....................... deps/Bartels/src/simple_psd_fix.cpp:5-22
//   Assembler / VecAssembler .......................... src/utils/sparse_utils.h:11-64, 66-120 (same pattern and scatter map as
//                                                       the mixed variable: both are built from T_ and free_map_)
//
// Per tet (one thread):  F = Ds Dm^-1;  g = dpsi/dF (9);  H = d2psi/dF2 (9x9), eigenvalues below 0 clamped to 0;
//   g_e = Jloc^T g vol h^2 (12);  H_e = Jloc^T H Jloc vol h^2 (12x12), Jloc = dphi^T (x) I_3;  both reduced into the
//   global vector / the reference-layout CSR through the element -> nonzero map.
//
// The reference's F-form code for Stable Neo-Hookean is symbolic-toolbox output (81 expanded polynomials); it is the
// expansion of   psi = mu/2 (|F|^2 - 3) - mu (J - 1) + la/2 (J - 1)^2,
//   g = mu F + (la (J - 1) - mu) cof F,   H = mu I + la vec(cof F) vec(cof F)^T + (la (J - 1) - mu) d(cof F)/dF,
// which is what is evaluated here (tests/test_oracle_stretch.py checks the closed form against finite differences of
// the energy expression of stable_neohookean.cpp:93-107).  Only Stable Neo-Hookean is implemented, for tets (config C2)
// and 2D triangles (config C1's material; stable_neohookean.cpp:274-324 is the same closed form with a 2x2 cofactor);
// ARAP (tets) is further down; other models return MFB_ERR_UNSUPPORTED.  This is a coverage path, not a tuned one: full (d(d+1))^2 scattered FP64
// reductions, the d^2 x d^2 Jacobi eigen-solve in local memory.
#include "mfb_ctx.cuh"
#include <cmath>

using namespace mfb;

#define CHECK_CTX(c) do { if (!(c)) { g_last_error = "null context"; return MFB_ERR_INVALID; } cudaSetDevice((c)->device); } while (0)

namespace {

inline unsigned nblk(int64_t n, int t = 256) { return (unsigned)((n + t - 1) / t); }

// cof F = dJ/dF, column-major [3c + r]
__device__ __forceinline__ void cofactor3(const double* F, double* C) {
  C[0] = F[4] * F[8] - F[7] * F[5]; C[1] = F[5] * F[6] - F[8] * F[3]; C[2] = F[3] * F[7] - F[6] * F[4];
  C[3] = F[7] * F[2] - F[1] * F[8]; C[4] = F[8] * F[0] - F[2] * F[6]; C[5] = F[6] * F[1] - F[0] * F[7];
  C[6] = F[1] * F[5] - F[4] * F[2]; C[7] = F[2] * F[3] - F[5] * F[0]; C[8] = F[0] * F[4] - F[3] * F[1];
}

// d2J / dF(r,c) dF(s,c2) = sum_{m,n} eps(r,s,m) eps(c,c2,n) F(m,n): non-zero only for r != s and c != c2
__device__ __forceinline__ double d2J(const double* F, int r, int c, int s, int c2) {
  if (r == s || c == c2) return 0.0;
  const int m = 3 - r - s, n = 3 - c - c2;
  const double sr = ((s - r + 3) % 3 == 1) ? 1.0 : -1.0;       // eps(r,s,m)
  const double sc = ((c2 - c + 3) % 3 == 1) ? 1.0 : -1.0;      // eps(c,c2,n)
  return sr * sc * F[3 * n + m];
}

// 2D twins: cof F = [F11 -F10; -F01 F00] (column-major [2c + r]), d2J = eps(r,s) eps(c,c2)
__device__ __forceinline__ void cofactor2(const double* F, double* C) { C[0] = F[3]; C[1] = -F[2]; C[2] = -F[1]; C[3] = F[0]; }
__device__ __forceinline__ double d2J2(int r, int c, int s, int c2) {
  if (r == s || c == c2) return 0.0;
  return ((s > r) ? 1.0 : -1.0) * ((c2 > c) ? 1.0 : -1.0);
}

// psi, J-dependent coefficient and cofactor of the Stable Neo-Hookean F-form for DIM = 2, 3
// (stable_neohookean.cpp:93-107 / :274-282: mu (1 - J) + la/2 (1 - J)^2 + mu/2 (|F|^2 - DIM))
template <int DIM>
__device__ __forceinline__ double snh_J_cof(const double* F, double* C) {
  if constexpr (DIM == 3) { cofactor3(F, C); return F[0] * C[0] + F[1] * C[1] + F[2] * C[2]; }
  else { cofactor2(F, C); return F[0] * F[3] - F[2] * F[1]; }
}

// ARAP F-form (src/energies/arap.cpp:134-162): psi = mu/2 (|F|^2 - 2 tr S + 3), g = mu (F - R), H = mu (I - dR/dF), NO psd fix.
// The reference takes R and dR/dF from iARAP::rotation (closed-form polar, src/svd/iARAP.cpp); the polar factor and its
// derivative are unique functions of F, here: the Newton polar of the mixed path started from the identity and polished
// (polar_from_identity), and  dR/dF(:, (r,c)) = [w x r_b]_b  with  w = (tr(Y) I - Y)^-1 (r_c x e_r),  Y = R F^T,  r_b = columns of R
// (differentiate F = R S, take the skew part of dF R^T).  Returns false if the polar iteration failed.
// Polar rotation of F from scratch: the mixed path's Newton iteration (reference tolerance 1e-6, with its back-tracking)
// from the identity, then two undamped Newton steps (quadratic: 1e-6 -> 1e-12 -> rounding).  Running the reference
// iteration itself to 1e-13 stalls on some elements: its back-tracking compares against -tr(R Y) (sic) and shrinks the
// step below the tolerance.  Y = R F^T on exit.
__device__ __forceinline__ int polar_from_identity(const double* F, double* R, double* Y) {
  int status = 0;
#pragma unroll
  for (int i = 0; i < 9; ++i) R[i] = (i % 4 == 0) ? 1.0 : 0.0;
  procrustes3(R, F, Y, 1e-6, 100, &status);
  for (int k = 0; k < 2; ++k) {
    const double gv[3] = {Y[5] - Y[7], Y[6] - Y[2], Y[1] - Y[3]};
    double inv[6], w[3], dR[9], t[9];
    syminv3(Y[4] + Y[8], -0.5 * (Y[3] + Y[1]), -0.5 * (Y[6] + Y[2]), Y[0] + Y[8], -0.5 * (Y[7] + Y[5]), Y[0] + Y[4], inv);
    symmul3(inv, gv, w);
    w[0] = -w[0]; w[1] = -w[1]; w[2] = -w[2];
    // exp([w]x) to second order (|w| <= 1e-6 here: error 1e-19); rodrigues3 keeps the reference's cut-off and returns
    // the identity below an angle of 1e-8, which would leave up to 1e-8 of rotation uncorrected
    dR[0] = 1.0 - 0.5 * (w[1] * w[1] + w[2] * w[2]); dR[4] = 1.0 - 0.5 * (w[0] * w[0] + w[2] * w[2]); dR[8] = 1.0 - 0.5 * (w[0] * w[0] + w[1] * w[1]);
    dR[1] = w[2] + 0.5 * w[0] * w[1];  dR[3] = -w[2] + 0.5 * w[0] * w[1];
    dR[2] = -w[1] + 0.5 * w[0] * w[2]; dR[6] = w[1] + 0.5 * w[0] * w[2];
    dR[5] = w[0] + 0.5 * w[1] * w[2];  dR[7] = -w[0] + 0.5 * w[1] * w[2];
    mul33(dR, R, t);
#pragma unroll
    for (int i = 0; i < 9; ++i) R[i] = t[i];
    mul33(dR, Y, t);
#pragma unroll
    for (int i = 0; i < 9; ++i) Y[i] = t[i];
  }
  return status;
}

__device__ __forceinline__ bool arap_f_form(const double* F, double mu, double* R, double* g, double* H /* packed lower 45 */) {
  double Y[9], Hinv3[6];
  const int status = polar_from_identity(F, R, Y);
  const bool ok = procrustes_hinv(Y, Hinv3) && !(status & ST_POLAR_MAXITER);
#pragma unroll
  for (int i = 0; i < 9; ++i) g[i] = mu * (F[i] - R[i]);
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int jc = 3 * c + r;
      double er[3] = {0.0, 0.0, 0.0};
      er[r] = 1.0;
      double a[3], w[3];
      cross3(&R[3 * c], er, a);
      symmul3(Hinv3, a, w);
#pragma unroll
      for (int b = 0; b < 3; ++b) {
        double v[3];
        cross3(w, &R[3 * b], v);
#pragma unroll
        for (int q = 0; q < 3; ++q) {
          const int i = 3 * b + q;
          if (i >= jc) H[symidx(i, jc)] = mu * ((i == jc ? 1.0 : 0.0) - v[q]);
        }
      }
    }
  return ok;
}

// 2D ARAP part (arap.cpp:96-132, fixed_corotational.cpp:153-195): R = U V^T with the flip fix of mfem::svd<double,2>, and
// dR/dF = 2/(s0 + s1) q0 q0^T, q0 = vec(U [0 -1; 1 0] V^T)/sqrt 2.  In closed form: R = Rot(atan2(c - b, a + d)),
// s0 + s1 = 2 Q = |(a + d, c - b)|, U Sk V^T = +-R Sk, so dR/dF = vec(R Sk) vec(R Sk)^T / (2 Q), vec(R Sk) = (-sr, cr, -cr, -sr).
__device__ __forceinline__ void arap2_parts(const double* F, double* R, double* q, double* inv_sum) {
  const double E = 0.5 * (F[0] + F[3]), Hh = 0.5 * (F[1] - F[2]);
  const double Q = sqrt(E * E + Hh * Hh);
  const double cr = Q > 0.0 ? E / Q : 1.0, sr = Q > 0.0 ? Hh / Q : 0.0;
  R[0] = cr; R[1] = sr; R[2] = -sr; R[3] = cr;
  q[0] = -sr; q[1] = cr; q[2] = -cr; q[3] = -sr;
  *inv_sum = Q > 0.0 ? 1.0 / (2.0 * Q) : 0.0;
}

// psd-fix tolerance of each F-form Hessian: Stable-Neohookean simple_psd_fix(H, 0.0) (stable_neohookean.cpp:228,322),
// Neohookean / FixedCorot simple_psd_fix(H) with the default 1e-6 (neohookean.cpp:227, fixed_corotational.cpp:210,358),
// ARAP none (arap.cpp:112-162): negative = no fix
__device__ __forceinline__ double fform_fix_tol(int model) { return model == 0 ? 0.0 : (model == 4 ? -1.0 : 1e-6); }

template <int DIM>
__global__ void __launch_bounds__(128) k_stretch_update(DevMesh m, const double* __restrict__ x, double h2, double* __restrict__ Avals,
                                                        double* __restrict__ grad, unsigned long long* __restrict__ counters) {
  constexpr int M = DIM * DIM, NPE = DIM + 1;
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= m.nE) return;
  const int64_t ld = m.ld;
  double dphi[M], F[M], C[M];
  elem_F<DIM>(m, e, x, dphi, F);
  const double J = snh_J_cof<DIM>(F, C);
  const MaterialDev mat = m.matid ? m.mats[m.matid[e]] : m.mat0;
  const double mu = mat.mu, la = mat.la;
  const double k1 = la * (J - 1.0) - mu;
  const double scale = m.vol[e] * h2;
  double g9[M], H[M * (M + 1) / 2];
  const int model = mat.model;
  // ---- the ARAP part of ARAP / FixedCorot: g = c (F - R), H = c (I - dR/dF); c = mu, except FixedCorot 2D: 2 mu
  if (model == 4 || model == 3) {
    if constexpr (DIM == 3) {
      double R[9];
      if (!arap_f_form(F, mu, R, g9, H)) atomicAdd(&counters[1], 1ull);     // polar_hfail
    } else {
      const double cm = (model == 3) ? 2.0 * mu : mu;
      double R[4], q[4], is;
      arap2_parts(F, R, q, &is);
#pragma unroll
      for (int i = 0; i < 4; ++i) g9[i] = cm * (F[i] - R[i]);
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j <= i; ++j) H[symidx(i, j)] = cm * ((i == j ? 1.0 : 0.0) - is * q[i] * q[j]);
    }
  } else {
#pragma unroll
    for (int i = 0; i < M; ++i) g9[i] = 0.0;
#pragma unroll
    for (int i = 0; i < M * (M + 1) / 2; ++i) H[i] = 0.0;
  }
  // ---- the part that goes through J:  g += a0 F + a1 cof F ;  H += a0 I + a2 cof cof^T + a1 d2J/dF2
  //   Stable-Neohookean  a0 = mu, a1 = la (J - 1) - mu,        a2 = la                        (stable_neohookean.cpp:109-226)
  //   Neohookean         a0 = mu, a1 = (la log J - mu) / J,     a2 = (la (1 - log J) + mu) / J^2  (neohookean.cpp:108-226)
  //   FixedCorot         a0 = 0,  a1 = la (J - 1),              a2 = la                        (fixed_corotational.cpp:233-357)
  if (model != 4) {
    double a0 = 0.0, a1, a2;
    if (model == 0) { a0 = mu; a1 = k1; a2 = la; }
    else if (model == 1) { const double lj = log(J), ij = 1.0 / J; a0 = mu; a1 = (la * lj - mu) * ij; a2 = (la * (1.0 - lj) + mu) * ij * ij; }
    else { a1 = la * (J - 1.0); a2 = la; }
#pragma unroll
    for (int i = 0; i < M; ++i) g9[i] += a0 * F[i] + a1 * C[i];
#pragma unroll
    for (int i = 0; i < M; ++i)
#pragma unroll
      for (int j = 0; j <= i; ++j) {
        double dd;
        if constexpr (DIM == 3) dd = d2J(F, i % 3, i / 3, j % 3, j / 3);
        else dd = d2J2(i % 2, i / 2, j % 2, j / 2);
        H[symidx(i, j)] += (i == j ? a0 : 0.0) + a2 * C[i] * C[j] + a1 * dd;
      }
  }
  // ---- sim::simple_psd_fix(H, tol): eigenvalues below tol are set to tol
  const double ftol = fform_fix_tol(model);
  if (ftol >= 0.0) {
    double V[M * M], ev[M];
    jacobi_eig<M>(H, V);
#pragma unroll
    for (int k = 0; k < M; ++k) ev[k] = H[symidx(k, k)] < ftol ? ftol : H[symidx(k, k)];
#pragma unroll
    for (int i = 0; i < M; ++i)
#pragma unroll
      for (int j = 0; j <= i; ++j) {
        double a = 0.0;
#pragma unroll
        for (int q = 0; q < M; ++q) a += V[q * M + i] * V[q * M + j] * ev[q];
        H[symidx(i, j)] = a;
      }
  }
  // gradient: Jloc^T g
  scatter_vertex<DIM>(m, e, dphi, g9, scale, grad);
  // column (k,s) of Hfix Jloc:  col[i] = sum_c2 dphi(k,c2) Hfix[i, DIM c2 + s] ;  K(j,r ; k,s) = sum_c dphi(j,c) col[DIM c + r]
  int sidx[NPE * NPE], stride[NPE];
#pragma unroll
  for (int q = 0; q < NPE * NPE; ++q) sidx[q] = m.smap[q * ld + e];
#pragma unroll
  for (int k = 0; k < NPE; ++k) stride[k] = m.sstride[k * ld + e];
  double d4[NPE][DIM];
#pragma unroll
  for (int c = 0; c < DIM; ++c) {
    double sum = 0.0;
#pragma unroll
    for (int k = 0; k < DIM; ++k) { d4[k + 1][c] = dphi[DIM * k + c]; sum += dphi[DIM * k + c]; }
    d4[0][c] = -sum;
  }
#pragma unroll
  for (int k = 0; k < NPE; ++k) {
#pragma unroll
    for (int s = 0; s < DIM; ++s) {
      double col[M];
#pragma unroll
      for (int i = 0; i < M; ++i) {
        double a = 0.0;
#pragma unroll
        for (int c2 = 0; c2 < DIM; ++c2) a += d4[k][c2] * H[symidx(i, DIM * c2 + s)];
        col[i] = a;
      }
#pragma unroll
      for (int j = 0; j < NPE; ++j) {
        const int base = sidx[j * NPE + k];
        if (base < 0) continue;
#pragma unroll
        for (int r = 0; r < DIM; ++r) {
          double v = 0.0;
#pragma unroll
          for (int c = 0; c < DIM; ++c) v += d4[j][c] * col[DIM * c + r];
          red_add(&Avals[(int64_t)base + (int64_t)r * stride[j] + s], scale * v);
        }
      }
    }
  }
}

template <int DIM>
__global__ void __launch_bounds__(128) k_stretch_energy(DevMesh m, const double* __restrict__ x, double* __restrict__ partials) {
  constexpr int M = DIM * DIM;
  __shared__ double sh[4];
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  double val = 0.0;
  if (e < m.nE) {
    double dphi[M], F[M], C[M];
    elem_F<DIM>(m, e, x, dphi, F);
    const double J = snh_J_cof<DIM>(F, C);
    const MaterialDev mat = m.matid ? m.mats[m.matid[e]] : m.mat0;
    double i1 = 0.0;
#pragma unroll
    for (int i = 0; i < M; ++i) i1 += F[i] * F[i];
    double psi = mat.mu * (1.0 - J) + 0.5 * mat.la * (1.0 - J) * (1.0 - J) + 0.5 * mat.mu * (i1 - (double)DIM);      // Stable-Neohookean
    if (mat.model == 1) { const double lj = log(J); psi = 0.5 * mat.mu * (i1 - 3.0) - mat.mu * lj + 0.5 * mat.la * lj * lj; }   // neohookean.cpp:93-106
    if (mat.model == 3 || mat.model == 4) {
      if constexpr (DIM == 3) {      // ARAP and (sic, no volume term) FixedCorot: mu/2 (|F|^2 - 2 tr S + 3), tr S = R : F (arap.cpp:134-141,
        double R[9], Y[9];           // fixed_corotational.cpp:215-223)
        polar_from_identity(F, R, Y);
        double trS = 0.0;
#pragma unroll
        for (int i = 0; i < 9; ++i) trS += R[i] * F[i];
        psi = 0.5 * mat.mu * (i1 - 2.0 * trS + 3.0);
      } else {                       // 2D: ARAP mu/2 |F - R|^2 (arap.cpp:82-94) ; FixedCorot mu |F - R|^2 + la/2 (J - 1)^2 (fixed_corotational.cpp:136-151)
        double R[4], q[4], is, d2 = 0.0;
        arap2_parts(F, R, q, &is);
#pragma unroll
        for (int i = 0; i < 4; ++i) d2 += (F[i] - R[i]) * (F[i] - R[i]);
        psi = (mat.model == 4) ? 0.5 * mat.mu * d2 : mat.mu * d2 + 0.5 * mat.la * (J - 1.0) * (J - 1.0);
      }
    }
    val = psi * m.vol[e];
  }
  const double r = block_sum_128(val, sh);
  if (threadIdx.x == 0) partials[blockIdx.x] = r;
}

struct StretchState {
  DBuf<double> Avals, grad;
  bool updated = false;
};

int stretch_get(mfb_ctx* c, StretchState** out) {
  if (c->kind == MFB_SHELL) { g_last_error = "Stretch variable: tetrahedra and 2D triangles only"; return MFB_ERR_UNSUPPORTED; }
  // every material of the mesh needs an F-form overload in the reference: material_model.h:43-62 and the per-model files
  for (size_t i = 0; i < c->h_mat_model.size(); ++i) {
    const int mdl = c->h_mat_model[i];
    if (!(mdl == 0 || mdl == 3 || mdl == 4 || (mdl == 1 && c->dim == 3))) {
      g_last_error = "Stretch variable: the reference has F-forms for Stable-Neohookean, ARAP, FixedCorot (2D, 3D) and Neohookean (3D) only "
                     "(Corot has none, Fung's returns zeros)";
      return MFB_ERR_UNSUPPORTED;
    }
  }
  if (!c->stretch) {
    StretchState* s = new StretchState();
    c->stretch = s;
    int rc;
    if ((rc = s->Avals.alloc(c->nnz > 0 ? (size_t)c->nnz : 1)) || (rc = s->grad.alloc((size_t)c->dim * c->nF + 1))) return rc;
  }
  *out = (StretchState*)c->stretch;
  return MFB_OK;
}

}  // namespace

namespace mfb {
void stretch_free(mfb_ctx* c) {
  if (c && c->stretch) { delete (StretchState*)c->stretch; c->stretch = nullptr; }
}
}  // namespace mfb

extern "C" int mfb_stretch_update_dev(mfb_ctx* c, const double* x_dev, double h) {
  CHECK_CTX(c);
  StretchState* s; int rc;
  if (!x_dev) return MFB_ERR_INVALID;
  if ((rc = stretch_get(c, &s))) return rc;
  cudaStream_t st = c->stream;
  MFB_CUDA(cudaMemsetAsync(s->Avals.p, 0, (size_t)(c->nnz > 0 ? c->nnz : 1) * 8, st));
  MFB_CUDA(cudaMemsetAsync(s->grad.p, 0, (size_t)c->dim * c->nF * 8, st));
  MFB_CUDA(cudaMemsetAsync(c->counters.p, 0, 64, st));
  if (c->dim == 3) k_stretch_update<3><<<c->grid(), 128, 0, st>>>(c->mesh(), x_dev, h * h, s->Avals.p, s->grad.p, c->counters.p);
  else k_stretch_update<2><<<c->grid(), 128, 0, st>>>(c->mesh(), x_dev, h * h, s->Avals.p, s->grad.p, c->counters.p);
  c->launches += 1;
  MFB_CUDA(cudaGetLastError());
  s->updated = true;
  return MFB_OK;
}

extern "C" int mfb_stretch_update(mfb_ctx* c, const double* x, double h) {
  CHECK_CTX(c);
  if (!x) return MFB_ERR_INVALID;
  MFB_CUDA(cudaMemcpyAsync(c->xdev.p, x, (size_t)c->dim * c->nV * 8, cudaMemcpyHostToDevice, c->stream));
  return mfb_stretch_update_dev(c, c->xdev.p, h);
}

extern "C" int mfb_stretch_lhs_values(mfb_ctx* c, double* values) {
  CHECK_CTX(c);
  StretchState* s; int rc;
  if (!values) return MFB_ERR_INVALID;
  if ((rc = stretch_get(c, &s))) return rc;
  if (!s->updated) { g_last_error = "mfb_stretch_lhs_values: call mfb_stretch_update first"; return MFB_ERR_STATE; }
  MFB_CUDA(cudaMemcpyAsync(values, s->Avals.p, (size_t)c->nnz * 8, cudaMemcpyDeviceToHost, c->stream));
  MFB_CUDA(cudaStreamSynchronize(c->stream));
  return MFB_OK;
}

extern "C" int mfb_stretch_dev(mfb_ctx* c, const double** lhs_values_dev, const double** gradient_dev) {
  CHECK_CTX(c);
  StretchState* s; int rc;
  if ((rc = stretch_get(c, &s))) return rc;
  if (!s->updated) { g_last_error = "mfb_stretch_dev: call mfb_stretch_update first"; return MFB_ERR_STATE; }
  if (lhs_values_dev) *lhs_values_dev = s->Avals.p;
  if (gradient_dev) *gradient_dev = s->grad.p;
  return MFB_OK;
}

extern "C" int mfb_stretch_gradient(mfb_ctx* c, double* out) {
  CHECK_CTX(c);
  StretchState* s; int rc;
  if (!out) return MFB_ERR_INVALID;
  if ((rc = stretch_get(c, &s))) return rc;
  if (!s->updated) { g_last_error = "mfb_stretch_gradient: call mfb_stretch_update first"; return MFB_ERR_STATE; }
  MFB_CUDA(cudaMemcpyAsync(out, s->grad.p, (size_t)c->dim * c->nF * 8, cudaMemcpyDeviceToHost, c->stream));
  MFB_CUDA(cudaStreamSynchronize(c->stream));
  return MFB_OK;
}

extern "C" int mfb_stretch_energy(mfb_ctx* c, const double* x, double* e) {
  CHECK_CTX(c);
  StretchState* s; int rc;
  if (!x || !e) return MFB_ERR_INVALID;
  if ((rc = stretch_get(c, &s))) return rc;
  cudaStream_t st = c->stream;
  MFB_CUDA(cudaMemcpyAsync(c->xdev.p, x, (size_t)c->dim * c->nV * 8, cudaMemcpyHostToDevice, st));
  if (c->dim == 3) k_stretch_energy<3><<<c->grid(), 128, 0, st>>>(c->mesh(), c->xdev.p, c->partials.p);
  else k_stretch_energy<2><<<c->grid(), 128, 0, st>>>(c->mesh(), c->xdev.p, c->partials.p);
  k_sum_partials<<<1, 256, 0, st>>>(c->partials.p, c->grid(), c->scal.p);
  c->launches += 2;
  MFB_CUDA(cudaMemcpyAsync(c->h_scal, c->scal.p, 8, cudaMemcpyDeviceToHost, st));
  MFB_CUDA(cudaStreamSynchronize(st));
  *e = c->h_scal[0];
  return MFB_OK;
}
