// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_linalg.hpp header).
//
// Per-element polar decomposition and its derivative, restated in the reference's operation order:
//   rodrigues            src/svd/newton_procrustes.h:87-106
//   newton_procrustes    src/svd/newton_procrustes.h:108-189   (constant generators :11-85)
//   polar_svd<3,6>       src/variables/mixed_stretch.cpp:18-53
//   polar_svd<2,3>       src/variables/mixed_stretch.cpp:54-89  (svd: src/svd/svd_eigen.h:7-31, dsvd: src/svd/dsvd.cpp:140-191)
//   flatten / flatten_multiply / flatten_multiply_right   deps/Bartels/include/flatten.h:66-71, flatten_multiply.h:30-121
// The 9x9 products are formed explicitly, as the reference does; the CUDA path uses a contracted form of the
// same algebra, so agreement between the two is a real check.
#pragma once
#include "orc_linalg.hpp"
#include <cstdio>
#include <cstdlib>

namespace orc {

typedef Mat<3, 3> M3;
typedef Mat<9, 9> M9;
typedef Mat<2, 2> M2;
typedef Mat<4, 4> M4;

struct PolarGen {
  M3 cpx, cpy, cpz, cpxx, cpyy, cpzz, cpxy, cpxz, cpyz;
  Mat<9, 3> skew_to_full;
  PolarGen() {
    cpx.zero(); cpy.zero(); cpz.zero();
    cpx(1, 2) = -1.0; cpx(2, 1) = 1.0;
    cpy(0, 2) = 1.0;  cpy(2, 0) = -1.0;
    cpz(0, 1) = -1.0; cpz(1, 0) = 1.0;
    cpxx = mul(cpx, cpx); cpyy = mul(cpy, cpy); cpzz = mul(cpz, cpz);
    cpxy = scaled(add(mul(cpx, cpy), mul(cpy, cpx)), 0.5);
    cpxz = scaled(add(mul(cpx, cpz), mul(cpz, cpx)), 0.5);
    cpyz = scaled(add(mul(cpy, cpz), mul(cpz, cpy)), 0.5);
    skew_to_full.zero();
    // rows are the column-major entries (0,0),(1,0),(2,0),(0,1),... of [w]x ; columns are w_x, w_y, w_z
    skew_to_full(1, 2) = 1.0;  skew_to_full(2, 1) = -1.0;
    skew_to_full(3, 2) = -1.0; skew_to_full(5, 0) = 1.0;
    skew_to_full(6, 1) = 1.0;  skew_to_full(7, 0) = -1.0;
  }
};
inline const PolarGen& gen() { static PolarGen g; return g; }

inline double tr_prod(const M3& A, const M3& B) {  // trace(A*B)
  double s = 0;
  for (int i = 0; i < 3; ++i) for (int k = 0; k < 3; ++k) s += A(i, k) * B(k, i);
  return s;
}

// newton_procrustes.h:87-106
inline void rodrigues(M3& R, const double* omega) {
  const double angle = std::sqrt(omega[0] * omega[0] + omega[1] * omega[1] + omega[2] * omega[2]);
  if (std::fabs(angle) < 1e-8) { R.identity(); return; }
  const double ax = omega[0] / angle, ay = omega[1] / angle, az = omega[2] / angle;
  M3 K; K.zero();
  K(0, 1) = -az; K(0, 2) = ay; K(1, 0) = az; K(1, 2) = -ax; K(2, 0) = -ay; K(2, 1) = ax;
  M3 KK = mul(K, K);
  const double sn = std::sin(angle), cs = 1.0 - std::cos(angle);
  R.identity();
  for (int i = 0; i < 9; ++i) R.a[i] += sn * K.a[i] + cs * KK.a[i];
}

inline void procrustes_hessian(const M3& Y, M3& H) {
  const PolarGen& G = gen();
  H(0, 0) = -tr_prod(G.cpxx, Y); H(0, 1) = -tr_prod(G.cpxy, Y); H(0, 2) = -tr_prod(G.cpxz, Y);
  H(1, 0) = -tr_prod(G.cpxy, Y); H(1, 1) = -tr_prod(G.cpyy, Y); H(1, 2) = -tr_prod(G.cpyz, Y);
  H(2, 0) = -tr_prod(G.cpxz, Y); H(2, 1) = -tr_prod(G.cpyz, Y); H(2, 2) = -tr_prod(G.cpzz, Y);
}

// flatten_multiply<B>(A): 9x9 (or 4x4) matrix of vec(X) -> vec(A*X)          = I (x) A
template <int D>
inline Mat<D * D, D * D> flatten_multiply(const Mat<D, D>& A) {
  Mat<D * D, D * D> out;
  for (int col = 0; col < D * D; ++col)
    for (int row = 0; row < D * D; ++row) {
      const int j = row / D, i = row % D, m = col / D, l = col % D;
      out(row, col) = (j != m) ? 0.0 : A(i, l);
    }
  return out;
}
// flatten_multiply_right<A>(B): matrix of vec(X) -> vec(X*B)                  = B^T (x) I
template <int D>
inline Mat<D * D, D * D> flatten_multiply_right(const Mat<D, D>& B) {
  Mat<D * D, D * D> out;
  for (int col = 0; col < D * D; ++col)
    for (int row = 0; row < D * D; ++row) {
      const int j = row / D, i = row % D, m = col / D, l = col % D;
      out(row, col) = (l != i) ? 0.0 : B(m, j);
    }
  return out;
}

// Status bits returned by newton_procrustes (the reference calls exit(1) on H FAILURE; the oracle reports it).
enum { POLAR_OK = 0, POLAR_MAXITER = 1, POLAR_HFAIL = 2 };

// newton_procrustes(R, A=I, B=F, compute_gradients, dRdF, tol, max_iter)   newton_procrustes.h:108-189
// Returns the number of Newton iterations taken; *status gets the bits above.
inline int newton_procrustes(M3& R, const M3& A, const M3& B, bool compute_gradients, M9& dRdF,
                             double tol, int max_iter, int* status) {
  const PolarGen& G = gen();
  M3 Y = mul(mul(R, A), transpose(B));
  double g[3], omega[3];
  M3 H, dR;
  double E0, E1;
  int itr = 0;
  *status = POLAR_OK;
  do {
    g[0] = -tr_prod(G.cpx, Y); g[1] = -tr_prod(G.cpy, Y); g[2] = -tr_prod(G.cpz, Y);
    if (std::sqrt(g[0] * g[0] + g[1] * g[1] + g[2] * g[2]) < tol) break;
    procrustes_hessian(Y, H);
    Mat<3, 1> gv; gv[0] = g[0]; gv[1] = g[1]; gv[2] = g[2];
    Mat<3, 1> sol = colpiv_qr_solve<3, 1>(H, gv);
    omega[0] = -sol[0]; omega[1] = -sol[1]; omega[2] = -sol[2];
    E0 = -mul(R, Y).trace();          // sic: R*Y, not Y (newton_procrustes.h:142)
    E1 = E0 + 1.0;
    double on;
    do {
      rodrigues(dR, omega);
      E1 = -mul(dR, Y).trace();
      omega[0] *= 0.6; omega[1] *= 0.6; omega[2] *= 0.6;
      on = std::sqrt(omega[0] * omega[0] + omega[1] * omega[1] + omega[2] * omega[2]);
    } while (E1 > E0 && on > tol);
    R = mul(dR, R);
    Y = mul(dR, Y);
    ++itr;
  } while (itr < max_iter);
  if (itr >= max_iter) *status |= POLAR_MAXITER;

  if (compute_gradients) {
    procrustes_hessian(Y, H);
    if (H.norm() < 1e-6) { *status |= POLAR_HFAIL; dRdF.zero(); return itr; }
    Mat<3, 9> dF;
    Y = mul(R, A);
    M3 t;
    t = mul(G.cpx, Y); for (int k = 0; k < 9; ++k) dF(0, k) = -t.a[k];
    t = mul(G.cpy, Y); for (int k = 0; k < 9; ++k) dF(1, k) = -t.a[k];
    t = mul(G.cpz, Y); for (int k = 0; k < 9; ++k) dF(2, k) = -t.a[k];
    Mat<3, 9> X = colpiv_qr_solve<3, 9>(H, dF);
    Mat<9, 9> left = mul(flatten_multiply_right<3>(R), mul(G.skew_to_full, X));
    for (int i = 0; i < 81; ++i) dRdF.a[i] = -left.a[i];
  }
  return itr;
}

// polar_svd<3,6>   mixed_stretch.cpp:18-53.  dsdF is 6x9 (rows = stretch dof), only written when compute_gradients.
inline int polar_svd3(M3& R, double* s, const M3& F, bool compute_gradients, Mat<6, 9>& dsdF, int* status) {
  M9 dRdF;
  M3 I; I.identity();
  const int iters = newton_procrustes(R, I, F, compute_gradients, dRdF, 1e-6, 100, status);
  M3 S = mul(transpose(R), F);
  s[0] = S(0, 0); s[1] = S(1, 1); s[2] = S(2, 2);
  s[3] = 0.5 * (S(1, 0) + S(0, 1));
  s[4] = 0.5 * (S(2, 0) + S(0, 2));
  s[5] = 0.5 * (S(2, 1) + S(1, 2));
  if (compute_gradients) {
    M9 I9; I9.identity();
    M9 J = mul(flatten_multiply<3>(transpose(R)), sub(I9, mul(flatten_multiply_right<3>(S), dRdF)));
    for (int c = 0; c < 9; ++c) {
      dsdF(0, c) = J(0, c); dsdF(1, c) = J(4, c); dsdF(2, c) = J(8, c);
      dsdF(3, c) = 0.5 * (J(1, c) + J(3, c));
      dsdF(4, c) = 0.5 * (J(2, c) + J(6, c));
      dsdF(5, c) = 0.5 * (J(5, c) + J(7, c));
    }
  }
  return iters;
}

// mfem::svd<double,2>   svd_eigen.h:7-31  (flip fix on the last column)
inline void svd2_flipfix(const M2& F, double* sig, M2& U, M2& V) {
  svd2(F, sig, U, V);
  const double detU = U(0, 0) * U(1, 1) - U(0, 1) * U(1, 0);
  if (detU < 0.0) { U(0, 1) = -U(0, 1); U(1, 1) = -U(1, 1); sig[1] = -sig[1]; }
  const double detV = V(0, 0) * V(1, 1) - V(0, 1) * V(1, 0);
  if (detV < 0.0) { V(0, 1) = -V(0, 1); V(1, 1) = -V(1, 1); sig[1] = -sig[1]; }
}

// polar_svd<2,3>   mixed_stretch.cpp:54-89 with dsvd (dsvd.cpp:140-191) inlined.  R input is ignored (not warm started).
inline void polar_svd2(M2& R, double* s, const M2& F, Mat<3, 4>& dsdF) {
  M2 U, V; double sv[2];
  svd2_flipfix(F, sv, U, V);
  // dsvd
  const double tol = 1e-4;
  double d01 = sv[1] * sv[1] - sv[0] * sv[0];
  d01 = (std::fabs(d01) < tol) ? 0.0 : 1.0 / d01;   // 1/inf
  M2 dR_dF[4];
  for (int r = 0; r < 2; ++r)
    for (int c = 0; c < 2; ++c) {
      M2 UVT;  // U.row(r)^T * V.row(c)
      for (int i = 0; i < 2; ++i) for (int j = 0; j < 2; ++j) UVT(i, j) = U(r, i) * V(c, j);
      UVT(0, 0) = 0.0; UVT(1, 1) = 0.0;   // UVT -= diag(UVT)
      // tmp = S*UVT + UVT^T*S
      M2 tmp;
      for (int i = 0; i < 2; ++i) for (int j = 0; j < 2; ++j) tmp(i, j) = sv[i] * UVT(i, j) + UVT(j, i) * sv[j];
      double w01 = tmp(0, 1) * d01;
      M2 W; W(0, 0) = 0; W(0, 1) = w01; W(1, 0) = -w01; W(1, 1) = 0;
      M2 dV = mul(V, W);
      for (int i = 0; i < 2; ++i) for (int j = 0; j < 2; ++j) tmp(i, j) = UVT(i, j) * sv[j] + sv[i] * UVT(j, i);
      w01 = tmp(0, 1) * d01;
      W(0, 1) = w01; W(1, 0) = -w01;
      M2 dU = mul(U, W);
      dR_dF[2 * c + r] = add(mul(dU, transpose(V)), mul(U, transpose(dV)));
    }
  R = mul(U, transpose(V));
  M2 S = mul(transpose(R), F);
  s[0] = S(0, 0); s[1] = S(1, 1); s[2] = 0.5 * (S(1, 0) + S(0, 1));
  M4 dRdF;
  for (int i = 0; i < 4; ++i) for (int k = 0; k < 4; ++k) dRdF(k, i) = dR_dF[i].a[k];
  M4 I4; I4.identity();
  M4 J = mul(flatten_multiply<2>(transpose(R)), sub(I4, mul(flatten_multiply_right<2>(S), dRdF)));
  for (int c = 0; c < 4; ++c) {
    dsdF(0, c) = J(0, c); dsdF(1, c) = J(3, c); dsdF(2, c) = 0.5 * (J(1, c) + J(2, c));
  }
}

// psd_fix_invert   src/utils/psd_fix.h:10-34
template <int N>
inline void psd_fix_invert(Mat<N, N>& A, Mat<N, N>& Ainv, double* evals, double tol) {
  Mat<N, N> V;
  sym_eig<N>(A, evals, V);
  for (int i = 0; i < N; ++i) if (evals[i] < tol) evals[i] = tol;
  for (int i = 0; i < N; ++i)
    for (int j = 0; j < N; ++j) {
      double s = 0, si = 0;
      for (int k = 0; k < N; ++k) { s += V(i, k) * evals[k] * V(j, k); si += V(i, k) * (1.0 / evals[k]) * V(j, k); }
      A(i, j) = s; Ainv(i, j) = si;
    }
}

}  // namespace orc
