// Per-element FP64 math of the mixed-stretch hot path, written for one-thread-per-element execution on sm_100a.
// Everything is straight-line or fully unrolled so that all small matrices live in registers.
//
// What each piece replaces in the reference (evouga/mixed-fem):
//   snh/nh/corot/fcr/arap/fung gradient+hessian ........ src/energies/*.cpp S-form overloads (6d and 3d)
//   procrustes3() + dsdf3() .............................. src/svd/newton_procrustes.h:108-189 and
//                                                          src/variables/mixed_stretch.cpp:18-53 (polar_svd<3,6>)
//   jacobi_eig<N>() + psd_fix_invert<N>() ................ src/utils/psd_fix.h:10-34 (SelfAdjointEigenSolver)
//   polar2() ............................................. src/variables/mixed_stretch.cpp:54-89, src/svd/dsvd.cpp:140-191
//
// The file also compiles as plain C++ (MFB_HD becomes `inline`): tests/ build a host harness from it to check the
// algebra on the CPU box.  That harness is test-only; the product library has no host execution path.
#pragma once
#include <math.h>

#if defined(__CUDACC__)
#define MFB_HD __host__ __device__ __forceinline__
#else
#define MFB_HD inline
#endif

namespace mfb {

enum Model { M_SNH = 0, M_NH = 1, M_COROT = 2, M_FCR = 3, M_ARAP = 4, M_FUNG = 5 };
enum StatusBits { ST_POLAR_MAXITER = 1, ST_POLAR_HFAIL = 2, ST_NONFINITE = 4, ST_EIG_MAXSWEEP = 8 };

// packed lower-triangular index for symmetric matrices, i >= j : idx = i*(i+1)/2 + j
MFB_HD constexpr int symidx(int i, int j) { return i >= j ? i * (i + 1) / 2 + j : j * (j + 1) / 2 + i; }

// -----------------------------------------------------------------------------------------------------------------
// Materials.  psi = a(I2) + b(J) + q |S-I|^2 + ltr/2 (tr S - d)^2 ; gradient / Hessian by the chain rule through
// I2 = tr(S^T S) and J = det S.  Hessian is returned as the packed LOWER triangle (what SelfAdjointEigenSolver
// reads, so the Neohookean (0,2) typo of neohookean.cpp:56 is unobservable on this path).
// -----------------------------------------------------------------------------------------------------------------
struct Coef { double a1, a2, b1, b2, q, ltr; };

MFB_HD Coef material_coefs(int model, double mu, double la, double I2, double J, int d) {
  Coef c; c.a1 = 0; c.a2 = 0; c.b1 = 0; c.b2 = 0; c.q = 0; c.ltr = 0;
  switch (model) {
    case M_SNH: c.a1 = 0.5 * mu; c.b1 = -mu + la * (J - 1.0); c.b2 = la; break;
    case M_NH: { const double lj = log(J); const double ij = 1.0 / J;
      c.a1 = 0.5 * mu; c.b1 = (-mu + la * lj) * ij; c.b2 = (la + mu - la * lj) * ij * ij; break; }
    case M_COROT: c.q = mu; c.ltr = la; break;
    case M_FCR: c.q = mu; c.b1 = la * (J - 1.0); c.b2 = la; break;
    case M_ARAP: c.q = 0.5 * mu; break;
    case M_FUNG: { const double cc = 4.0; const double ex = exp(0.5 * cc * (I2 - d)); const double k = (mu + cc * mu) / la;
      c.a1 = 0.5 * mu + 0.25 * mu * cc * ex; c.a2 = 0.125 * mu * cc * cc * ex; c.b1 = la * (J - 1.0 - k); c.b2 = la; break; }
  }
  return c;
}

MFB_HD double material_energy6(int model, double mu, double la, const double* s) {
  const double I2 = s[0] * s[0] + s[1] * s[1] + s[2] * s[2] + 2.0 * (s[3] * s[3] + s[4] * s[4] + s[5] * s[5]);
  const double J = s[0] * s[1] * s[2] + 2.0 * s[3] * s[4] * s[5] - s[0] * s[5] * s[5] - s[1] * s[4] * s[4] - s[2] * s[3] * s[3];
  const double tr = s[0] + s[1] + s[2];
  const double quad = (s[0] - 1.0) * (s[0] - 1.0) + (s[1] - 1.0) * (s[1] - 1.0) + (s[2] - 1.0) * (s[2] - 1.0) +
                      2.0 * (s[3] * s[3] + s[4] * s[4] + s[5] * s[5]);
  switch (model) {
    case M_SNH: return 0.5 * mu * (I2 - 3.0) - mu * (J - 1.0) + 0.5 * la * (J - 1.0) * (J - 1.0);
    case M_NH: { const double lj = log(J); return 0.5 * mu * (I2 - 3.0) - mu * lj + 0.5 * la * lj * lj; }
    case M_COROT: return 0.5 * la * (tr - 3.0) * (tr - 3.0) + mu * quad;
    case M_FCR: return 0.5 * la * (J - 1.0) * (J - 1.0) + mu * quad;
    case M_ARAP: return 0.5 * mu * quad;
    case M_FUNG: { const double cc = 4.0; const double k = (mu + cc * mu) / la;
      return 0.5 * mu * (I2 - 3.0) + 0.5 * mu * (exp(0.5 * cc * (I2 - 3.0)) - 1.0) + 0.5 * la * (J - 1.0 - k) * (J - 1.0 - k); }
  }
  return 0.0;
}
MFB_HD double material_energy3(int model, double mu, double la, const double* s) {
  const double I2 = s[0] * s[0] + s[1] * s[1] + 2.0 * s[2] * s[2];
  const double J = s[0] * s[1] - s[2] * s[2];
  const double tr = s[0] + s[1];
  const double quad = (s[0] - 1.0) * (s[0] - 1.0) + (s[1] - 1.0) * (s[1] - 1.0) + 2.0 * s[2] * s[2];
  switch (model) {
    case M_SNH: return 0.5 * mu * (I2 - 2.0) - mu * (J - 1.0) + 0.5 * la * (J - 1.0) * (J - 1.0);
    case M_COROT: return 0.5 * la * (tr - 2.0) * (tr - 2.0) + mu * quad;
    case M_FCR: return 0.5 * la * (J - 1.0) * (J - 1.0) + mu * quad;
    case M_ARAP: return 0.5 * mu * quad;
  }
  return 0.0;
}

// g (6) and packed-lower H (21), both multiplied by `scale` (= h^2 on the path, mixed_stretch.cpp:197,201)
MFB_HD void material_grad_hess6(int model, double mu, double la, const double* s, double scale, double* g, double* H) {
  const double I2 = s[0] * s[0] + s[1] * s[1] + s[2] * s[2] + 2.0 * (s[3] * s[3] + s[4] * s[4] + s[5] * s[5]);
  const double J = s[0] * s[1] * s[2] + 2.0 * s[3] * s[4] * s[5] - s[0] * s[5] * s[5] - s[1] * s[4] * s[4] - s[2] * s[3] * s[3];
  const double tr = s[0] + s[1] + s[2];
  double dJ[6];
  dJ[0] = s[1] * s[2] - s[5] * s[5];
  dJ[1] = s[0] * s[2] - s[4] * s[4];
  dJ[2] = s[0] * s[1] - s[3] * s[3];
  dJ[3] = 2.0 * (s[4] * s[5] - s[2] * s[3]);
  dJ[4] = 2.0 * (s[3] * s[5] - s[1] * s[4]);
  dJ[5] = 2.0 * (s[3] * s[4] - s[0] * s[5]);
  const Coef c = material_coefs(model, mu, la, I2, J, 3);
  double dI[6];
#pragma unroll
  for (int i = 0; i < 3; ++i) dI[i] = 2.0 * s[i];
#pragma unroll
  for (int i = 3; i < 6; ++i) dI[i] = 4.0 * s[i];
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    const double dq = (i < 3) ? 2.0 * (s[i] - 1.0) : 4.0 * s[i];
    const double dtr = (i < 3) ? c.ltr * (tr - 3.0) : 0.0;
    g[i] = scale * (c.a1 * dI[i] + c.b1 * dJ[i] + c.q * dq + dtr);
  }
#pragma unroll
  for (int i = 0; i < 6; ++i)
#pragma unroll
    for (int j = 0; j <= i; ++j) {
      double h = c.a2 * dI[i] * dI[j] + c.b2 * dJ[i] * dJ[j];
      if (i == j) h += 2.0 * (c.a1 + c.q) * (i < 3 ? 1.0 : 2.0);
      if (i < 3) h += c.ltr;  // j <= i < 3
      H[symidx(i, j)] = h;
    }
  // + b1 * d2J/ds2 (sparse, linear in s)
  H[symidx(1, 0)] += c.b1 * s[2]; H[symidx(2, 0)] += c.b1 * s[1]; H[symidx(2, 1)] += c.b1 * s[0];
  H[symidx(5, 0)] += c.b1 * (-2.0 * s[5]); H[symidx(4, 1)] += c.b1 * (-2.0 * s[4]); H[symidx(3, 2)] += c.b1 * (-2.0 * s[3]);
  H[symidx(3, 3)] += c.b1 * (-2.0 * s[2]); H[symidx(4, 4)] += c.b1 * (-2.0 * s[1]); H[symidx(5, 5)] += c.b1 * (-2.0 * s[0]);
  H[symidx(4, 3)] += c.b1 * (2.0 * s[5]); H[symidx(5, 3)] += c.b1 * (2.0 * s[4]); H[symidx(5, 4)] += c.b1 * (2.0 * s[3]);
#pragma unroll
  for (int i = 0; i < 21; ++i) H[i] *= scale;
}

MFB_HD void material_grad_hess3(int model, double mu, double la, const double* s, double scale, double* g, double* H) {
  const double I2 = s[0] * s[0] + s[1] * s[1] + 2.0 * s[2] * s[2];
  const double J = s[0] * s[1] - s[2] * s[2];
  const double tr = s[0] + s[1];
  const double dJ[3] = {s[1], s[0], -2.0 * s[2]};
  const double dI[3] = {2.0 * s[0], 2.0 * s[1], 4.0 * s[2]};
  const Coef c = material_coefs(model, mu, la, I2, J, 2);
#pragma unroll
  for (int i = 0; i < 3; ++i) {
    const double dq = (i < 2) ? 2.0 * (s[i] - 1.0) : 4.0 * s[i];
    const double dtr = (i < 2) ? c.ltr * (tr - 2.0) : 0.0;
    g[i] = scale * (c.a1 * dI[i] + c.b1 * dJ[i] + c.q * dq + dtr);
  }
#pragma unroll
  for (int i = 0; i < 3; ++i)
#pragma unroll
    for (int j = 0; j <= i; ++j) {
      double h = c.a2 * dI[i] * dI[j] + c.b2 * dJ[i] * dJ[j];
      if (i == j) h += 2.0 * (c.a1 + c.q) * (i < 2 ? 1.0 : 2.0);
      if (i < 2) h += c.ltr;
      H[symidx(i, j)] = h;
    }
  H[symidx(1, 0)] += c.b1;
  H[symidx(2, 2)] += -2.0 * c.b1;
#pragma unroll
  for (int i = 0; i < 6; ++i) H[i] *= scale;
}

// -----------------------------------------------------------------------------------------------------------------
// Symmetric eigen-solver: cyclic Jacobi with fully unrolled static pivot order (all indices compile-time).
// A is the packed lower triangle (in/out: diagonal ends up holding the eigenvalues), V gets the eigenvectors
// (column-major N x N).  Returns the number of sweeps.
// -----------------------------------------------------------------------------------------------------------------
template <int N, int P, int Q>
MFB_HD void jacobi_rotate(double* A, double* V) {
  const double apq = A[symidx(Q, P)];
  const double app = A[symidx(P, P)], aqq = A[symidx(Q, Q)];
  // skip negligible pivots (keeps the rotation well defined)
  if (fabs(apq) <= 1e-300 || fabs(apq) <= 1e-19 * (fabs(app) + fabs(aqq))) { A[symidx(Q, P)] = 0.0; return; }
  const double theta = 0.5 * (aqq - app) / apq;
  const double t = (theta >= 0.0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
  const double c = 1.0 / sqrt(t * t + 1.0);
  const double s = t * c;
  const double tau = s / (1.0 + c);
  A[symidx(P, P)] = app - t * apq;
  A[symidx(Q, Q)] = aqq + t * apq;
  A[symidx(Q, P)] = 0.0;
#pragma unroll
  for (int r = 0; r < N; ++r) {
    if (r != P && r != Q) {
      const double arp = A[symidx(r, P)], arq = A[symidx(r, Q)];
      A[symidx(r, P)] = arp - s * (arq + tau * arp);
      A[symidx(r, Q)] = arq + s * (arp - tau * arq);
    }
  }
#pragma unroll
  for (int r = 0; r < N; ++r) {
    const double vrp = V[P * N + r], vrq = V[Q * N + r];
    V[P * N + r] = vrp - s * (vrq + tau * vrp);
    V[Q * N + r] = vrq + s * (vrp - tau * vrq);
  }
}

template <int N, int P, int Q>
struct JacobiSweep {
  static MFB_HD void run(double* A, double* V) {
    jacobi_rotate<N, P, Q>(A, V);
    if constexpr (Q + 1 < N) JacobiSweep<N, P, Q + 1>::run(A, V);
    else if constexpr (P + 2 < N) JacobiSweep<N, P + 1, P + 2>::run(A, V);
  }
};

template <int N>
MFB_HD int jacobi_eig(double* A, double* V) {
  constexpr int NP = N * (N + 1) / 2;
#pragma unroll
  for (int i = 0; i < N * N; ++i) V[i] = 0.0;
#pragma unroll
  for (int i = 0; i < N; ++i) V[i * N + i] = 1.0;
  int sweep = 0;
  for (; sweep < 30; ++sweep) {
    double off = 0.0, tot = 0.0;
#pragma unroll
    for (int i = 0; i < N; ++i)
#pragma unroll
      for (int j = 0; j <= i; ++j) {
        const double a = A[symidx(i, j)];
        if (i == j) tot += a * a; else off += a * a;
      }
    tot += 2.0 * off;
    if (!(off > 1e-28 * tot)) break;   // also exits on NaN
    JacobiSweep<N, 0, 1>::run(A, V);
  }
  (void)NP;
  return sweep;
}

// -----------------------------------------------------------------------------------------------------------------
// Parallel-order Jacobi for N = 6: the 15 pivots of a sweep are visited as 5 rounds of 3 DISJOINT pairs (round-robin
// tournament).  The three rotation parameter sets of a round depend only on their own 2x2 diagonal blocks, which the
// other two rotations of the round never touch, so they are computed up front: three independent div/sqrt chains in
// flight instead of one (the kernel is latency-bound, not FP64-throughput-bound).  Same arithmetic as a sequential
// cyclic Jacobi with this pivot order.  Branch-free: a negligible pivot gets the identity rotation.
// -----------------------------------------------------------------------------------------------------------------
struct JRot { double t, s, c; };

// Reciprocal and reciprocal square root without the IEEE slow paths: hardware seed (MUFU.RCP64H / MUFU.RSQ64H, ~20
// bits) + two Newton steps, relative error a few ulp, for positive normal arguments only (what Jacobi feeds them).
MFB_HD double fast_rcp(double x) {
#if defined(__CUDA_ARCH__)
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  double e = fma(-x, y, 1.0);
  y = fma(y, e, y);
  e = fma(-x, y, 1.0);
  return fma(y, e, y);
#else
  return 1.0 / x;
#endif
}
MFB_HD double fast_rsqrt(double x) {
#if defined(__CUDA_ARCH__)
  double y;
  asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
  double h = 0.5 * x;
  double e = fma(-h * y, y, 0.5);
  y = fma(y, e, y);
  e = fma(-h * y, y, 0.5);
  return fma(y, e, y);
#else
  return 1.0 / sqrt(x);
#endif
}

// tan / sin / cos of the Jacobi angle for the pivot (p,q):  t = 2 apq sgn(d) / (|d| + sqrt(d^2 + 4 apq^2)), d = aqq - app
MFB_HD JRot jacobi_params(double app, double aqq, double apq) {
  const bool tiny = !(fabs(apq) > 1e-300) || fabs(apq) <= 1e-19 * (fabs(app) + fabs(aqq));
  const double d = aqq - app;
  const double q = d * d + 4.0 * apq * apq;
  const double qs = tiny ? 1.0 : q;
  const double root = qs * fast_rsqrt(qs);              // sqrt(q)
  const double num = (d >= 0.0) ? 2.0 * apq : -2.0 * apq;
  double t = num * fast_rcp(fabs(d) + root);
  t = tiny ? 0.0 : t;
  const double c = fast_rsqrt(fma(t, t, 1.0));
  JRot r; r.t = t; r.c = c; r.s = t * c;
  return r;
}

template <int N, int P, int Q>
MFB_HD void jacobi_apply(double* A, double* V, const JRot r) {
  const double apq = A[symidx(Q, P)];
  A[symidx(P, P)] -= r.t * apq;
  A[symidx(Q, Q)] += r.t * apq;
  A[symidx(Q, P)] = 0.0;
#pragma unroll
  for (int k = 0; k < N; ++k) {
    if (k != P && k != Q) {
      const double akp = A[symidx(k, P)], akq = A[symidx(k, Q)];
      A[symidx(k, P)] = fma(r.c, akp, -r.s * akq);
      A[symidx(k, Q)] = fma(r.s, akp, r.c * akq);
    }
  }
#pragma unroll
  for (int k = 0; k < N; ++k) {
    const double vkp = V[P * N + k], vkq = V[Q * N + k];
    V[P * N + k] = fma(r.c, vkp, -r.s * vkq);
    V[Q * N + k] = fma(r.s, vkp, r.c * vkq);
  }
}

template <int P0, int Q0, int P1, int Q1, int P2, int Q2>
MFB_HD void jacobi_round6(double* A, double* V) {
  const JRot r0 = jacobi_params(A[symidx(P0, P0)], A[symidx(Q0, Q0)], A[symidx(Q0, P0)]);
  const JRot r1 = jacobi_params(A[symidx(P1, P1)], A[symidx(Q1, Q1)], A[symidx(Q1, P1)]);
  const JRot r2 = jacobi_params(A[symidx(P2, P2)], A[symidx(Q2, Q2)], A[symidx(Q2, P2)]);
  jacobi_apply<6, P0, Q0>(A, V, r0);
  jacobi_apply<6, P1, Q1>(A, V, r1);
  jacobi_apply<6, P2, Q2>(A, V, r2);
}

MFB_HD int jacobi_eig6_parallel(double* A, double* V) {
#pragma unroll
  for (int i = 0; i < 36; ++i) V[i] = 0.0;
#pragma unroll
  for (int i = 0; i < 6; ++i) V[i * 6 + i] = 1.0;
  int sweep = 0;
#pragma unroll 1
  for (; sweep < 30; ++sweep) {
    double off = 0.0, tot = 0.0;
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int j = 0; j <= i; ++j) {
        const double a = A[symidx(i, j)];
        if (i == j) tot += a * a; else off += a * a;
      }
    tot += 2.0 * off;
    if (!(off > 1e-28 * tot)) break;
    jacobi_round6<0, 5, 1, 4, 2, 3>(A, V);
    jacobi_round6<0, 4, 3, 5, 1, 2>(A, V);
    jacobi_round6<0, 3, 2, 4, 1, 5>(A, V);
    jacobi_round6<0, 2, 1, 3, 4, 5>(A, V);
    jacobi_round6<0, 1, 2, 5, 3, 4>(A, V);
  }
  return sweep;
}

// psd_fix_invert (src/utils/psd_fix.h:10-34): H <- V max(L,tol) V^T, Hinv <- V max(L,tol)^-1 V^T, returns lambda_max
// (after clamping).  H, Hinv packed lower.  *sweeps gets the Jacobi sweep count.
template <int N>
MFB_HD double psd_fix_invert(double* H, double* Hinv, double tol, int* sweeps) {
  double V[N * N];
  if constexpr (N == 6) *sweeps = jacobi_eig6_parallel(H, V);
  else *sweeps = jacobi_eig<N>(H, V);
  double ev[N], iv[N];
  double lmax = -1e300;
#pragma unroll
  for (int k = 0; k < N; ++k) {
    double e = H[symidx(k, k)];
    e = (e < tol) ? tol : e;
    ev[k] = e; iv[k] = 1.0 / e;
    lmax = e > lmax ? e : lmax;
  }
#pragma unroll
  for (int i = 0; i < N; ++i)
#pragma unroll
    for (int j = 0; j <= i; ++j) {
      double a = 0.0, b = 0.0;
#pragma unroll
      for (int k = 0; k < N; ++k) {
        const double vv = V[k * N + i] * V[k * N + j];
        a += vv * ev[k]; b += vv * iv[k];
      }
      H[symidx(i, j)] = a; Hinv[symidx(i, j)] = b;
    }
  return lmax;
}

// -----------------------------------------------------------------------------------------------------------------
// psd_fix_invert FAST PATH for N = 6 (src/utils/psd_fix.h:10-34 when no eigenvalue is clamped).
// If every eigenvalue of H exceeds tol, the reference's  V max(L,tol) V^T  is H itself and  V max(L,tol)^-1 V^T  is
// H^-1: no eigenvectors are needed.  What is still needed from the spectrum is lambda_max (evals_).
//   1. Cholesky H = L L^T with relative pivot check (pivot > relpiv * H_jj, i.e. kappa(H) <~ 1/relpiv);
//   2. H^-1 = L^-T L^-1; certificate lambda_min(H) >= 1/trace(H^-1) > tol;
//   3. lambda_max: Householder tridiagonalisation (no Q), then Newton on the characteristic polynomial from the
//      upper bound |H|_F: monotone from above for a real-rooted polynomial, quadratic for a separated top eigenvalue.
// Returns false (caller must take the Jacobi path) when a pivot fails, the certificate fails, Newton has not
// converged in 24 iterations (a cluster of eigenvalues at the top: the root is then ill-conditioned as a
// polynomial root although not as an eigenvalue) or anything is non-finite.
// Outputs: L (packed lower Cholesky factor), Hinv (packed lower), *lmax.
// -----------------------------------------------------------------------------------------------------------------
template <int K>
MFB_HD void tridiag6_step(double* A) {
  // Householder step K of the reduction of the packed-lower symmetric 6x6 A to tridiagonal form
  constexpr int m = 5 - K;             // length of the column below the diagonal
  double v[m];
#pragma unroll
  for (int i = 0; i < m; ++i) v[i] = A[symidx(K + 1 + i, K)];
  double sig = 0.0;
#pragma unroll
  for (int i = 1; i < m; ++i) sig = fma(v[i], v[i], sig);
  const bool act = sig > 0.0;
  const double q = act ? fma(v[0], v[0], sig) : 1.0;
  const double nrm = q * fast_rsqrt(q);
  const double alpha = act ? (v[0] >= 0.0 ? -nrm : nrm) : v[0];
  v[0] -= alpha;                        // 0 when inactive
  const double vv = act ? fma(v[0], v[0], sig) : 1.0;
  const double beta = act ? 2.0 * fast_rcp(vv) : 0.0;
  double p[m];
#pragma unroll
  for (int i = 0; i < m; ++i) {
    double a = 0.0;
#pragma unroll
    for (int j = 0; j < m; ++j) a = fma(A[symidx(K + 1 + i, K + 1 + j)], v[j], a);
    p[i] = beta * a;
  }
  double kk = 0.0;
#pragma unroll
  for (int i = 0; i < m; ++i) kk = fma(v[i], p[i], kk);
  kk *= 0.5 * beta;
#pragma unroll
  for (int i = 0; i < m; ++i) p[i] = fma(-kk, v[i], p[i]);       // w
#pragma unroll
  for (int i = 0; i < m; ++i)
#pragma unroll
    for (int j = 0; j <= i; ++j) A[symidx(K + 1 + i, K + 1 + j)] -= fma(v[i], p[j], p[i] * v[j]);
  A[symidx(K + 1, K)] = alpha;
}

// steps 1 and 2: Cholesky factor, inverse, lambda_min certificate
MFB_HD bool psd_fast6_factor(const double* H, double* L, double* Hinv, double tol, double relpiv) {
  bool ok = true;
  // ---- 1. Cholesky with pivot check; id[j] = 1 / L_jj
  double id[6];
#pragma unroll
  for (int i = 0; i < 21; ++i) L[i] = H[i];
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    double d = L[symidx(j, j)];
#pragma unroll
    for (int k = 0; k < j; ++k) d = fma(-L[symidx(j, k)], L[symidx(j, k)], d);
    const bool good = d > relpiv * H[symidx(j, j)] && d > 0.0;    // false on NaN as well
    ok = ok && good;
    d = good ? d : 1.0;
    const double r = fast_rsqrt(d);
    id[j] = r;
    L[symidx(j, j)] = d * r;
#pragma unroll
    for (int i = j + 1; i < 6; ++i) {
      double a = L[symidx(i, j)];
#pragma unroll
      for (int k = 0; k < j; ++k) a = fma(-L[symidx(i, k)], L[symidx(j, k)], a);
      L[symidx(i, j)] = a * r;
    }
  }
  // ---- 2. inverse: W = L^-1 (lower), Hinv = W^T W
  double W[21];
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    W[symidx(j, j)] = id[j];
#pragma unroll
    for (int i = j + 1; i < 6; ++i) {
      double a = 0.0;
#pragma unroll
      for (int k = j; k < i; ++k) a = fma(L[symidx(i, k)], W[symidx(k, j)], a);
      W[symidx(i, j)] = -a * id[i];
    }
  }
  double tr = 0.0;
#pragma unroll
  for (int i = 0; i < 6; ++i)
#pragma unroll
    for (int j = 0; j <= i; ++j) {
      double a = 0.0;
#pragma unroll
      for (int k = i; k < 6; ++k) a = fma(W[symidx(k, i)], W[symidx(k, j)], a);
      Hinv[symidx(i, j)] = a;
      if (i == j) tr += a;
    }
  ok = ok && (tr * tol < 1.0);          // lambda_min >= 1 / trace(H^-1) > tol ; false on NaN
  return ok;
}

// step 3: lambda_max
MFB_HD bool psd_fast6_lmax(const double* H, double* lmax, int* newton_iters) {
  bool ok = true;
  double A[21];
  double fro = 0.0;
#pragma unroll
  for (int i = 0; i < 6; ++i)
#pragma unroll
    for (int j = 0; j <= i; ++j) { const double a = H[symidx(i, j)]; fro = fma((i == j) ? a : 2.0 * a, a, fro); }
  const bool fok = fro > 0.0 && fro < 1e300;
  ok = ok && fok;
  fro = fok ? fro : 1.0;
  const double inrm = fast_rsqrt(fro), nrm = fro * inrm;
#pragma unroll
  for (int i = 0; i < 21; ++i) A[i] = H[i] * inrm;
  tridiag6_step<0>(A); tridiag6_step<1>(A); tridiag6_step<2>(A); tridiag6_step<3>(A);
  double d[6], e2[5];
#pragma unroll
  for (int i = 0; i < 6; ++i) d[i] = A[symidx(i, i)];
#pragma unroll
  for (int i = 0; i < 5; ++i) e2[i] = A[symidx(i + 1, i)] * A[symidx(i + 1, i)];
  // Deflation: a negligible off-diagonal (|e| <= 1e-12 |A|_F) splits T.  A 1x1 block IS an eigenvalue and is taken out
  // of the polynomial (factor 1): constant Hessians (ARAP: diagonal; Corot and SNH at rest: the shear block 4 mu I)
  // have a multiple top eigenvalue made of such blocks, on which Newton would only converge linearly.
  double w[6], liso = -1.0;
  int nfree = 0;
#pragma unroll
  for (int i = 0; i < 5; ++i) e2[i] = (e2[i] <= 1e-24) ? 0.0 : e2[i];
#pragma unroll
  for (int i = 0; i < 6; ++i) {
    const bool iso = (i == 0 || e2[i - 1] == 0.0) && (i == 5 || e2[i] == 0.0);
    w[i] = iso ? 0.0 : 1.0;
    liso = iso ? fmax(liso, d[i]) : liso;
    nfree += iso ? 0 : 1;
  }
  double sg = 1.0;                      // |A|_F = 1 >= lambda_max(A)
  bool conv = false;
  int it = 0;
  if (nfree == 0) { sg = liso; conv = true; }
  if (nfree == 6) {                     // the common case: nothing deflated, plain recurrence
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
    for (; it < 24; ++it) {
      double p0 = 1.0, p1 = sg - d[0], q0 = 0.0, q1 = 1.0;
#pragma unroll
      for (int i = 1; i < 6; ++i) {
        const double t = sg - d[i];
        const double p2 = fma(t, p1, -e2[i - 1] * p0);
        const double q2 = fma(t, q1, fma(-e2[i - 1], q0, p1));
        p0 = p1; p1 = p2; q0 = q1; q1 = q2;
      }
      if (!(q1 > 0.0)) break;
      const double dl = p1 * fast_rcp(q1);
      sg -= dl;
      if (!(dl > 1e-15)) { conv = dl > -1e-12; ++it; break; }
    }
  } else
#if defined(__CUDA_ARCH__)
#pragma unroll 1
#endif
  for (; it < 24 && nfree > 0; ++it) {
    double p0 = 1.0, p1 = (w[0] != 0.0) ? sg - d[0] : 1.0, q0 = 0.0, q1 = w[0];
#pragma unroll
    for (int i = 1; i < 6; ++i) {
      const double t = (w[i] != 0.0) ? sg - d[i] : 1.0;
      const double p2 = fma(t, p1, -e2[i - 1] * p0);
      const double q2 = fma(t, q1, fma(-e2[i - 1], q0, w[i] * p1));
      p0 = p1; p1 = p2; q0 = q1; q1 = q2;
    }
    if (!(q1 > 0.0)) break;             // cannot happen above the largest root; guards NaN
    const double dl = p1 * fast_rcp(q1);
    sg -= dl;
    if (!(dl > 1e-15)) { conv = dl > -1e-12; ++it; break; }   // a large negative step = noise near a clustered root
  }
  sg = fmax(sg, liso);
  ok = ok && conv;
  *lmax = sg * nrm;
  *newton_iters = it;
  return ok;
}


MFB_HD bool psd_fast6(const double* H, double* L, double* Hinv, double tol, double relpiv, double* lmax, int* newton_iters) {
  const bool ok1 = psd_fast6_factor(H, L, Hinv, tol, relpiv);
  const bool ok2 = psd_fast6_lmax(H, lmax, newton_iters);
  return ok1 && ok2;
}

// -----------------------------------------------------------------------------------------------------------------
// 3x3 helpers.  Matrices are column-major double[9]: m[3*c + r].
// -----------------------------------------------------------------------------------------------------------------
MFB_HD void cross3(const double* a, const double* b, double* o) {
  o[0] = a[1] * b[2] - a[2] * b[1]; o[1] = a[2] * b[0] - a[0] * b[2]; o[2] = a[0] * b[1] - a[1] * b[0];
}
// C = A*B
MFB_HD void mul33(const double* A, const double* B, double* C) {
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int r = 0; r < 3; ++r) C[3 * c + r] = A[r] * B[3 * c] + A[3 + r] * B[3 * c + 1] + A[6 + r] * B[3 * c + 2];
}
// inverse of the symmetric 3x3 [h00 h01 h02; . h11 h12; . . h22] -> packed (i00,i01,i02,i11,i12,i22)
MFB_HD void syminv3(double h00, double h01, double h02, double h11, double h12, double h22, double* inv) {
  const double c00 = h11 * h22 - h12 * h12;
  const double c01 = h02 * h12 - h01 * h22;
  const double c02 = h01 * h12 - h02 * h11;
  const double det = h00 * c00 + h01 * c01 + h02 * c02;
  const double id = 1.0 / det;
  inv[0] = c00 * id; inv[1] = c01 * id; inv[2] = c02 * id;
  inv[3] = (h00 * h22 - h02 * h02) * id;
  inv[4] = (h01 * h02 - h00 * h12) * id;
  inv[5] = (h00 * h11 - h01 * h01) * id;
}
MFB_HD void symmul3(const double* inv, const double* v, double* o) {
  o[0] = inv[0] * v[0] + inv[1] * v[1] + inv[2] * v[2];
  o[1] = inv[1] * v[0] + inv[3] * v[1] + inv[4] * v[2];
  o[2] = inv[2] * v[0] + inv[4] * v[1] + inv[5] * v[2];
}

// rodrigues (newton_procrustes.h:87-106)
MFB_HD void rodrigues3(const double* w, double* R) {
  const double angle = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
  if (fabs(angle) < 1e-8) {
    R[0] = 1; R[1] = 0; R[2] = 0; R[3] = 0; R[4] = 1; R[5] = 0; R[6] = 0; R[7] = 0; R[8] = 1;
    return;
  }
  const double ia = 1.0 / angle;
  const double x = w[0] * ia, y = w[1] * ia, z = w[2] * ia;
  double sn, cs;
#if defined(__CUDA_ARCH__)
  sincos(angle, &sn, &cs);
#else
  sn = sin(angle); cs = cos(angle);
#endif
  const double oc = 1.0 - cs;
  // I + sn*K + oc*K^2 with K = [w]x / angle ; K^2 = a a^T - I
  R[0] = 1.0 + oc * (x * x - 1.0); R[4] = 1.0 + oc * (y * y - 1.0); R[8] = 1.0 + oc * (z * z - 1.0);
  R[1] = sn * z + oc * x * y;  R[3] = -sn * z + oc * x * y;
  R[2] = -sn * y + oc * x * z; R[6] = sn * y + oc * x * z;
  R[5] = sn * x + oc * y * z;  R[7] = -sn * x + oc * y * z;
}

// Warm-started Newton polar decomposition: newton_procrustes(R, I, F, ...) of newton_procrustes.h:108-189 with the
// generator traces written out:  Y = R F^T, g = axl-like differences of Y, H = tr(Y) I - sym(Y), omega = -H^-1 g,
// back-tracking on E1 = -tr(dR Y) against E0 = -tr(R Y) (sic), R <- dR R, Y <- dR Y.
// On exit Y holds the iterated R F^T.  Returns the iteration count.
MFB_HD int procrustes3(double* R, const double* F, double* Y, double tol, int max_iter, int* status) {
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int r = 0; r < 3; ++r) Y[3 * c + r] = R[r] * F[c] + R[3 + r] * F[3 + c] + R[6 + r] * F[6 + c];
  int itr = 0;
  do {
    const double g0 = Y[5] - Y[7], g1 = Y[6] - Y[2], g2 = Y[1] - Y[3];   // Y21-Y12, Y02-Y20, Y10-Y01
    if (sqrt(g0 * g0 + g1 * g1 + g2 * g2) < tol) break;
    double inv[6];
    syminv3(Y[4] + Y[8], -0.5 * (Y[3] + Y[1]), -0.5 * (Y[6] + Y[2]), Y[0] + Y[8], -0.5 * (Y[7] + Y[5]), Y[0] + Y[4], inv);
    double w[3];
    const double gv[3] = {g0, g1, g2};
    symmul3(inv, gv, w);
    w[0] = -w[0]; w[1] = -w[1]; w[2] = -w[2];
    // E0 = -trace(R*Y)
    double E0 = 0.0;
#pragma unroll
    for (int i = 0; i < 3; ++i) E0 -= R[i] * Y[3 * i] + R[3 + i] * Y[3 * i + 1] + R[6 + i] * Y[3 * i + 2];
    double dR[9], E1, wn;
    do {
      rodrigues3(w, dR);
      E1 = 0.0;
#pragma unroll
      for (int i = 0; i < 3; ++i) E1 -= dR[i] * Y[3 * i] + dR[3 + i] * Y[3 * i + 1] + dR[6 + i] * Y[3 * i + 2];
      w[0] *= 0.6; w[1] *= 0.6; w[2] *= 0.6;
      wn = sqrt(w[0] * w[0] + w[1] * w[1] + w[2] * w[2]);
    } while (E1 > E0 && wn > tol);
    double t[9];
    mul33(dR, R, t);
#pragma unroll
    for (int i = 0; i < 9; ++i) R[i] = t[i];
    mul33(dR, Y, t);
#pragma unroll
    for (int i = 0; i < 9; ++i) Y[i] = t[i];
    ++itr;
  } while (itr < max_iter);
  if (itr >= max_iter) *status |= ST_POLAR_MAXITER;
  return itr;
}

// S = R^T F and its 6-vector (mixed_stretch.cpp:36-40)
MFB_HD void stretch6(const double* R, const double* F, double* s) {
  double S[9];
#pragma unroll
  for (int c = 0; c < 3; ++c)
#pragma unroll
    for (int r = 0; r < 3; ++r) S[3 * c + r] = R[3 * r] * F[3 * c] + R[3 * r + 1] * F[3 * c + 1] + R[3 * r + 2] * F[3 * c + 2];
  s[0] = S[0]; s[1] = S[4]; s[2] = S[8];
  s[3] = 0.5 * (S[1] + S[3]); s[4] = 0.5 * (S[2] + S[6]); s[5] = 0.5 * (S[5] + S[7]);
}

// ds/dF in contracted form.  With Hp = tr(Y) I - sym(Y) at the final Y, C_ij = f_j x r_i (f, r = columns of F, R):
//   Js[n][3b+a] = P_n(a,b) + ( r_b x Hp^-1 c_n )_a ,   c_n = C_ii (n=i<3) or (C_ij + C_ji)/2 (n = 3,4,5 <-> (0,1),(0,2),(1,2)),
//   P_n(a,b) = R(a,i) delta(b,i)  or  (R(a,i) delta(b,j) + R(a,j) delta(b,i))/2.
// This is algebraically identical to  J = (I(x)R^T)(I - (S^T(x)I) dRdF),  dRdF = -(R^T(x)I) Skew H^-1 dF  of
// mixed_stretch.cpp:43-52 / newton_procrustes.h:161-187 (the oracle forms those 9x9 products explicitly).
// Output: column n of Js, 9 values indexed [3b+a].  Hinv3 = packed inverse of Hp.
MFB_HD void dsdf3_column(int n, const double* R, const double* F, const double* Hinv3, double* col) {
  const int i = (n < 3) ? n : (n == 5 ? 1 : 0);
  const int j = (n < 3) ? n : (n == 3 ? 1 : 2);
  double c[3];
  if (n < 3) {
    cross3(&F[3 * i], &R[3 * i], c);
  } else {
    double c1[3], c2[3];
    cross3(&F[3 * j], &R[3 * i], c1);   // C_ij = f_j x r_i
    cross3(&F[3 * i], &R[3 * j], c2);   // C_ji = f_i x r_j
    c[0] = 0.5 * (c1[0] + c2[0]); c[1] = 0.5 * (c1[1] + c2[1]); c[2] = 0.5 * (c1[2] + c2[2]);
  }
  double d[3];
  symmul3(Hinv3, c, d);
#pragma unroll
  for (int b = 0; b < 3; ++b) {
    double v[3];
    cross3(&R[3 * b], d, v);
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      double p = 0.0;
      if (n < 3) { if (b == i) p = R[3 * i + a]; }
      else { if (b == j) p = 0.5 * R[3 * i + a]; if (b == i) p = 0.5 * R[3 * j + a]; }
      col[3 * b + a] = p + v[a];
    }
  }
}

// Same column for a RUN-TIME n (the caller keeps its loop over n rolled: one body instead of six in the instruction
// stream).  Columns of R and F are picked with selects; for n < 3 (i == j) the symmetrised expressions reduce to the
// unsymmetrised ones exactly (0.5 (c + c) = c, 0.5 r + 0.5 r = r), so the result is bit-identical to dsdf3_column.
MFB_HD void dsdf3_column_rt(int n, const double* R, const double* F, const double* Hinv3, double* col) {
  const int i = (n < 3) ? n : (n == 5 ? 1 : 0);
  const int j = (n < 3) ? n : (n == 3 ? 1 : 2);
  double ri[3], rj[3], fi[3], fj[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    ri[a] = (i == 0) ? R[a] : (i == 1) ? R[3 + a] : R[6 + a];
    rj[a] = (j == 0) ? R[a] : (j == 1) ? R[3 + a] : R[6 + a];
    fi[a] = (i == 0) ? F[a] : (i == 1) ? F[3 + a] : F[6 + a];
    fj[a] = (j == 0) ? F[a] : (j == 1) ? F[3 + a] : F[6 + a];
  }
  double c1[3], c2[3], c[3], d[3];
  cross3(fj, ri, c1);
  cross3(fi, rj, c2);
  c[0] = 0.5 * (c1[0] + c2[0]); c[1] = 0.5 * (c1[1] + c2[1]); c[2] = 0.5 * (c1[2] + c2[2]);
  symmul3(Hinv3, c, d);
#pragma unroll
  for (int b = 0; b < 3; ++b) {
    double v[3];
    cross3(&R[3 * b], d, v);
#pragma unroll
    for (int a = 0; a < 3; ++a) col[3 * b + a] = (0.5 * ((b == j) ? ri[a] : 0.0) + 0.5 * ((b == i) ? rj[a] : 0.0)) + v[a];
  }
}

// Hp^-1 from the final Y; returns false on the reference's "H FAILURE" condition (|H|_F < 1e-6)
MFB_HD bool procrustes_hinv(const double* Y, double* Hinv3) {
  const double h00 = Y[4] + Y[8], h11 = Y[0] + Y[8], h22 = Y[0] + Y[4];
  const double h01 = -0.5 * (Y[3] + Y[1]), h02 = -0.5 * (Y[6] + Y[2]), h12 = -0.5 * (Y[7] + Y[5]);
  const double n2 = h00 * h00 + h11 * h11 + h22 * h22 + 2.0 * (h01 * h01 + h02 * h02 + h12 * h12);
  syminv3(h00, h01, h02, h11, h12, h22, Hinv3);
  return !(n2 < 1e-12);
}

// -----------------------------------------------------------------------------------------------------------------
// 2D polar via closed-form 2x2 SVD (mixed_stretch.cpp:54-89).  R = U V^T is the rotation of the polar factor (flip
// fix makes det R = +1 and lets sigma_1 go negative); ds/dF (3x4) from dU, dV of dsvd.cpp:140-191.
// F, R column-major double[4]; Js is 3x4 as Js[n][k], k = 2b + a.
// -----------------------------------------------------------------------------------------------------------------
MFB_HD void polar2(const double* F, double* R, double* s, double Js[3][4]) {
  const double a = F[0], c = F[1], b = F[2], d = F[3];
  const double E = 0.5 * (a + d), Fm = 0.5 * (a - d), G = 0.5 * (c + b), Hh = 0.5 * (c - b);
  const double Q = sqrt(E * E + Hh * Hh), Rr = sqrt(Fm * Fm + G * G);
  const double s0 = Q + Rr, s1 = Q - Rr;           // signed: s1 < 0 <=> det F < 0 (after the flip fix)
  // rotation angle of R = Rot(a2): cos = E/Q, sin = Hh/Q
  double cr, sr;
  if (Q > 0.0) { cr = E / Q; sr = Hh / Q; } else { cr = 1.0; sr = 0.0; }
  R[0] = cr; R[1] = sr; R[2] = -sr; R[3] = cr;
  // S = R^T F
  const double S00 = cr * a + sr * c, S01 = cr * b + sr * d, S10 = -sr * a + cr * c, S11 = -sr * b + cr * d;
  s[0] = S00; s[1] = S11; s[2] = 0.5 * (S10 + S01);
  // dR/dF: R depends on F only through the angle a2 = atan2(Hh, E):  dR = Rot'(a2) * d(a2),
  // d(a2)/dF = (E dHh - Hh dE)/Q^2.  The reference's dsvd route gives the same derivative wherever
  // |s1^2 - s0^2| >= 1e-4 and ZERO inside that band (dsvd.cpp:163-167): reproduce the band.
  const double gap = s1 * s1 - s0 * s0;
  const bool dead = fabs(gap) < 1e-4;
  const double iq2 = (Q > 0.0 && !dead) ? 1.0 / (Q * Q) : 0.0;
  // dE/dF = (1/2,0,0,1/2) for (a,c,b,d) ; dHh/dF = (0,1/2,-1/2,0)  in storage order k = (a:0, c:1, b:2, d:3)
  const double dth[4] = {-0.5 * Hh * iq2, 0.5 * E * iq2, -0.5 * E * iq2, -0.5 * Hh * iq2};
  // dS = R^T dF - R^T dR S, with R^T dR = [0 -1; 1 0] dth
  const double SS[4] = {S00, S10, S01, S11};   // column-major S
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int aa = k & 1, bb = k >> 1;           // dF = e_aa e_bb^T
    double dS[4] = {0, 0, 0, 0};
    // R^T e_aa e_bb^T : column bb gets R^T(:,aa) = (R(aa,0), R(aa,1))
    dS[2 * bb + 0] += R[0 + aa];   // R(aa,0) = R[2*0+aa]
    dS[2 * bb + 1] += R[2 + aa];   // R(aa,1)
    // - [0 -1;1 0] S dth : row0 = +S(1,:), row1 = -S(0,:)
    dS[0] += dth[k] * SS[1]; dS[2] += dth[k] * SS[3];
    dS[1] -= dth[k] * SS[0]; dS[3] -= dth[k] * SS[2];
    Js[0][k] = dS[0]; Js[1][k] = dS[3]; Js[2][k] = 0.5 * (dS[1] + dS[2]);
  }
}

}  // namespace mfb
