// ORACLE — TEST INFRASTRUCTURE ONLY (see orc_linalg.hpp header).
//
// S-form material models of the reference, restated from their definitions (scripts/derivs.m:20-94) rather
// than from the MATLAB-generated expressions, and pinned against the reference's own generated code through
// tests/golden/materials_golden.json (made by tests/golden/make_materials_golden.py, which evaluates the
// expression text of src/energies/*.cpp at fixed probe points).
//
//   s (3D) = [S00 S11 S22 S01 S02 S12],  s (2D) = [S00 S11 S01]          (src/variables/mixed_stretch.cpp:37-40,76)
//   I2 = tr(S^T S),  J = det S
//   Stable-Neohookean  psi = mu/2 (I2-d) - mu (J-1) + la/2 (J-1)^2       src/energies/stable_neohookean.cpp:8-91,232-272
//   Neohookean         psi = mu/2 (I2-3) - mu log J + la/2 log^2 J       src/energies/neohookean.cpp:8-91 (3D only)
//   Corot              psi = la/2 (tr S - d)^2 + mu |S-I|^2              src/energies/corotational.cpp:7-96
//   FixedCorot         psi = la/2 (J-1)^2 + mu |S-I|^2                   src/energies/fixed_corotational.cpp:10-134
//   ARAP               psi = mu/2 |S-I|^2                                src/energies/arap.cpp:9-83
//   Fung (c = 4)       psi = mu/2 (I2-3) + la/2 (J-1-(mu+c mu)/la)^2 + mu/2 (exp(c/2 (I2-3)) - 1)
//                                                                        src/energies/fung.cpp:9-94 (3D only)
// Quirk kept: Neohookean::hessian entry (0,2) carries the reference's typo "S2*3" (neohookean.cpp:56); it sits
// in the strict upper triangle, which psd_fix_invert (SelfAdjointEigenSolver, lower triangle) never reads.
#pragma once
#include <cmath>

namespace orc {

enum Model { M_SNH = 0, M_NH = 1, M_COROT = 2, M_FCR = 3, M_ARAP = 4, M_FUNG = 5 };
static const double kFungC = 4.0;  // src/energies/fung.cpp:7

struct Material { int model; double mu, la; };

// ---- invariants of the 6-vector stretch ----
struct Inv6 {
  double I2, J, tr;
  double dJ[6];      // dJ/ds
  double K[6][6];    // d2J/ds2
  double dI[6];      // dI2/ds
};
inline Inv6 invariants6(const double* s) {
  Inv6 v;
  v.tr = s[0] + s[1] + s[2];
  v.I2 = s[0] * s[0] + s[1] * s[1] + s[2] * s[2] + 2.0 * (s[3] * s[3] + s[4] * s[4] + s[5] * s[5]);
  v.J = s[0] * s[1] * s[2] + 2.0 * s[3] * s[4] * s[5] - s[0] * s[5] * s[5] - s[1] * s[4] * s[4] - s[2] * s[3] * s[3];
  v.dJ[0] = s[1] * s[2] - s[5] * s[5];
  v.dJ[1] = s[0] * s[2] - s[4] * s[4];
  v.dJ[2] = s[0] * s[1] - s[3] * s[3];
  v.dJ[3] = 2.0 * (s[4] * s[5] - s[2] * s[3]);
  v.dJ[4] = 2.0 * (s[3] * s[5] - s[1] * s[4]);
  v.dJ[5] = 2.0 * (s[3] * s[4] - s[0] * s[5]);
  for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) v.K[i][j] = 0.0;
  auto set = [&](int i, int j, double x) { v.K[i][j] = x; v.K[j][i] = x; };
  set(0, 1, s[2]); set(0, 2, s[1]); set(1, 2, s[0]);
  set(0, 5, -2.0 * s[5]); set(1, 4, -2.0 * s[4]); set(2, 3, -2.0 * s[3]);
  set(3, 3, -2.0 * s[2]); set(4, 4, -2.0 * s[1]); set(5, 5, -2.0 * s[0]);
  set(3, 4, 2.0 * s[5]); set(3, 5, 2.0 * s[4]); set(4, 5, 2.0 * s[3]);
  for (int i = 0; i < 3; ++i) v.dI[i] = 2.0 * s[i];
  for (int i = 3; i < 6; ++i) v.dI[i] = 4.0 * s[i];
  return v;
}
static const double kSym6[6] = {1, 1, 1, 2, 2, 2};

// psi = a(I2) + b(J) + quadratic(S - I) + la_tr/2 (tr - 3)^2
struct Coef { double a, a1, a2, b, b1, b2, q, ltr; };
inline Coef coefs(const Material& m, double I2, double J, int d) {
  Coef c{0, 0, 0, 0, 0, 0, 0, 0};
  const double mu = m.mu, la = m.la;
  switch (m.model) {
    case M_SNH:
      c.a = 0.5 * mu * (I2 - d); c.a1 = 0.5 * mu;
      c.b = -mu * (J - 1.0) + 0.5 * la * (J - 1.0) * (J - 1.0); c.b1 = -mu + la * (J - 1.0); c.b2 = la; break;
    case M_NH: {
      const double lj = std::log(J);
      c.a = 0.5 * mu * (I2 - d); c.a1 = 0.5 * mu;
      c.b = -mu * lj + 0.5 * la * lj * lj; c.b1 = (-mu + la * lj) / J; c.b2 = (la + mu - la * lj) / (J * J); break; }
    case M_COROT: c.q = mu; c.ltr = la; break;
    case M_FCR:
      c.q = mu; c.b = 0.5 * la * (J - 1.0) * (J - 1.0); c.b1 = la * (J - 1.0); c.b2 = la; break;
    case M_ARAP: c.q = 0.5 * mu; break;
    case M_FUNG: {
      const double ex = std::exp(0.5 * kFungC * (I2 - d));
      const double k = (mu + kFungC * mu) / la;
      c.a = 0.5 * mu * (I2 - d) + 0.5 * mu * (ex - 1.0);
      c.a1 = 0.5 * mu + 0.25 * mu * kFungC * ex; c.a2 = 0.125 * mu * kFungC * kFungC * ex;
      c.b = 0.5 * la * (J - 1.0 - k) * (J - 1.0 - k); c.b1 = la * (J - 1.0 - k); c.b2 = la; break; }
  }
  return c;
}

inline double energy6(const Material& m, const double* s) {
  Inv6 v = invariants6(s);
  Coef c = coefs(m, v.I2, v.J, 3);
  double quad = 0;
  for (int i = 0; i < 3; ++i) quad += (s[i] - 1.0) * (s[i] - 1.0);
  for (int i = 3; i < 6; ++i) quad += 2.0 * s[i] * s[i];
  return c.a + c.b + c.q * quad + 0.5 * c.ltr * (v.tr - 3.0) * (v.tr - 3.0);
}
inline void gradient6(const Material& m, const double* s, double* g) {
  Inv6 v = invariants6(s);
  Coef c = coefs(m, v.I2, v.J, 3);
  for (int i = 0; i < 6; ++i) {
    double dq = (i < 3) ? 2.0 * (s[i] - 1.0) : 4.0 * s[i];
    double dtr = (i < 3) ? c.ltr * (v.tr - 3.0) : 0.0;
    g[i] = c.a1 * v.dI[i] + c.b1 * v.dJ[i] + c.q * dq + dtr;
  }
}
// H is 6x6 column-major (symmetric except the documented NH typo entry)
inline void hessian6(const Material& m, const double* s, double* H) {
  Inv6 v = invariants6(s);
  Coef c = coefs(m, v.I2, v.J, 3);
  for (int i = 0; i < 6; ++i)
    for (int j = 0; j < 6; ++j) {
      double h = c.a2 * v.dI[i] * v.dI[j] + c.b2 * v.dJ[i] * v.dJ[j] + c.b1 * v.K[i][j];
      if (i == j) h += 2.0 * c.a1 * kSym6[i] + 2.0 * c.q * kSym6[i];
      if (i < 3 && j < 3) h += c.ltr;
      H[j * 6 + i] = h;
    }
  if (m.model == M_NH) {
    // neohookean.cpp:56 : the la*dJ2*dJ0/J^2 term of H(0,2) is written with (S2*3 - S6*S6) instead of (S2*S3 - S6*S6)
    const double J2 = v.J * v.J;
    H[2 * 6 + 0] += m.la * v.dJ[2] * ((s[1] * 3.0 - s[5] * s[5]) - v.dJ[0]) / J2;
  }
}

// ---- 2D (3-vector) forms: only SNH, Corot, FixedCorot, ARAP have them in the reference ----
inline bool has_form3(int model) { return model == M_SNH || model == M_COROT || model == M_FCR || model == M_ARAP; }
struct Inv3 { double I2, J, tr, dJ[3], dI[3]; };
inline Inv3 invariants3(const double* s) {
  Inv3 v;
  v.tr = s[0] + s[1];
  v.I2 = s[0] * s[0] + s[1] * s[1] + 2.0 * s[2] * s[2];
  v.J = s[0] * s[1] - s[2] * s[2];
  v.dJ[0] = s[1]; v.dJ[1] = s[0]; v.dJ[2] = -2.0 * s[2];
  v.dI[0] = 2.0 * s[0]; v.dI[1] = 2.0 * s[1]; v.dI[2] = 4.0 * s[2];
  return v;
}
static const double kSym3[3] = {1, 1, 2};
inline double energy3(const Material& m, const double* s) {
  Inv3 v = invariants3(s);
  Coef c = coefs(m, v.I2, v.J, 2);
  double quad = (s[0] - 1.0) * (s[0] - 1.0) + (s[1] - 1.0) * (s[1] - 1.0) + 2.0 * s[2] * s[2];
  return c.a + c.b + c.q * quad + 0.5 * c.ltr * (v.tr - 2.0) * (v.tr - 2.0);
}
inline void gradient3(const Material& m, const double* s, double* g) {
  Inv3 v = invariants3(s);
  Coef c = coefs(m, v.I2, v.J, 2);
  for (int i = 0; i < 3; ++i) {
    double dq = (i < 2) ? 2.0 * (s[i] - 1.0) : 4.0 * s[i];
    double dtr = (i < 2) ? c.ltr * (v.tr - 2.0) : 0.0;
    g[i] = c.a1 * v.dI[i] + c.b1 * v.dJ[i] + c.q * dq + dtr;
  }
}
inline void hessian3(const Material& m, const double* s, double* H) {
  Inv3 v = invariants3(s);
  Coef c = coefs(m, v.I2, v.J, 2);
  double K[3][3] = {{0, 1, 0}, {1, 0, 0}, {0, 0, -2}};
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) {
      double h = c.a2 * v.dI[i] * v.dI[j] + c.b2 * v.dJ[i] * v.dJ[j] + c.b1 * K[i][j];
      if (i == j) h += 2.0 * c.a1 * kSym3[i] + 2.0 * c.q * kSym3[i];
      if (i < 2 && j < 2) h += c.ltr;
      H[j * 3 + i] = h;
    }
}

}  // namespace orc
