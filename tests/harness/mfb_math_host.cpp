// TEST-ONLY harness: compiles mixed-fem_b200/csrc/mfb_math.cuh as plain C++ so the per-element algebra of the CUDA
// kernels can be checked against the oracle on the CPU box (no GPU here).  Not part of the product library.
#include "../../mixed-fem_b200/csrc/mfb_math.cuh"
#include <cstring>
using namespace mfb;
extern "C" {
void hh_material6(int model, double mu, double la, const double* s, double scale, double* g, double* Hfull, double* e) {
  double H[21];
  material_grad_hess6(model, mu, la, s, scale, g, H);
  for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) Hfull[j * 6 + i] = H[symidx(i, j)];
  *e = material_energy6(model, mu, la, s);
}
void hh_material3(int model, double mu, double la, const double* s, double scale, double* g, double* Hfull, double* e) {
  double H[6];
  material_grad_hess3(model, mu, la, s, scale, g, H);
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) Hfull[j * 3 + i] = H[symidx(i, j)];
  *e = material_energy3(model, mu, la, s);
}
int hh_psd6(double* Hfull, double* Hinvfull, double tol, double* lmax) {
  double H[21], Hi[21]; int sw;
  for (int i = 0; i < 6; ++i) for (int j = 0; j <= i; ++j) H[symidx(i, j)] = Hfull[j * 6 + i];
  *lmax = psd_fix_invert<6>(H, Hi, tol, &sw);
  for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) { Hfull[j * 6 + i] = H[symidx(i, j)]; Hinvfull[j * 6 + i] = Hi[symidx(i, j)]; }
  return sw;
}
// fast path: returns 1 when accepted; Hfull in (col-major), outputs Lfull (lower, col-major), Hinvfull, lmax, iters
int hh_psd_fast6(const double* Hfull, double* Lfull, double* Hinvfull, double tol, double relpiv, double* lmax, int* iters) {
  double H[21], L[21], Hi[21];
  for (int i = 0; i < 6; ++i) for (int j = 0; j <= i; ++j) H[symidx(i, j)] = Hfull[j * 6 + i];
  const bool ok = psd_fast6(H, L, Hi, tol, relpiv, lmax, iters);
  for (int i = 0; i < 6; ++i) for (int j = 0; j < 6; ++j) { Lfull[j * 6 + i] = (i >= j) ? L[symidx(i, j)] : 0.0; Hinvfull[j * 6 + i] = Hi[symidx(i, j)]; }
  return ok ? 1 : 0;
}
int hh_psd3(double* Hfull, double* Hinvfull, double tol, double* lmax) {
  double H[6], Hi[6]; int sw;
  for (int i = 0; i < 3; ++i) for (int j = 0; j <= i; ++j) H[symidx(i, j)] = Hfull[j * 3 + i];
  *lmax = psd_fix_invert<3>(H, Hi, tol, &sw);
  for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) { Hfull[j * 3 + i] = H[symidx(i, j)]; Hinvfull[j * 3 + i] = Hi[symidx(i, j)]; }
  return sw;
}
// R in/out (col-major), s6 out, Js out as 6x9 column-major (Js[c*6+n])
int hh_polar3(double* R, const double* F, double* s6, double* Js_6x9, int* status) {
  double Y[9], Hinv3[6];
  *status = 0;
  int it = procrustes3(R, F, Y, 1e-6, 100, status);
  stretch6(R, F, s6);
  if (!procrustes_hinv(Y, Hinv3)) *status |= ST_POLAR_HFAIL;
  for (int n = 0; n < 6; ++n) { double col[9]; dsdf3_column(n, R, F, Hinv3, col); for (int k = 0; k < 9; ++k) Js_6x9[k * 6 + n] = col[k]; }
  return it;
}
// number of entries where the run-time-index variant of the ds/dF column differs from the unrolled one (must be 0)
int hh_dsdf_rt_mismatches(const double* R, const double* F, const double* Hinv3) {
  int bad = 0;
  for (int n = 0; n < 6; ++n) {
    double a[9], b[9];
    dsdf3_column(n, R, F, Hinv3, a);
    dsdf3_column_rt(n, R, F, Hinv3, b);
    for (int k = 0; k < 9; ++k) bad += !(a[k] == b[k]);
  }
  return bad;
}
void hh_polar2(const double* F, double* R, double* s3, double* Js_3x4) {
  double Js[3][4];
  polar2(F, R, s3, Js);
  for (int n = 0; n < 3; ++n) for (int k = 0; k < 4; ++k) Js_3x4[k * 3 + n] = Js[n][k];
}
}
