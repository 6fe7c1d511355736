// Runs the host C++ mirror (mixed-fem_b200/host/mixed_stretch.hpp) on a GPU, call for call as the reference drives a
// MixedVariable<3> (src/variables/mixed_variable.h:11-69): reset / value() / lambda() / update / rhs / lhs / solve /
// delta / energy / constraint_value / gradient*, then NewtonStep::step.  Input and output are raw binary files written /
// read by tests/test_gpu_cpp_shim.py, which compares every vector with the oracle.
//   in : int32 nV, nE ; int32 T[nE*4] ; double V[nV*3] ; int32 free_map[nV] ; double mu, la, dt ;
//        double x[3 nV] ; double s[6 nE] ; double lam[6 nE] ; int32 nF ; double dx[3 nF]
//   out: sequence of (int64 n ; double data[n]) records, in the order they are written below
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "mixed-fem_b200/host/mixed_stretch.hpp"

template <class T>
static std::vector<T> rd(FILE* f, size_t n) {
  std::vector<T> v(n);
  if (n && fread(v.data(), sizeof(T), n, f) != n) { fprintf(stderr, "short read\n"); exit(2); }
  return v;
}
static void wr(FILE* f, const std::vector<double>& v) {
  const long long n = (long long)v.size();
  fwrite(&n, sizeof(n), 1, f);
  if (n) fwrite(v.data(), sizeof(double), (size_t)n, f);
}

int main(int argc, char** argv) {
  if (argc < 3) { fprintf(stderr, "usage: shim_run in.bin out.bin\n"); return 2; }
  FILE* fi = fopen(argv[1], "rb");
  FILE* fo = fopen(argv[2], "wb");
  if (!fi || !fo) { fprintf(stderr, "cannot open files\n"); return 2; }
  const auto hdr = rd<int32_t>(fi, 2);
  const int nV = hdr[0], nE = hdr[1];
  const auto T = rd<int32_t>(fi, (size_t)nE * 4);
  const auto V = rd<double>(fi, (size_t)nV * 3);
  const auto fm = rd<int32_t>(fi, (size_t)nV);
  const auto par = rd<double>(fi, 3);
  const auto x = rd<double>(fi, (size_t)nV * 3);
  const auto s0 = rd<double>(fi, (size_t)nE * 6);
  const auto la0 = rd<double>(fi, (size_t)nE * 6);
  const int nF = rd<int32_t>(fi, 1)[0];
  const auto dx = rd<double>(fi, (size_t)nF * 3);
  fclose(fi);
  try {
    mfb::MeshView m;
    m.dim = 3; m.nV = nV; m.nE = nE; m.T = T.data(); m.Vref = V.data(); m.free_map = fm.data();
    m.materials.push_back({MFB_SNH, par[0], par[1], 1000.0});
    mfb::MixedStretch<3> var(m);
    var.reset();
    wr(fo, var.value());                       // 0: s_ after reset()
    var.value() = s0;                          // the callers write through the references (linesearch.h:120-125)
    var.lambda() = la0;
    var.update(x, par[2]);
    wr(fo, var.rhs());                         // 1
    const auto A = var.lhs();
    wr(fo, std::vector<double>(A.values, A.values + A.nnz));   // 2
    wr(fo, var.gradient());                    // 3
    wr(fo, var.gradient_mixed());              // 4
    wr(fo, var.gradient_dual());               // 5
    var.solve(dx);
    wr(fo, var.delta());                       // 6
    wr(fo, var.lambda());                      // 7
    std::vector<double> st = var.value();
    for (size_t i = 0; i < st.size(); ++i) st[i] += 0.5 * var.delta()[i];
    wr(fo, {var.energy(x, st), var.constraint_value(x, st)});   // 8
    std::vector<double> c;
    var.evaluate_constraint(x, c);
    wr(fo, c);                                 // 9
    wr(fo, var.max_stresses());                // 10
    var.post_solve();
    wr(fo, var.lambda());                      // 11: zeros
    // one device-resident time step from the rest state (NewtonOptimizer::step)
    var.reset();
    mfb::NewtonStep<3> nt(var, m, 1000.0, 0.02);
    mfb::PcgSolver solver = mfb::PcgSolver::by_name("eigen-pcg-block_jacobi");
    solver.tol = 1e-13; solver.max_iters = 5000;
    mfb_set_option(var.ctx(), "cg_zero_guess", 1);
    const mfb_newton_log lg = nt.step(solver, 3, 20, 1e-10);
    std::vector<double> hist;
    hist.push_back((double)lg.n_iters);
    for (int i = 0; i < lg.n_iters; ++i) { hist.push_back(lg.energy[i]); hist.push_back(lg.decrement[i]); hist.push_back(lg.alpha[i]); }
    wr(fo, hist);                              // 12
    wr(fo, nt.positions());                    // 13
    const mfb_status status = var.status();
    wr(fo, {(double)status.nonfinite, (double)status.polar_maxiter});   // 14
  } catch (const std::exception& e) {
    fprintf(stderr, "shim_run: %s\n", e.what());
    return 1;
  }
  fclose(fo);
  return 0;
}
