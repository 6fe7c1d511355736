// Host-side C++ mirror of mfem::MixedStretch<DIM> (src/variables/mixed_stretch.h:12-164) over the C ABI of
// include/mfem_b200.h.  Same method names, argument meaning and ownership as the reference class so that it can be
// registered with MixedVariableFactory under "mixed-stretch" (src/factories/variable_factory.cpp:13-19); see
// INTEGRATION.md for the Eigen-typed adapter a maintainer would add inside the reference tree.
//
// Eigen is not required: vectors are std::vector<double> with the reference layouts (x[d*v+l], N*e+k).  The members
// the reference's callers mutate in place (value(), delta(), lambda(): src/linesearch.h:120-125,
// src/linear_solvers/linear_system.h:97,189) are host mirrors; they are pushed to the device before every call that
// reads them and pulled back after every call that writes them, exactly the host<->device traffic a drop-in costs.
// No CPU fallback: every method forwards to the CUDA library and throws std::runtime_error on failure.
#pragma once
#include <cstdint>
#include <stdexcept>
#include <string>
#include <vector>
#include "../../include/mfem_b200.h"

namespace mfb {

struct MeshView {               // what MixedStretch pulls from mfem::Mesh (src/mesh/mesh.h:82-105,151-171)
  int dim = 3;
  int nV = 0, nE = 0;
  const int32_t* T = nullptr;        // nE x (dim+1) row-major
  const double* Vref = nullptr;      // nV x dim row-major
  const int32_t* free_map = nullptr; // nV
  const int32_t* mat_id = nullptr;   // nE or null
  std::vector<mfb_material> materials;
  int device = 0;
  // TriMesh (src/mesh/tri_mesh.cpp): triangles in 3D, dim = 3, T is nE x 3, per-face rest normals and a thickness
  bool shell = false;
  const double* normals = nullptr;   // nE x 3 row-major
  double thickness = 1e-3;           // MaterialConfig::thickness (config.h:150)
};

template <int DIM>
class MixedStretch {
 public:
  static std::string name() { return "mixed-stretch"; }
  static constexpr int N() { return DIM == 3 ? 6 : 3; }
  static constexpr int M() { return DIM * DIM; }

  explicit MixedStretch(const MeshView& mesh) {
    if (mesh.dim != DIM) throw std::runtime_error("MixedStretch: mesh dimension mismatch");
    mfb_mesh_desc d{};
    if (mesh.shell && DIM != 3) throw std::runtime_error("MixedStretch: a TriMesh shell is a DIM = 3 variable");
    d.elem_kind = mesh.shell ? MFB_SHELL : (DIM == 3 ? MFB_TET : MFB_TRI2D);
    d.normals = mesh.normals; d.thickness = mesh.thickness;
    d.nV = mesh.nV; d.nE = mesh.nE; d.T = mesh.T; d.Vref = mesh.Vref; d.free_map = mesh.free_map; d.mat_id = mesh.mat_id;
    d.n_mat = (int32_t)mesh.materials.size(); d.materials = mesh.materials.data(); d.device = mesh.device;
    check(mfb_create(&ctx_, &d));
    int32_t nF = 0; int64_t nnz = 0;
    mfb_dims(ctx_, nullptr, nullptr, &nelem_, &nF, nullptr, &nnz);
    nfree_ = nF; nnz_ = nnz;
  }
  ~MixedStretch() { mfb_destroy(ctx_); }
  MixedStretch(const MixedStretch&) = delete;
  MixedStretch& operator=(const MixedStretch&) = delete;

  // --- Variable / MixedVariable virtuals -------------------------------------------------------------------
  double energy(const std::vector<double>& x, const std::vector<double>& s) {
    double e = 0; check(mfb_energy(ctx_, x.data(), s.data(), &e)); return e;
  }
  double constraint_value(const std::vector<double>& x, const std::vector<double>& s) {
    push_state(); double e = 0; check(mfb_constraint_value(ctx_, x.data(), s.data(), &e)); return e;
  }
  void update(const std::vector<double>& x, double dt) {
    push_state();
    check(mfb_update(ctx_, x.data(), dt));
    lhs_valid_ = false;
  }
  void reset() {
    check(mfb_reset(ctx_, nullptr));
    s_.resize((size_t)N() * nelem_); ds_.assign((size_t)N() * nelem_, 0.0); la_.assign((size_t)N() * nelem_, 0.0);
    check(mfb_get(ctx_, MFB_F_S, s_.data()));
    s_dirty_ = false; la_dirty_ = false;
  }
  void pre_solve() {}
  void post_solve() { la_.assign(la_.size(), 0.0); check(mfb_post_solve(ctx_)); la_dirty_ = false; }   // mixed_stretch.cpp:340-343
  std::vector<double> rhs() {
    push_state();
    std::vector<double> r((size_t)DIM * nfree_); check(mfb_rhs(ctx_, r.data())); return r;
  }
  std::vector<double> gradient() { return get(MFB_F_GRAD_X); }
  std::vector<double> gradient_mixed() { return get(MFB_F_GRAD); }
  std::vector<double> gradient_dual() { return get(MFB_F_GRAD_LA); }
  void solve(const std::vector<double>& dx) {
    push_state();
    check(mfb_solve(ctx_, dx.data()));
    check(mfb_get(ctx_, MFB_F_DS, ds_.data()));
    check(mfb_get(ctx_, MFB_F_LA, la_.data()));
    la_dirty_ = false;
  }
  // The reference hands these out by non-const reference and its callers write through them (linesearch.h:120-125,
  // linear_system.h:97,189).  A non-const access marks the host copy as newer than the device's; the next call that reads
  // the state uploads it once (no O(nE) comparison per call).  const access does not.
  std::vector<double>& delta() { return ds_; }
  std::vector<double>& value() { s_dirty_ = true; return s_; }
  std::vector<double>& lambda() { la_dirty_ = true; return la_; }
  const std::vector<double>& value() const { return s_; }
  const std::vector<double>& lambda() const { return la_; }
  int size() const { return N() * nelem_; }
  int size_dual() const { return N() * nelem_; }
  std::vector<double> max_stresses() { return get(MFB_F_EVALS); }
  void evaluate_constraint(const std::vector<double>& x, std::vector<double>& c) {
    push_state(); c.resize((size_t)N() * nelem_); check(mfb_evaluate_constraint(ctx_, x.data(), c.data()));
  }
  // lhs(): RowMajor CSR (outer, inner, values) of A_ ; pattern pointers stay valid for the object's lifetime
  struct Csr { const int32_t* outer; const int32_t* inner; const double* values; int64_t nnz; int32_t rows; };
  Csr lhs() {
    Csr m{};
    check(mfb_csr_pattern(ctx_, &m.outer, &m.inner, &m.nnz, &m.rows));
    if (!lhs_valid_) { lhs_vals_.resize((size_t)nnz_); check(mfb_lhs_values(ctx_, lhs_vals_.data())); lhs_valid_ = true; }
    m.values = lhs_vals_.data();
    return m;
  }
  void product_hessian(const std::vector<double>& x, std::vector<double>& out) { check(mfb_product(ctx_, 0, x.data(), out.data(), 0)); }
  void product_hessian_inv(const std::vector<double>& x, std::vector<double>& out) { check(mfb_product(ctx_, 1, x.data(), out.data(), 0)); }
  void product_jacobian_mixed(const std::vector<double>& x, std::vector<double>& out) { check(mfb_product(ctx_, 2, x.data(), out.data(), 0)); }
  void product_jacobian_x(const std::vector<double>& x, std::vector<double>& out, bool transposed) {
    check(mfb_product(ctx_, 3, x.data(), out.data(), transposed ? 1 : 0));
  }
  // hessian()/hessian_inv()/jacobian_mixed() block diagonals (mixed_stretch.cpp:368-404) as per-element N x N blocks
  std::vector<double> hessian_blocks() { return get(MFB_F_HLOC); }
  std::vector<double> hessian_inv_blocks() {
    std::vector<double> h = get(MFB_F_HINV), v = get(MFB_F_VOLS);
    for (int e = 0; e < nelem_; ++e) for (int k = 0; k < N() * N(); ++k) h[(size_t)e * N() * N() + k] *= v[e];
    return h;
  }
  std::vector<double> dSdF_blocks() { return get(MFB_F_DSDF); }
  // jacobian_x(): -mesh_->jacobian() * blockdiag(dSdF_) (mixed_stretch.cpp:385-390) as triplets (pinned rows: row -1)
  struct Triplets { std::vector<int32_t> rows, cols; std::vector<double> vals; int32_t nrows, ncols; };
  Triplets jacobian_x() {
    Triplets t;
    const size_t n = (size_t)mfb_jacobian_x_nnz(ctx_);
    t.rows.resize(n); t.cols.resize(n); t.vals.resize(n);
    check(mfb_jacobian_x_coo(ctx_, t.rows.data(), t.cols.data(), t.vals.data()));
    t.nrows = DIM * nfree_; t.ncols = N() * nelem_;
    return t;
  }

  mfb_ctx* ctx() { return ctx_; }
  mfb_status status() { mfb_status st{}; check(mfb_last_status(ctx_, &st)); return st; }

 private:
  void check(int rc) const { if (rc != MFB_OK) throw std::runtime_error(std::string("mfem_b200: ") + mfb_last_error_string()); }
  std::vector<double> get(int f) {
    std::vector<double> v((size_t)mfb_field_size(ctx_, f)); check(mfb_get(ctx_, f, v.data())); return v;
  }
  // value()/lambda() were handed out by non-const reference since the last upload: push them to the device
  void push_state() {
    if (s_dirty_) { check(mfb_set(ctx_, MFB_F_S, s_.data())); s_dirty_ = false; }
    if (la_dirty_) { check(mfb_set(ctx_, MFB_F_LA, la_.data())); la_dirty_ = false; }
  }
  mfb_ctx* ctx_ = nullptr;
  int32_t nelem_ = 0, nfree_ = 0;
  int64_t nnz_ = 0;
  std::vector<double> s_, ds_, la_, lhs_vals_;
  bool lhs_valid_ = false, s_dirty_ = false, la_dirty_ = false;
};

// The linear solve of the condensed system as the reference's solver factory names it (linear_solver_factory.cpp:107-173):
// lhs = M + A on the Assembler pattern, Eigen's ConjugateGradient recurrence, device vectors.
struct PcgSolver {
  int precond = 1;          // "eigen-pcg-diag" ; 2 = "eigen-pcg-block_jacobi" ; 0 = identity
  double tol = 1e-4;        // SimConfig::itr_tol
  int max_iters = 500;      // SimConfig::max_iterative_solver_iters
  static PcgSolver by_name(const std::string& n) {
    PcgSolver s;
    if (n == "eigen-pcg-diag") s.precond = 1;
    else if (n == "eigen-pcg-block_jacobi") s.precond = 2;
    else if (n == "eigen-pcg-identity") s.precond = 0;
    else if (n == "mfb-pcg-chebyshev") s.precond = 3;          // not a reference solver name: Chebyshev over block-Jacobi (mfb_pcg_dev)
    else throw std::runtime_error("PcgSolver: no device implementation of solver '" + n + "'");
    return s;
  }
};

// NewtonOptimizer<DIM>::step (src/optimizers/newton_optimizer.cpp:14-139) for one mesh, the mixed-stretch variable and no
// contact, resident on the device: Displacement + BDF1 + SystemMatrixPD + line search (include/mfem_b200.h, mfb_newton_*).
template <int DIM>
class NewtonStep {
 public:
  NewtonStep(MixedStretch<DIM>& var, const MeshView& mesh, double density, double h, const double* force = nullptr,
             const double* x0 = nullptr, const double* v0 = nullptr) : var_(var), nx_((size_t)DIM * mesh.nV) {
    check(mfb_newton_init(var_.ctx(), mesh.free_map, density, h, force, x0, v0));
  }
  // one density per ELEMENT: the density of the material each element's OBJECT was created with (meshes.cpp:100-108)
  NewtonStep(MixedStretch<DIM>& var, const MeshView& mesh, const std::vector<double>& density_e, double h, const double* force = nullptr,
             const double* x0 = nullptr, const double* v0 = nullptr) : var_(var), nx_((size_t)DIM * mesh.nV) {
    if ((int)density_e.size() != mesh.nE) throw std::runtime_error("NewtonStep: one density per element expected");
    check(mfb_newton_init_density(var_.ctx(), mesh.free_map, density_e.data(), h, force, x0, v0));
  }
  ~NewtonStep() { mfb_newton_free(var_.ctx()); }
  // one time step; returns the OptimizerData columns (newton_optimizer.cpp:92-97)
  mfb_newton_log step(const PcgSolver& solver = PcgSolver(), int outer_steps = 5, int ls_iters = 20, double newton_tol = 1e-10) {
    mfb_newton_opts o{};
    o.outer_steps = outer_steps; o.ls_iters = ls_iters; o.newton_tol = newton_tol;
    o.precond = solver.precond; o.cg_tol = solver.tol; o.cg_max_iters = solver.max_iters;
    mfb_newton_log log{};
    check(mfb_newton_step(var_.ctx(), &o, &log));
    return log;
  }
  std::vector<double> positions() { std::vector<double> x(nx_); check(mfb_newton_positions(var_.ctx(), x.data())); return x; }

 private:
  void check(int rc) const { if (rc != MFB_OK) throw std::runtime_error(std::string("mfem_b200: ") + mfb_last_error_string()); }
  MixedStretch<DIM>& var_;
  size_t nx_;
};

// The two saddle-point systems of src/linear_solvers/linear_system.h on the device (include/mfem_b200.h, mfb_kkt_* /
// mfb_dual_*).  Same member names as the reference's system classes; vectors in the reference's ordering.  They read
// the Displacement / BDF1 state, so a NewtonStep on the same variable must exist.
template <int DIM>
class SystemMatrixIndefinite {          // linear_system.h:64-118 + BlockMatrix product, block_matrix.h:74-125
 public:
  explicit SystemMatrixIndefinite(MixedStretch<DIM>& var) : var_(var) {}
  size_t rows() const { return (size_t)mfb_kkt_size(var_.ctx()); }
  std::vector<double> b() { std::vector<double> r(rows()); check(mfb_kkt_rhs(var_.ctx(), r.data())); return r; }
  std::vector<double> product(const std::vector<double>& v) { std::vector<double> o(rows()); check(mfb_kkt_product(var_.ctx(), v.data(), o.data())); return o; }
  // "minres-indefinite-*" without the factorisation-based preconditioners: pre_solve + MINRES + post_solve
  std::vector<double> solve(double tol, int max_iters, int* iters = nullptr, double* error = nullptr) {
    std::vector<double> x(rows(), 0.0);
    check(mfb_kkt_solve_minres(var_.ctx(), x.data(), tol, max_iters, iters, error));
    return x;
  }
 private:
  void check(int rc) const { if (rc != MFB_OK) throw std::runtime_error(std::string("mfem_b200: ") + mfb_last_error_string()); }
  MixedStretch<DIM>& var_;
};

template <int DIM>
class DualCondensedSystem {             // linear_system.h:120-239, lhs applied matrix-free
 public:
  explicit DualCondensedSystem(MixedStretch<DIM>& var, size_t n_stretch) : var_(var), n_(n_stretch) {}
  std::vector<double> b() { std::vector<double> r(n_); check(mfb_dual_rhs(var_.ctx(), r.data())); return r; }
  std::vector<double> product(const std::vector<double>& la) { std::vector<double> o(n_); check(mfb_dual_product(var_.ctx(), la.data(), o.data())); return o; }
  // "subspace-CG" without the Gauss-Seidel preconditioner: pre_solve + CG + post_solve; returns lambda()
  std::vector<double> solve(double tol, int max_iters, int* iters = nullptr, double* error = nullptr) {
    std::vector<double> la(n_, 0.0);
    check(mfb_dual_solve_cg(var_.ctx(), la.data(), tol, max_iters, iters, error, nullptr, nullptr));
    return la;
  }
 private:
  void check(int rc) const { if (rc != MFB_OK) throw std::runtime_error(std::string("mfem_b200: ") + mfb_last_error_string()); }
  MixedStretch<DIM>& var_;
  size_t n_;
};

}  // namespace mfb
