/* mfem_b200.h — C ABI of the B200-native mixed-stretch hot path (drop-in for mixed-fem's MixedStretch<DIM>).
 *
 * Every entry point cites the reference interface it replaces (paths relative to evouga/mixed-fem).
 * Conventions: return 0 on success, a negative mfb_error otherwise (never exit()); the caller owns every input
 * buffer; the context owns all device memory; one context is used from one host thread; calls are ordered on the
 * context's CUDA stream and are synchronous at the ABI unless the name ends in _dev (device pointers, asynchronous
 * on the stream).  All reals are FP64, all indices int32 (Eigen's default StorageIndex).
 *
 * Host-side layouts are the reference's: x[d*v+l]; per-element vectors N*e+k (N = 6 in 3D, 3 in 2D);
 * R, dSdF, H, Hinv column-major per element (src/variables/mixed_stretch.h:150-159).
 */
#ifndef MFEM_B200_H
#define MFEM_B200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct mfb_ctx mfb_ctx;

enum mfb_error {
  MFB_OK = 0,
  MFB_ERR_INVALID = -1,      /* bad argument */
  MFB_ERR_CUDA = -2,         /* CUDA runtime failure (mfb_last_error_string has the text) */
  MFB_ERR_NO_DEVICE = -3,    /* no CUDA device: there is NO CPU fallback */
  MFB_ERR_TOO_LARGE = -4,    /* nnz does not fit int32 */
  MFB_ERR_UNSUPPORTED = -5,  /* e.g. Neohookean/Fung on a 2D mesh (the reference has no 3-vector form either) */
  MFB_ERR_STATE = -6         /* call order (e.g. solve() before rhs()) */
};

/* src/factories/material_model_factory.cpp:11-47 names: "Stable-Neohookean","Neohookean","Corot","FixedCorot","ARAP","Fung" */
enum mfb_model { MFB_SNH = 0, MFB_NH = 1, MFB_COROT = 2, MFB_FCR = 3, MFB_ARAP = 4, MFB_FUNG = 5 };
enum mfb_elem { MFB_TET = 0, MFB_TRI2D = 1, MFB_SHELL = 2 };   /* TetrahedralMesh, Tri2DMesh, TriMesh (src/mesh/*.h) */

typedef struct {          /* src/config.h:143-151 MaterialConfig (the fields the path reads) */
  int32_t model;
  double mu, la;
  double density;         /* MaterialConfig::density (config.h:149); <= 0 means the default 1000.  Read by mfb_newton_init only */
} mfb_material;

typedef struct {          /* what MixedStretch pulls from Mesh: src/mesh/mesh.h:82-105,151-171 */
  int32_t elem_kind;      /* MFB_TET (TetrahedralMesh), MFB_TRI2D (Tri2DMesh) or MFB_SHELL (TriMesh: T is nE x 3, Vref is nV x 3) */
  int32_t nV, nE;
  const int32_t* T;       /* nE x (d+1) row-major vertex ids      (Mesh::T_) */
  const double* Vref;     /* nV x d row-major reference positions (Mesh::Vref_) */
  const int32_t* free_map;/* nV, -1 = pinned, else ascending free id (Mesh::free_map_, mesh.cpp:95-101) */
  const int32_t* mat_id;  /* nE or NULL (all 0)                   (Mesh::elements_[i].material_) */
  int32_t n_mat;
  const mfb_material* materials;
  int32_t device;         /* CUDA device ordinal */
  const double* normals;  /* MFB_SHELL only: nE x 3 row-major per-face rest normals (TriMesh::N_, tri_mesh.cpp:31-36) */
  double thickness;       /* MFB_SHELL only: MaterialConfig::thickness (volumes = doublearea*thickness/2, tri_mesh.cpp:38-41) */
} mfb_mesh_desc;

typedef struct {          /* SURVEY.md §5 failure detection: counts since the last update/reset */
  int64_t polar_maxiter;  /* elements whose procrustes hit max_iter (newton_procrustes.h:158) */
  int64_t polar_hfail;    /* elements with |H| < 1e-6, where the reference calls exit(1) (newton_procrustes.h:171-174) */
  int64_t nonfinite;      /* elements that produced a non-finite value */
  int64_t polar_iters;    /* total procrustes iterations */
  int64_t eig_sweeps;     /* total Jacobi sweeps */
  int64_t psd_slow;       /* elements whose Hessian needed the eigen-decomposition (fast Cholesky path refused) */
} mfb_status;

/* per-element / per-dof fields for mfb_get / mfb_set / mfb_field_dev (names = members of MixedStretch, mixed_stretch.h:141-162) */
enum mfb_field {
  MFB_F_S = 0,      /* s_      N*nE   (value())   settable */
  MFB_F_DS,         /* ds_     N*nE   (delta())   settable */
  MFB_F_LA,         /* la_     N*nE   (lambda())  settable */
  MFB_F_R,          /* R_      d*d*nE col-major   settable */
  MFB_F_SROT,       /* S_      N*nE */
  MFB_F_G,          /* g_      N*nE */
  MFB_F_H,          /* H_      N*N*nE */
  MFB_F_HINV,       /* Hinv_   N*N*nE */
  MFB_F_HLOC,       /* Hloc_   N*N*nE   (= vol^2 Sym H_ Sym) */
  MFB_F_DSDF,       /* dSdF_   M*N*nE col-major M x N */
  MFB_F_EVALS,      /* evals_  nE     (max_stresses()) */
  MFB_F_GL,         /* gl_     N*nE */
  MFB_F_GRAD,       /* grad_   N*nE   (gradient_mixed()) */
  MFB_F_GRAD_LA,    /* grad_la_ N*nE  (gradient_dual()) */
  MFB_F_GRAD_X,     /* grad_x_ d*nF   (gradient()) */
  MFB_F_RHS,        /* rhs_    d*nF */
  MFB_F_VOLS,       /* Mesh::volumes() nE */
  MFB_F_ITERS,      /* procrustes iterations of the last update, as doubles, nE */
  MFB_F_COUNT
};

/* ---- lifetime -------------------------------------------------------------------------------------------------
 * mfb_create: MixedStretch(mesh) ctor + the Mesh::init() products the variable reads (mesh.cpp:44-83: volumes,
 * dphi/dX of tet_mesh.cpp:43-75 / tri2d_mesh.cpp:44-77) + Assembler ctor (sparse_utils.cpp:19-117).  One-time
 * host->device upload happens here. */
int mfb_create(mfb_ctx** out, const mfb_mesh_desc* desc);
void mfb_destroy(mfb_ctx* ctx);
int mfb_set_stream(mfb_ctx* ctx, void* cuda_stream);          /* default: stream 0 */
const char* mfb_last_error_string(void);
const char* mfb_version(void);

/* ---- sizes ---- */
int mfb_dims(const mfb_ctx* ctx, int32_t* dim, int32_t* nV, int32_t* nE, int32_t* nF, int32_t* N, int64_t* nnz);

/* ---- Assembler products: Assembler::A pattern (sparse_utils.cpp:101-116) and the element->nonzero map implied by
 * update_matrix (sparse_utils.cpp:140-163).  Host pointers owned by ctx, valid until destroy.
 * map[e*(d+1)^2 + j*(d+1) + k] = value index of entry (0,0) of block (j,k) (-1 if either vertex is pinned);
 * entry (r,c) of that block lives at map + r*stride[e*(d+1)+j] + c. */
int mfb_csr_pattern(mfb_ctx* ctx, const int32_t** outer, const int32_t** inner, int64_t* nnz, int32_t* nrows);
int mfb_scatter_map(mfb_ctx* ctx, const int32_t** map, const int32_t** stride);

/* ---- MixedVariable<DIM> virtuals (src/variables/mixed_variable.h:11-69, mixed_stretch.h:25-83) ---- */
int mfb_reset(mfb_ctx* ctx, const double* x_or_null);          /* reset()  mixed_stretch.cpp:294-338 (NULL: x = vec(Vref)) */
int mfb_update(mfb_ctx* ctx, const double* x, double dt);      /* update() mixed_stretch.cpp:158-164 ; x is host, d*nV */
int mfb_update_dev(mfb_ctx* ctx, const double* x_dev, double dt);
int mfb_rhs(mfb_ctx* ctx, double* out);                        /* rhs()    mixed_stretch.cpp:245-275 ; out host d*nF */
int mfb_rhs_dev(mfb_ctx* ctx, const double** out_dev);
int mfb_solve(mfb_ctx* ctx, const double* dx);                 /* solve()  mixed_stretch.cpp:277-292 ; dx host d*nF */
int mfb_solve_dev(mfb_ctx* ctx, const double* dx_dev);
int mfb_post_solve(mfb_ctx* ctx);                              /* post_solve() mixed_stretch.cpp:340-343 : la_ = 0 */
int mfb_energy(mfb_ctx* ctx, const double* x, const double* s, double* e);            /* mixed_stretch.cpp:93-107 */
int mfb_constraint_value(mfb_ctx* ctx, const double* x, const double* s, double* e);  /* mixed_stretch.cpp:109-136 */
int mfb_energy_dev(mfb_ctx* ctx, const double* s_dev_soa_or_null, double* e);
int mfb_constraint_value_dev(mfb_ctx* ctx, const double* x_dev, const double* s_dev_soa_or_null, double* e);
int mfb_evaluate_constraint(mfb_ctx* ctx, const double* x, double* c);                /* mixed_stretch.cpp:345-366 */
int mfb_lhs_values(mfb_ctx* ctx, double* values);              /* lhs().valuePtr() mixed_stretch.h:37-39 ; host, nnz */
int mfb_lhs_values_dev(mfb_ctx* ctx, const double** values_dev);
/* the operator (A, or A + M for device M values on the same pattern, else NULL) as block-CSR with d x d ROW-MAJOR blocks:
 * ptr[nF+1], col[nb], val[nb*d*d] (host) - the layout of amgcl's crs<static_matrix<double,d,d>> that the reference's
 * AMGCL solver consumes (src/linear_solvers/amgcl_solver.h:40-75); nb = nnz / d^2 */
int mfb_lhs_bsr(mfb_ctx* ctx, const double* M_vals_dev, int32_t* ptr, int32_t* col, double* val);

/* matrix-free operators, out += ... / out -= ... exactly as mixed_stretch.cpp:406-468.  which: 0 product_hessian,
 * 1 product_hessian_inv, 2 product_jacobian_mixed, 3 product_jacobian_x (transposed selects the branch). host buffers. */
int mfb_product(mfb_ctx* ctx, int which, const double* x, double* out, int transposed);
/* jacobian_x() = -PJW_ * blockdiag(dSdF_) (mixed_stretch.cpp:385-390) as COO triplets for the indefinite / dual-condensed
 * systems (linear_system.h:64-239): mfb_jacobian_x_nnz() = nE*N*(verts per element)*d entries; row = d*free_id + r or -1
 * for a pinned vertex (value 0: drop it), col = N*e + n.  The block-diagonal hessian() / hessian_inv() /
 * jacobian_mixed() are mfb_get(MFB_F_HLOC), vol * MFB_F_HINV and vol * Sym per element. */
int64_t mfb_jacobian_x_nnz(mfb_ctx* ctx);
int mfb_jacobian_x_coo(mfb_ctx* ctx, int32_t* rows, int32_t* cols, double* vals);

/* ---- state access ---- */
int64_t mfb_field_size(const mfb_ctx* ctx, int field);
int mfb_get(mfb_ctx* ctx, int field, double* out);             /* host copy in the reference layout */
int mfb_set(mfb_ctx* ctx, int field, const double* in);        /* S, DS, LA, R only */
/* device pointer of the structure-of-arrays storage: component c of element e at ptr[c*ld + e] */
int mfb_field_dev(mfb_ctx* ctx, int field, double** ptr, int64_t* ld, int32_t* ncomp);
int mfb_last_status(mfb_ctx* ctx, mfb_status* st);

/* ---- linear-solver side (src/linear_solvers/linear_system.h:9-62 SystemMatrixPD, eigen_iterative_solver.h:27-65,
 * pcg.h:11-80).  All pointers are DEVICE pointers; the matrix is  M_vals (same CSR pattern, may be NULL) + A. ---- */
int mfb_spmv_dev(mfb_ctx* ctx, const double* M_vals_dev, const double* x_dev, double* y_dev);
/* precond: 0 none, 1 diagonal (Eigen DiagonalPreconditioner), 2 block-Jacobi d x d (preconditioners/block_jacobi.h:45-60),
 * 3 Chebyshev polynomial over block-Jacobi (NOT in the reference, whose stronger preconditioners are factorisations / AMG:
 *   linear_solver_factory.cpp:119-173): degree "cheb_degree" (option, default 6) on [lmax / "cheb_ratio" (120), lmax], lmax of
 *   Dinv L by 12 power iterations x 1.2, capped by the rigorous bound vertices-per-element.  Cuts CG iterations 4x at config C3
 *   (500 -> 137) but costs 1.5x the SpMVs: slower on one GPU (44.8 vs 31.0 ms per Newton iteration), meant for several.
 * Solves (M + A) x = b from the initial guess in x; stops when |r| <= tol*|b| (Eigen ConjugateGradient criterion). */
int mfb_pcg_dev(mfb_ctx* ctx, const double* M_vals_dev, const double* b_dev, double* x_dev, int precond,
                double tol, int max_iters, int* iters_out, double* relres_out);

/* ---- device-resident Newton time step around the variable.  Replaces, for a scene with ONE mesh,
 * the mixed-stretch variable and no contact:  NewtonOptimizer<DIM>::step (src/optimizers/newton_optimizer.cpp:14-139),
 * linesearch_backtracking (src/linesearch.h:80-129), Displacement<DIM> (src/variables/displacement.cpp:21-96),
 * BDF<1> (src/time_integrators/BDF.cpp:7-27), SystemMatrixPD (src/linear_solvers/linear_system.h:18-44) and the
 * ConjugateGradient call of EigenIterativeSolver::solve (src/linear_solvers/eigen_iterative_solver.h:27-65) with the
 * consistent mass matrix of deps/Bartels/src/linear_tet_mass_matrix.cpp:8-43 / linear_tri2d_mass_matrix.cpp:8-36. ---- */
#define MFB_NEWTON_LOG_MAX 64
typedef struct mfb_newton_opts {
  int32_t outer_steps;    /* SimConfig::outer_steps (config.h:112, default 5) */
  int32_t ls_iters;       /* SimConfig::ls_iters (config.h:113, default 20) */
  double newton_tol;      /* SimConfig::newton_tol (config.h:127, default 1e-10) */
  int32_t precond;        /* 0 none, 1 diagonal ("eigen-pcg-diag"), 2 block-Jacobi ("eigen-pcg-block_jacobi"), 3 Chebyshev (see mfb_pcg_dev) */
  double cg_tol;          /* SimConfig::itr_tol (config.h:131, default 1e-4) */
  int32_t cg_max_iters;   /* SimConfig::max_iterative_solver_iters (default 500) */
} mfb_newton_opts;
typedef struct mfb_newton_log {   /* the OptimizerData columns of newton_optimizer.cpp:92-97 per Newton iteration */
  int32_t n_iters;
  int64_t cg_iters_total;
  double energy[MFB_NEWTON_LOG_MAX], energy_res[MFB_NEWTON_LOG_MAX], decrement[MFB_NEWTON_LOG_MAX], alpha[MFB_NEWTON_LOG_MAX];
  double cg_error[MFB_NEWTON_LOG_MAX];
  int32_t cg_iters[MFB_NEWTON_LOG_MAX], ls_iters[MFB_NEWTON_LOG_MAX];
  /* option "convergence_log" = 1: what apps/convergence.cpp:21-51 records through Optimizer::callback at the TOP of every
   * Newton iteration (newton_optimizer.cpp:37) - |x.rhs() + Stretch.rhs()|_2 of the NON-mixed system (the Stretch<DIM>
   * variable with each element's F-form material) at the current mixed iterate; else untouched */
  double fem_gradient_norm[MFB_NEWTON_LOG_MAX];
} mfb_newton_log;
/* Displacement::reset + BDF1 state: free_map as given to mfb_create; h = SimConfig::h; force = body force d-vector
 * (NULL: gravity (0,-9.8,0)); x0 / v0 host d*nV or NULL (vec(Vref^T), zero).  Also builds PMP^T on the Assembler pattern.
 * density > 0: one MaterialConfig::density for the whole mesh; density <= 0: every element takes the density of ITS
 * material (mfb_material::density through mat_id).
 * mfb_newton_init_density: one density per ELEMENT (host, nE).  This is what reproduces a multi-object scene exactly:
 * the reference scales each object's mass matrix by the density of the material the OBJECT was created with
 * (src/mesh/tet_mesh.cpp:36-40, tri2d_mesh.cpp:36-41, tri_mesh.cpp:67-68; block-diagonal over objects, meshes.cpp:100-108),
 * which for an object with a per-tet material_ids file is materials[0] (mesh.cpp:14-19), not the per-tet material. */
int mfb_newton_init(mfb_ctx* ctx, const int32_t* free_map, double density, double h, const double* force, const double* x0, const double* v0);
int mfb_newton_init_density(mfb_ctx* ctx, const int32_t* free_map, const double* density_e, double h, const double* force, const double* x0,
                            const double* v0);
int mfb_newton_step(mfb_ctx* ctx, const mfb_newton_opts* opts, mfb_newton_log* log);   /* one time step = NewtonOptimizer::step() */
int mfb_newton_positions(mfb_ctx* ctx, double* x_full);                                 /* host d*nV */
int mfb_newton_positions_dev(mfb_ctx* ctx, double** x_full_dev, const double** mass_values_dev);
void mfb_newton_free(mfb_ctx* ctx);

/* ---- the two saddle-point formulations of the same Newton system, matrix-free on the device.
 * They need the Displacement / BDF1 state of mfb_newton_init and an mfb_update at the current positions.
 * Host vectors use the reference's ordering: KKT vector = [dx (d*nF) ; ds (N*nE, entry N*e+n) ; la (N*nE)].
 *
 * SystemMatrixIndefinite (src/linear_solvers/linear_system.h:64-118) with the BlockMatrix product
 * (src/linear_solvers/block_matrix.h:74-125):  out = [PMP^T dx + B la ; Hloc ds + C la ; B^T dx + C ds],
 * B = jacobian_x(), C = jacobian_mixed();  b = [-x.gradient() ; -gradient_mixed() ; -gradient_dual()]. */
int64_t mfb_kkt_size(mfb_ctx* ctx);                                        /* d*nF + 2*N*nE */
int mfb_kkt_product(mfb_ctx* ctx, const double* v, double* out);          /* out = A v (host, reference ordering) */
int mfb_kkt_rhs(mfb_ctx* ctx, double* out);                               /* SystemMatrixIndefinite::pre_solve, linear_system.h:71-87 */
/* device vectors: [x part | ds component-major N x ld | la component-major N x ld] at the offsets given here; the
 * padding (ld > nE, and between the parts) must be zero in the input and is zero in the output */
int mfb_kkt_layout_dev(mfb_ctx* ctx, int64_t* off_x, int64_t* off_s, int64_t* off_la, int64_t* total);
int mfb_kkt_product_dev(mfb_ctx* ctx, const double* v_dev, double* out_dev);
/* pre_solve + MINRES (identity preconditioner; Eigen MINRES recurrence and stop test |r| < tol |b|) from the guess in
 * sol + post_solve (linear_system.h:89-100: x.delta() = dx, delta() = ds, lambda() = la).  sol: host, in/out. */
int mfb_kkt_solve_minres(mfb_ctx* ctx, double* sol, double tol, int max_iters, int* iters_out, double* relres_out);

/* DualCondensedSystem (src/linear_solvers/linear_system.h:120-239):  lhs = G^T G + C*C*H with G = Ml^-1/2 B,
 * Ml = lumped PMP^T, H = hessian_inv();  rhs = B^T Ml^-1 (-x.gradient()) - C H gradient_mixed() + gradient_dual().
 * The reference builds lhs as an explicit sparse matrix; here it is applied matrix-free (two element passes). */
int mfb_dual_product(mfb_ctx* ctx, const double* la, double* out);        /* out = lhs * la (host, N*nE) */
int mfb_dual_product_dev(mfb_ctx* ctx, const double* la_dev_soa, double* out_dev_soa);   /* component-major N x ld */
int mfb_dual_rhs(mfb_ctx* ctx, double* out);                              /* linear_system.h:158-165 */
/* pre_solve + CG (Eigen ConjugateGradient recurrence, identity preconditioner) from the guess in la + post_solve
 * (linear_system.h:178-209: lambda() = la, x.delta() = Ml^-1(-gx - B la), delta() = H(-gradient_mixed() - C la)).
 * la: host in/out; dx_out (d*nF), ds_out (N*nE): host or NULL. */
int mfb_dual_solve_cg(mfb_ctx* ctx, double* la, double tol, int max_iters, int* iters_out, double* relres_out, double* dx_out,
                      double* ds_out);

/* ---- the non-mixed Stretch<DIM> variable (src/variables/stretch.cpp:16-91) on the same mesh, Assembler
 * pattern and scatter map: F-form material (material_model.h:43-62), H_e = Jloc^T d2psi/dF2 Jloc vol h^2 with
 * sim::simple_psd_fix, g_e = Jloc^T dpsi/dF vol h^2.  One material per mesh, every F-form overload the reference has: Stable-Neohookean
 * (tets, 2D triangles; fix at 0), Neohookean (tets; fix at 1e-6), FixedCorot (tets, 2D triangles; ARAP part + volume term, fix at 1e-6, the
 * reference's quirks kept: no volume term in the 3D energy, mu in 3D where 2D has 2 mu), ARAP (tets, 2D triangles; H = mu (I - dR/dF), no
 * fix).  Corot has no F-form in the reference and Fung's returns zeros: MFB_ERR_UNSUPPORTED.  x: d*nV. ---- */
int mfb_stretch_update(mfb_ctx* ctx, const double* x, double h);             /* Stretch::update, stretch.cpp:33-60 (x host) */
int mfb_stretch_update_dev(mfb_ctx* ctx, const double* x_dev, double h);
int mfb_stretch_lhs_values(mfb_ctx* ctx, double* values);                    /* lhs().valuePtr(), nnz values on mfb_csr_pattern */
int mfb_stretch_gradient(mfb_ctx* ctx, double* out);                         /* gradient(), d*nF ; rhs() = -gradient() (stretch.cpp:62-70) */
int mfb_stretch_dev(mfb_ctx* ctx, const double** lhs_values_dev, const double** gradient_dev);
int mfb_stretch_energy(mfb_ctx* ctx, const double* x, double* e);            /* Stretch::energy, stretch.cpp:16-28 */

/* ---- multi-GPU (the reference has no counterpart: it is a single-address-space OpenMP code).
 * One process per GPU, each with a context on ITS slab of elements (locally renumbered vertices).  The assembled
 * matrix stays sub-assembled; vectors are summed at the free vertices shared between slabs with one ncclAllReduce over
 * a packed buffer laid out over ALL shared vertices.  After mfb_comm_init + mfb_comm_set_interface, mfb_pcg_dev and
 * mfb_newton_step run distributed: every rank calls them collectively with consistent (already summed) vectors. ---- */
/* Partitioner (host code, no device needed).  Every rank calls it with the GLOBAL mesh and its own rank and gets its part:
 * method 0 = slabs of equal element count along the longest axis (stable order of the element centroids), method 1 =
 * recursive coordinate bisection on the centroids.  Local vertices ascend in the global numbering (local free ids keep the
 * order of the global free_map_); local ELEMENTS are ordered interface-first (those touching a free vertex shared with
 * another part), so that mfb_update_dev can exchange their vertex sums while the interior elements run. */
typedef struct mfb_partition mfb_partition;
enum mfb_part_array {
  MFB_PART_ELEMS = 0,   /* int32 nE_l : local element -> global element */
  MFB_PART_VMAP,        /* int32 nV_l : local vertex -> global vertex (ascending) */
  MFB_PART_T,           /* int32 nE_l x npe : local vertex ids, for mfb_mesh_desc::T */
  MFB_PART_V,           /* double nV_l x d */
  MFB_PART_FREE_MAP,    /* int32 nV_l : local free ids */
  MFB_PART_FREE_L2G,    /* int64 nF_l : local free id -> global free id (ascending) */
  MFB_PART_IF_LOCAL,    /* int32 : local free ids of this part's shared vertices ... */
  MFB_PART_IF_POS,      /* int32 : ... and their slots among the n_shared shared vertices of the whole partition */
  MFB_PART_DOT_WEIGHT,  /* double nF_l : 1 interior / owned (lowest rank touching the vertex), 0 copy owned elsewhere */
  MFB_PART_NBR_RANK,    /* int32 n_nbr : ranks this part shares vertices with, ascending */
  MFB_PART_NBR_COUNT,   /* int32 n_nbr : shared vertices per neighbour */
  MFB_PART_NBR_LOCAL,   /* int32 sum(count) : local free ids per neighbour, ascending GLOBAL id (both sides agree on the order) */
  MFB_PART_ELEM_PART    /* int32 nE global : part of every element of the global mesh */
};
int mfb_partition_create(mfb_partition** out, int32_t dim, int32_t npe, int32_t nV, int32_t nE, const int32_t* T, const double* V,
                         const int32_t* free_map, int32_t nparts, int32_t rank, int32_t method);
void mfb_partition_destroy(mfb_partition* p);
/* sizes[8] = nV_l, nE_l, nF_l, n_if_elems, n_shared, n_if_local, n_nbr, n_nbr_entries */
int mfb_partition_sizes(const mfb_partition* p, int32_t* sizes);
int mfb_partition_get(const mfb_partition* p, int what, const void** ptr, int64_t* count);   /* pointers owned by p */

int mfb_comm_unique_id(char* id128);                                  /* rank 0: ncclGetUniqueId; broadcast the 128 bytes */
int mfb_comm_init(mfb_ctx* ctx, const char* id128, int rank, int world);   /* ncclCommInitRank on the context's device */
/* if_local[i] = local free vertex id of this rank's i-th interface vertex, if_pos[i] = its slot among the n_shared
 * shared vertices (same numbering on all ranks); dot_weight[nF local] = 1 for interior vertices and for interface
 * vertices this rank owns, 0 for copies owned by another rank (so that global dot products count every vertex once) */
int mfb_comm_set_interface(mfb_ctx* ctx, const int32_t* if_local, const int32_t* if_pos, int32_t n_if_local, int32_t n_shared,
                           const double* dot_weight);
/* The whole interface description from a partition of THIS context's mesh (mfb_create was given MFB_PART_T / _V / _FREE_MAP):
 * = mfb_comm_set_interface + the per-neighbour lists + the number of leading interface elements.  With it the exchange is
 * point-to-point (grouped ncclSend / ncclRecv of each rank's own cuts, summed on arrival) instead of an all-reduce over
 * every shared vertex of the partition, the CG writes / reads its halo inside its own kernels, and option
 * "update_exchange" = 1 makes mfb_update_dev sum grad_x_ / rhs_ over ranks itself, overlapped with the interior elements. */
int mfb_comm_set_partition(mfb_ctx* ctx, const mfb_partition* p);
/* With a partition set and all ranks on one node, the distributed mfb_pcg_dev (preconditioners 0..2) leaves NCCL out of
 * the iteration: every rank exports one mailbox (cudaMalloc + CUDA IPC, opened by the others on first use), the SpMV
 * kernel stores its interface rows straight into the neighbours' mailboxes over NVLink, the last block of each kernel
 * posts the rank's partial dot products and a sequence number to every rank, and the next kernel of the iteration waits
 * for the sequence numbers in its own memory and adds the partial sums in rank order (bit-identical on all ranks).
 * Spins are bounded; a timeout makes mfb_pcg_dev return MFB_ERR_CUDA.  Option "peer" = 0 keeps the NCCL exchange; so
 * does any rank that cannot export or open a mailbox (decided collectively).  1: in use, -1: unavailable, 0: not tried. */
int mfb_comm_peer_state(mfb_ctx* ctx);
/* a (ncomp_a doubles per local free vertex) and optionally b <- sum over ranks at interface vertices, one collective */
int mfb_interface_sum_dev(mfb_ctx* ctx, double* a_dev, int ncomp_a, double* b_dev, int ncomp_b);
int mfb_allreduce_dev(mfb_ctx* ctx, double* v_dev, int n, int op);   /* op 0 sum, 1 max; stream-ordered */

/* ---- allocation-free entry point.  NO COUNTERPART IN THE REFERENCE: added for
 * callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * mfb_energy_noalloc_dev = mfb_energy_dev (MixedStretch::energy, sum of psi(s_e) vol_e) with the scratch of its two-stage sum
 * in a file-scope static __device__ array of MFB_NOALLOC_MAX_PARTIALS doubles (1 MiB, held by every GPU that loads the
 * library until exit; one double per 128 elements, i.e. at most mfb_noalloc_max_elements() = 16 777 216 elements - larger
 * contexts get MFB_ERR_TOO_LARGE) and the result written to a DEVICE pointer, stream-ordered on the context stream with no
 * host synchronisation.  It calls no cudaMalloc / cudaFree / event creation and no allocating library routine; the first
 * call on a device looks the array's address up once (cudaGetSymbolAddress): make it before starting a capture.  All calls
 * on a device share the array: calls that overlap on the GPU race - keep them on one stream or order them with events.
 * s_dev_soa_or_null: component-major N x ld stretches (NULL: the context's s_), e_dev: one double on the device. ---- */
#define MFB_NOALLOC_MAX_PARTIALS 131072
int mfb_energy_noalloc_dev(mfb_ctx* ctx, const double* s_dev_soa_or_null, double* e_dev);
int64_t mfb_noalloc_max_elements(void);

/* ---- instrumentation ---- */
int mfb_num_kernel_launches(mfb_ctx* ctx, int64_t* n, int reset);   /* kernels this library launched on ctx */
int mfb_set_option(mfb_ctx* ctx, const char* name, int value);      /* "fuse_rhs" (default 1), "profile" (0), "cg_zero_guess" (0:
                                                                       mfb_newton_step starts CG from the reference's guess),
                                                                       "update_exchange" (0; multi-GPU, see mfb_comm_set_partition),
                                                                       "p2p" (1: use the neighbour exchange when a partition is set),
                                                                       "peer" (1: peer mailboxes, see mfb_comm_peer_state).
                                                                       Solver / optimizer: "cg_graph" (0), "cheb_degree" (6),
                                                                       "cheb_ratio" (120), "convergence_log" (0: fill
                                                                       mfb_newton_log.fem_gradient_norm, one GPU).
                                                                       Update-kernel variants kept for A/B measurement, all
                                                                       results identical to rounding: "asm_mode" (1 reductions
                                                                       into the block-CSR upper triangle; 0 into the scalar CSR;
                                                                       2 deterministic gather, bit-stable), "stage_finalize"
                                                                       (0; 1 = finalize per element chunk on a side stream,
                                                                       measured slower), "pipe" (0; persistent pipelined kernel,
                                                                       measured slower), "exp" (what-if timing bits, WRONG
                                                                       results); "cg_persist" (2: persistent cooperative CG
                                                                       kernel for small systems, 1 always, 0 never).
                                                                       Unknown names return MFB_ERR_INVALID. */
/* device time of the update kernel alone, CUDA events on the ctx stream, accumulated while option "profile" is 1 */
int mfb_kernel_time_ms(mfb_ctx* ctx, double* total_ms, int* count);
/* Phase timers under the reference's own Timer keys (src/optimizers/optimizer_data.cpp:127-256; MixedStretch brackets its
 * phases at mixed_stretch.cpp:95,112,168,192,207,217,247,279, the optimizer at newton_optimizer.cpp:122,140,162,164,86,88):
 * CUDA-event spans on the ctx stream, recorded while option "profile" is 1, also emitted as NVTX ranges.  On the device
 * rotations + derivatives + Aloc + assemble are ONE fused launch sequence, reported under "update" (= the optimizer's key
 * for the same span).  ms / counts: MFB_PHASE_COUNT entries; reset != 0 clears the accumulators. */
enum mfb_phase { MFB_PH_UPDATE = 0, MFB_PH_RHS, MFB_PH_LINSOLVE, MFB_PH_SOLVE, MFB_PH_LS, MFB_PH_ENERGY, MFB_PH_CONSTRAINT, MFB_PHASE_COUNT };
const char* mfb_phase_name(int phase);      /* "update","rhs","linsolve","solve","LS","energy","constraint" */
int mfb_phase_times(mfb_ctx* ctx, double* ms, int32_t* counts, int reset);

#ifdef __cplusplus
}
#endif
#endif /* MFEM_B200_H */
