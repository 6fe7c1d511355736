"""GPU: y = (M + A) x and the PCG loop against scipy / the Eigen-CG restatement (oracle/pcg.py); full-size
property checks of the assembled matrix at BASELINE sizes where the oracle would take too long."""
import numpy as np
import pytest
import scipy.sparse as sp

from oracle import pcg as opcg
from tests.common import DT, LA, MU, TOL, mesh, pair, vec_rel

pytestmark = pytest.mark.gpu


def _setup(dim, n, model="Stable-Neohookean"):
    import torch
    V, T, fm, h = mesh(dim, n)
    gpu, cpu, st = pair(V, T, fm, [(model, MU, LA)], h=h)
    for o in (gpu, cpu):
        o.update(st["x"], DT)
    outer, inner, vals = cpu.csr()
    nrow = outer.size - 1
    A = sp.csr_matrix((vals, inner, outer), shape=(nrow, nrow))
    # mass-like term on the same pattern: PMP^T has the pattern of A; here rho*vol lumped on the diagonal
    Mv = np.zeros_like(vals)
    diagpos = np.array([outer[r] + np.searchsorted(inner[outer[r]:outer[r + 1]], r) for r in range(nrow)])
    Mv[diagpos] = 1e-3 * (1.0 + 0.1 * np.arange(nrow) % 7)
    M = sp.csr_matrix((Mv, inner, outer), shape=(nrow, nrow))
    dev = torch.device("cuda")
    return gpu, cpu, st, A, M, torch.as_tensor(Mv, device=dev), dev


@pytest.mark.parametrize("dim,n", [(3, (7, 5, 4)), (2, (15, 11, 0))])
def test_spmv_matches_scipy(built, dim, n):
    import torch
    gpu, cpu, st, A, M, Mv_d, dev = _setup(dim, n)
    x = np.random.default_rng(0).standard_normal(A.shape[0])
    xd = torch.as_tensor(x, device=dev)
    yd = torch.zeros_like(xd)
    gpu.spmv_dev(xd.data_ptr(), yd.data_ptr())
    torch.cuda.synchronize()
    assert vec_rel(yd.cpu().numpy(), A @ x) <= 1e-13
    gpu.spmv_dev(xd.data_ptr(), yd.data_ptr(), M_ptr=Mv_d.data_ptr())
    torch.cuda.synchronize()
    assert vec_rel(yd.cpu().numpy(), (A + M) @ x) <= 1e-13
    gpu.close()


@pytest.mark.parametrize("persist", [1, 0])
@pytest.mark.parametrize("dim,n,precond", [(3, (7, 5, 4), 1), (3, (7, 5, 4), 2), (3, (6, 4, 3), 0), (2, (15, 11, 0), 1), (2, (15, 11, 0), 2)])
def test_pcg_matches_eigen_cg_recurrence(built, dim, n, precond, persist):
    """persist 1: the loop as one persistent cooperative kernel (grid barriers between the phases; what small systems get by
    default), 0: three launches per iteration - the same recurrence, each against Eigen's"""
    import torch
    gpu, cpu, st, A, M, Mv_d, dev = _setup(dim, n)
    gpu.set_option("cg_persist", persist)
    K = (A + M).tocsr()
    b = cpu.rhs()
    x0 = 1e-3 * np.random.default_rng(1).standard_normal(b.size)
    tol, maxit = 1e-8, 500
    xr, itr, err = opcg.conjugate_gradient(K, b, x0, opcg.make_precond(K, precond, dim), tol, maxit)
    bd = torch.as_tensor(b, device=dev)
    xd = torch.as_tensor(x0, device=dev).clone()
    it, rr = gpu.pcg_dev(bd.data_ptr(), xd.data_ptr(), M_ptr=Mv_d.data_ptr(), precond=precond, tol=tol, max_iters=maxit)
    torch.cuda.synchronize()
    # rounding (different dot-product summation order) moves the stopping test by an iteration or two on ~400-iteration runs
    assert abs(it - itr) <= max(2, itr // 50), (it, itr)
    assert rr <= tol and err <= tol
    xg = xd.cpu().numpy()
    assert np.linalg.norm(K @ xg - b) <= 2 * tol * np.linalg.norm(b)
    assert vec_rel(xg, xr) <= 1e-6
    # zero right-hand side: Eigen returns x = 0 with 0 iterations
    zd = torch.zeros_like(bd)
    it0, rr0 = gpu.pcg_dev(zd.data_ptr(), xd.data_ptr(), M_ptr=Mv_d.data_ptr(), precond=precond, tol=tol, max_iters=maxit)
    torch.cuda.synchronize()
    assert it0 == 0 and not xd.cpu().numpy().any()
    gpu.close()


def test_persistent_cg_continues_and_stops_like_the_three_launch_loop(built):
    """same system, tight and loose tolerance, an iteration cap that is hit, and a warm start from the other path's result:
    iteration counts equal to +-1, solutions to 1e-9 (the partial sums are grouped differently)"""
    import torch
    gpu, cpu, st, A, M, Mv_d, dev = _setup(3, (12, 8, 6))
    b = cpu.rhs()
    bd = torch.as_tensor(b, device=dev)
    res = {}
    for persist in (1, 0):
        gpu.set_option("cg_persist", persist)
        for tol, maxit in ((1e-10, 2000), (1e-3, 2000), (1e-10, 7)):
            xd = torch.zeros_like(bd)
            it, rr = gpu.pcg_dev(bd.data_ptr(), xd.data_ptr(), M_ptr=Mv_d.data_ptr(), precond=1, tol=tol, max_iters=maxit)
            res[(persist, tol, maxit)] = (it, rr, xd.cpu().numpy())
    for tol, maxit in ((1e-10, 2000), (1e-3, 2000), (1e-10, 7)):
        a, c = res[(1, tol, maxit)], res[(0, tol, maxit)]
        assert abs(a[0] - c[0]) <= 1 and (maxit != 7 or a[0] == c[0] == 7), (a[0], c[0])
        # same iterates up to the grouping of the partial sums; one iteration more or less moves x by about tol
        assert vec_rel(a[2], c[2]) <= (1e-9 if a[0] == c[0] else 10 * tol), (tol, maxit, vec_rel(a[2], c[2]))
    # warm start: the converged solution of one path is a zero-iteration start for the other
    gpu.set_option("cg_persist", 1)
    xd = torch.as_tensor(res[(0, 1e-10, 2000)][2], device=dev).clone()
    it, rr = gpu.pcg_dev(bd.data_ptr(), xd.data_ptr(), M_ptr=Mv_d.data_ptr(), precond=1, tol=1e-8, max_iters=100)
    assert it == 0 and rr <= 1e-8
    gpu.close()


@pytest.mark.parametrize("dim,n,degree", [(3, (9, 6, 5), 4), (3, (9, 6, 5), 2), (3, (7, 5, 4), 1), (2, (15, 11, 0), 3)])
def test_pcg_chebyshev_preconditioner_matches_its_numpy_twin(built, dim, n, degree):
    """precond 3 (Chebyshev polynomial over block-Jacobi, not in the reference): same iteration counts and solution as
    oracle/pcg.py::make_chebyshev inside the Eigen-CG restatement, fewer iterations than block-Jacobi alone, and
    degree 1 degenerates to (scaled) block-Jacobi."""
    import torch
    gpu, cpu, st, A, M, Mv_d, dev = _setup(dim, n)
    K = (A + M).tocsr()
    b = cpu.rhs()
    x0 = np.zeros(b.size)
    tol, maxit = 1e-9, 2000
    xr, itr, err = opcg.conjugate_gradient(K, b, x0, opcg.make_chebyshev(K, dim, b, degree=degree), tol, maxit)
    _, it_bj, _ = opcg.conjugate_gradient(K, b, x0, opcg.make_precond(K, 2, dim), tol, maxit)
    gpu.set_option("cheb_degree", degree)
    bd = torch.as_tensor(b, device=dev)
    xd = torch.zeros_like(bd)
    it, rr = gpu.pcg_dev(bd.data_ptr(), xd.data_ptr(), M_ptr=Mv_d.data_ptr(), precond=3, tol=tol, max_iters=maxit)
    torch.cuda.synchronize()
    xg = xd.cpu().numpy()
    assert abs(it - itr) <= max(2, itr // 25), (it, itr)
    assert rr <= tol and np.linalg.norm(K @ xg - b) <= 2 * tol * np.linalg.norm(b)
    assert vec_rel(xg, xr) <= 1e-6
    if degree == 1:
        assert abs(it - it_bj) <= 2
    else:
        assert it < it_bj, (it, it_bj)
    gpu.close()


def test_chebyshev_cuts_cg_iterations_at_C3(built):
    """CG iterations of the Newton step at config C3 (998 250 tets, ARAP, itr_tol 1e-4) - 380 / 462 / 500 (cap)
    with eigen-pcg-diag - below 150 with the Chebyshev preconditioner (defaults: degree 6 on [lmax / 120, lmax]), same
    energies to the solver tolerance.  Fewer iterations, NOT less time on one GPU: the polynomial costs 1.5x the SpMVs
    (tools/sweep_cheb.py: 44.8 against 31.0 ms per Newton iteration); it is there for runs where reductions dominate."""
    import mixed_fem_b200 as mf
    from mixed_fem_b200 import meshgen
    V, T = meshgen.kuhn_grid(55, 55, 55)
    fm = meshgen.pinned_free_map(V)
    logs = {}
    for precond in (1, 3):
        ms = mf.MixedStretch(V, T, fm, [("ARAP", MU, LA)]); ms.reset()
        ms.newton_init(density=1000.0, h=DT)
        logs[precond] = ms.newton_step(outer_steps=3, precond=precond, cg_tol=1e-4, cg_max_iters=500)
        ms.close()
    print("C3 cg iterations diag", logs[1]["cg_iters"], "chebyshev", logs[3]["cg_iters"])
    assert logs[3]["cg_iters"].max() < 150 and logs[1]["cg_iters"].max() > 300
    assert np.all(logs[3]["cg_error"] <= 1e-4)
    assert np.all(np.abs(logs[3]["energy"] - logs[1]["energy"]) <= 1e-3 * np.abs(logs[1]["energy"]))


@pytest.mark.parametrize("pipe", [0, 1])
def test_full_size_properties_C2(built, pipe):
    """BASELINE config C2 (64x16x16x6 = 98 304 tets): full parity against the oracle, which finishes in seconds.
    pipe = 1: the persistent software-pipelined update kernel (768 tiles over 296 CTAs: every CTA walks 2-3 tiles), twice in
    a row so that the second update starts from the rotations the first one stored."""
    from tests.test_gpu_parity import check_update
    V, T, fm, h = mesh(3, (64, 16, 16))
    gpu, cpu, st = pair(V, T, fm, [("Stable-Neohookean", MU, LA)], h=h)
    gpu.set_option("pipe", pipe)
    for xx in (st["x"], 0.5 * (st["x"] + st["xprev"])) if pipe else (st["x"],):
        for o in (gpu, cpu):
            o.update(xx, DT)
        check_update(gpu, cpu, 3)
    go, gi = gpu.csr_pattern()
    co, ci, _ = cpu.csr(values=False)
    assert np.array_equal(go, co) and np.array_equal(gi, ci)
    gpu.close()


def test_full_size_properties_C3_1M(built):
    """BASELINE config C3 size (55^3 x 6 = 998 250 tets, ARAP): size-independent properties - symmetry x.(Ay) = y.(Ax);
    idempotence (a second update at the same x takes zero polar iterations and reproduces every field);
    product_jacobian_x(gl) reproduces rhs; counters clean.  Full oracle parity at this size: the next test."""
    import torch
    from mixed_fem_b200 import meshgen
    import mixed_fem_b200 as mf
    V, T = meshgen.kuhn_grid(55, 55, 55)
    fm = meshgen.pinned_free_map(V)
    gpu = mf.MixedStretch(V, T, fm, [("ARAP", MU, LA)])
    xa = meshgen.deformed_positions(V, 1.0 / 55, scale=1.0)
    xb = meshgen.deformed_positions(V, 1.0 / 55, scale=0.5)
    gpu.reset(xb)
    rng = np.random.default_rng(3)
    gpu.set("la", rng.uniform(-1, 1, size=6 * T.shape[0]) * MU * 1e-3)
    gpu.update(xa, DT)
    st1 = gpu.status()
    assert st1["polar_maxiter"] == 0 and st1["polar_hfail"] == 0 and st1["nonfinite"] == 0 and st1["polar_iters"] > 0
    rhs1, S1, A1 = gpu.rhs(), gpu.get("S"), gpu.lhs()[2]
    dev = torch.device("cuda")
    n = rhs1.size
    x = torch.as_tensor(rng.standard_normal(n), device=dev); y = torch.as_tensor(rng.standard_normal(n), device=dev)
    Ax = torch.empty_like(x); Ay = torch.empty_like(x)
    gpu.spmv_dev(x.data_ptr(), Ax.data_ptr()); gpu.spmv_dev(y.data_ptr(), Ay.data_ptr())
    torch.cuda.synchronize()
    a, b = float(x @ Ay), float(y @ Ax)
    assert abs(a - b) <= 1e-11 * max(abs(a), abs(b))
    assert float(x @ Ax) > 0                                  # PSD-fixed blocks => positive on a generic vector
    out = gpu.product("jacobian_x", gpu.get("gl"), np.zeros(n), transposed=True)   # out -= PJW dSdF gl  == rhs
    assert vec_rel(out, rhs1) <= 1e-12
    gpu.update(xa, DT)                                         # idempotence
    st2 = gpu.status()
    assert st2["polar_iters"] == 0
    assert vec_rel(gpu.get("S"), S1) <= 1e-14 and vec_rel(gpu.rhs(), rhs1) <= 1e-12 and vec_rel(gpu.lhs()[2], A1) <= 1e-12
    gpu.close()


def test_full_oracle_parity_C3_1M(built):
    """BASELINE config C3 as a parity case: 55^3 x 6 = 998 250 tets, ARAP (and Corot, the other material the config
    names), every update field, the assembled matrix, rhs and the CSR pattern against the oracle (a few seconds of CPU)."""
    from tests.test_gpu_parity import check_update
    V, T, fm, h = mesh(3, (55, 55, 55))
    for model in ("ARAP", "Corot"):
        gpu, cpu, st = pair(V, T, fm, [(model, MU, LA)], h=h)
        for o in (gpu, cpu):
            o.update(st["x"], DT)
        errs = check_update(gpu, cpu, 3)
        print("C3", model, errs)
        if model == "ARAP":
            go, gi = gpu.csr_pattern()
            co, ci, _ = cpu.csr(values=False)
            assert np.array_equal(go, co) and np.array_equal(gi, ci)
        gpu.close()
        del cpu


@pytest.mark.parametrize("dim", [3, 2])
def test_block_csr_export_equals_the_scalar_csr(built, dim):
    """mfb_lhs_bsr: the amgcl hand-off layout (row-major d x d blocks) against the reference-layout scalar CSR."""
    import scipy.sparse as sp
    import mixed_fem_b200 as mf
    from mixed_fem_b200 import meshgen
    from tests.common import LA, MU
    V, T = meshgen.kuhn_grid(5, 3, 2) if dim == 3 else meshgen.tri_grid(7, 5)
    fm = meshgen.pinned_free_map(V, ratio=0.05)
    ms = mf.MixedStretch(V, T, fm, [("Stable-Neohookean", MU, LA)])
    ms.reset()
    ms.update((V * 1.03).reshape(-1), 0.0333)
    outer, inner, vals = ms.lhs()
    n = outer.size - 1
    A = sp.csr_matrix((vals, inner, outer), shape=(n, n))
    ptr, col, blk = ms.lhs_bsr()
    B = sp.bsr_matrix((blk, col, ptr), shape=(n, n))
    assert abs(A - B).max() == 0.0
    ms.newton_init(h=0.0333)
    _, Mptr = ms.newton_positions_dev()
    ptr, col, blk = ms.lhs_bsr(Mptr)
    L = sp.bsr_matrix((blk, col, ptr), shape=(n, n)).tocsr()
    assert abs(L - L.T).max() <= 1e-12 * abs(L).max()
    assert (L - A).diagonal().min() > 0                 # the mass matrix sits on the diagonal blocks (and beyond)
    ms.close()
