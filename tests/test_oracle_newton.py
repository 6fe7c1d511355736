"""CPU checks of the oracle's Newton time step (oracle/newton.py): what pins it, since the reference ships no history
fixtures: (1) the consistent mass matrix against its defining integral properties, (2) direct and
CG solves of the condensed system giving the same history, (3) monotone energy decrease
and Newton-like contraction of the decrement on a hanging beam."""
import numpy as np
from mixed_fem_b200 import meshgen
from oracle import newton as onewton
from oracle import orc
from tests.common import LA, MU, compare_histories


def _beam(dim):
    if dim == 3:
        V, T = meshgen.kuhn_grid(6, 2, 2)
    else:
        V, T = meshgen.tri_grid(10, 4)
    fm = meshgen.pinned_free_map(V, ratio=0.01)
    return V, T, fm


def test_mass_matrix_properties():
    for dim in (3, 2):
        V, T, fm = _beam(dim)
        var = orc.OracleMixedStretch(V, T, fm, [("Stable-Neohookean", MU, LA)])
        var.reset()
        vols = var.get("vols")
        M = onewton.mass_matrix(V, T, vols, 1000.0)
        assert abs(M - M.T).max() < 1e-12
        one = np.tile(np.eye(dim)[0], V.shape[0])                     # rigid translation along x
        total = 1000.0 * vols.sum()
        # tets: 4*(1/10) + 12*(1/20) = 1 of the element mass ; triangles: 3/6 + 6/12 = 1
        assert abs(one @ (M @ one) - total) <= 1e-12 * total
        assert np.linalg.eigvalsh(M.toarray()).min() > 0


def test_direct_and_cg_solves_give_the_same_history():
    """the condensed system itself is pinned against the explicit KKT elimination in tests/test_oracle_system.py; here the
    two linear-solve options of the oracle step (SuperLU, Eigen-CG restatement to 1e-12) must give the same history."""
    V, T, fm = _beam(3)
    logs = []
    for cg in (None, (1, 1e-12, 2000)):
        var = orc.OracleMixedStretch(V, T, fm, [("Stable-Neohookean", MU, LA)])
        var.reset()
        nt = onewton.OracleNewton(var, V, T, fm, h=0.02)
        logs.append(nt.step(outer_steps=4, cg=cg))
    compare_histories(logs[0], logs[1])


def test_newton_history_behaves_like_newton():
    for dim in (3, 2):
        V, T, fm = _beam(dim)
        var = orc.OracleMixedStretch(V, T, fm, [("Stable-Neohookean", MU, LA)])
        var.reset()
        nt = onewton.OracleNewton(var, V, T, fm, h=0.01)
        log = nt.step(outer_steps=8, newton_tol=1e-9)
        assert log["n_iters"] >= 2
        assert np.all(np.isfinite(log["energy"])) and np.all(log["alpha"] > 0)
        assert np.all(np.diff(log["energy"]) <= 1e-9 * np.abs(log["energy"][:-1]))      # the line search never goes uphill
        assert log["decrement"][-1] < 1e-2 * log["decrement"][0]
        # gravity pulls the free end down
        y = nt.x.reshape(-1, dim)[:, 1] - V[:, 1]
        assert y[fm >= 0].mean() < 0
        # second step starts from the new velocity
        log2 = nt.step(outer_steps=3)
        assert np.isfinite(log2["energy"]).all()


def test_shell_newton_step_on_the_oracle():
    """config C5 (TriMesh shell) through the time step: triangle mass weights on vol = doublearea*thickness/2
    (tri_mesh.cpp:38-69), energy decreases monotonically, the decrement contracts, direct and CG solves agree."""
    V, T, N = meshgen.shell_grid(8, 5)
    fm = meshgen.pinned_free_map(V, ratio=0.05)
    mu, la = meshgen.enu_to_lame(0.57e6, 0.3)
    logs = []
    for cg in (None, (2, 1e-13, 5000)):
        var = orc.OracleMixedStretch(V, T, fm, [("Stable-Neohookean", mu, la)], normals=N, thickness=1e-3)
        var.reset()
        vols = var.get("vols")
        M = onewton.mass_matrix(V, T, vols, 100.0)
        one = np.tile(np.eye(3)[1], V.shape[0])
        assert abs(one @ (M @ one) - 100.0 * vols.sum()) <= 1e-12 * 100.0 * vols.sum()
        nt = onewton.OracleNewton(var, V, T, fm, density=100.0, h=0.01)
        logs.append(nt.step(outer_steps=6, newton_tol=1e-10, cg=cg))
    lo = logs[0]
    assert np.all(np.diff(lo["energy"]) <= 1e-9 * np.abs(lo["energy"][:-1]))
    assert lo["decrement"][-1] < 1e-3 * lo["decrement"][0]
    compare_histories(logs[1], logs[0])
