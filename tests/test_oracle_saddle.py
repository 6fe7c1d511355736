"""CPU: the oracle's restatement of the saddle-point systems (oracle/saddle.py).  Nothing in the reference pins these
(SURVEY.md §8c), so the KKT operator is held against an independent fact: eliminating ds and la from
SystemMatrixIndefinite's system must give the step of SystemMatrixPD ((PMP^T + A) dx = -grad_x + rhs_), which the
assembled-matrix path of the oracle computes through entirely different code (Assembler + rhs())."""
import numpy as np
import pytest
import scipy.sparse as sp
import scipy.sparse.linalg as spla

from mixed_fem_b200 import meshgen
from oracle import newton as onewton
from oracle import orc
from oracle import saddle as osaddle
from tests.common import LA, MU


def _state(dim, model, n, h=0.02):
    if dim == 3:
        V, T = meshgen.kuhn_grid(*n)
    else:
        V, T = meshgen.tri_grid(n[0], n[1])
    fm = meshgen.pinned_free_map(V, ratio=0.01)
    var = orc.OracleMixedStretch(V, T, fm, [(model, MU, LA)])
    var.reset()
    nt = onewton.OracleNewton(var, V, T, fm, density=1000.0, h=h)
    nt.step(outer_steps=1)             # a deformed state with la_ = 0 and s_ != S(x)
    var.update(nt.x, h)
    return var, nt


@pytest.mark.parametrize("dim,model,n", [(3, "Stable-Neohookean", (3, 2, 2)), (3, "ARAP", (3, 2, 2)), (2, "Stable-Neohookean", (5, 4, 0))])
def test_kkt_elimination_gives_the_condensed_newton_step(built, dim, model, n):
    var, nt = _state(dim, model, n)
    sd = osaddle.OracleSaddle(var, nt)
    K = sd.kkt_matrix()
    assert np.abs(K - K.T).max() <= 1e-12 * np.abs(K).max()
    b = sd.kkt_rhs()
    # update() leaves grad_la_ = +Sym (S - s) vol: the line `grad_la_ = -grad_la_` is commented out in the reference
    # (mixed_stretch.cpp:158-164), so SystemMatrixIndefinite's third block row carries the
    # constraint residual with the opposite sign of the linearisation the condensed path uses.  The oracle keeps the
    # reference's sign; the elimination identity holds for the flipped one.
    b[sd.n + sd.m:] *= -1.0
    v = np.linalg.solve(K, b)
    outer, inner, vals = var.csr()
    A = sp.csr_matrix((vals, inner, outer), shape=(sd.n, sd.n))
    dx = spla.spsolve((nt.PMP + A).tocsc(), -sd.grad_x() + var.rhs())
    assert np.linalg.norm(v[:sd.n] - dx) <= 1e-7 * np.linalg.norm(dx)
    # post_solve of the condensed path reproduces ds and la of the KKT solution (mixed_stretch.cpp:277-292)
    var.solve(dx)
    assert np.linalg.norm(v[sd.n:sd.n + sd.m] - var.get("ds")) <= 1e-6 * np.linalg.norm(var.get("ds"))
    # the BlockMatrix carries +B, +C where the Lagrangian la.Sym(S - s) vol has -B, -C: its multiplier is -lambda of solve()
    assert np.linalg.norm(v[sd.n + sd.m:] + var.get("la")) <= 1e-6 * np.linalg.norm(var.get("la"))


def test_minres_restatement_solves_an_indefinite_system():
    rng = np.random.default_rng(3)
    Q, _ = np.linalg.qr(rng.standard_normal((40, 40)))
    A = Q @ np.diag(np.concatenate([np.linspace(1, 5, 25), -np.linspace(1, 4, 15)])) @ Q.T
    b = rng.standard_normal(40)
    x, it, err = osaddle.minres(A, b, np.zeros(40), 1e-12, 500)
    assert it < 200 and np.linalg.norm(A @ x - b) <= 1e-10 * np.linalg.norm(b)
    assert abs(err - np.linalg.norm(A @ x - b) / np.linalg.norm(b)) <= 1e-9


def test_dual_condensed_matrix_matches_its_factors(built):
    """lhs = G^T G + C*C*H assembled explicitly equals the matrix-free application B^T Ml^-1 B la + C C H la"""
    var, nt = _state(3, "Stable-Neohookean", (3, 2, 2))
    sd = osaddle.OracleSaddle(var, nt)
    lhs, rhs, parts = sd.dual_system()
    la = np.random.default_rng(0).standard_normal(sd.m)
    t = var.product("jacobian_x", la, np.zeros(sd.n), transposed=True)
    w = parts["Minv"] * t
    out = var.product("jacobian_x", w, np.zeros(sd.m), transposed=False)
    out += parts["C"] @ (parts["C"] @ (parts["H"] @ la))
    assert np.linalg.norm(lhs @ la - out) <= 1e-12 * np.linalg.norm(out)
    assert np.isfinite(rhs).all()
