"""GPU: the saddle-point systems on the device (mfb_kkt_* = SystemMatrixIndefinite + BlockMatrix product, mfb_dual_* =
DualCondensedSystem, mixed-fem_b200/csrc/mfb_saddle.cu) against the oracle's restatement (oracle/saddle.py): operator
applications and right-hand sides at 1e-10, and the two iterative solvers against the same recurrences in numpy."""
import numpy as np
import pytest

import mixed_fem_b200 as mf
from mixed_fem_b200 import meshgen
from oracle import newton as onewton
from oracle import orc
from oracle import pcg as opcg
from oracle import saddle as osaddle
from tests.common import LA, MU, vec_rel

pytestmark = pytest.mark.gpu


def _pair(kind, model, n, h=0.02):
    kw, density, mats = {}, 1000.0, [(model, MU, LA)]
    if kind == "tet":
        V, T = meshgen.kuhn_grid(*n)
    elif kind == "tri2d":
        V, T = meshgen.tri_grid(n[0], n[1])
    else:                                   # cloth parameters of apps/cloth_sim.cpp:64-71, as tests/test_gpu_newton.py
        V, T, N = meshgen.shell_grid(n[0], n[1])
        kw = dict(normals=N, thickness=1e-3)
        mats, density, h = [(model,) + tuple(meshgen.enu_to_lame(0.57e6, 0.3))], 100.0, 0.01
    fm = meshgen.pinned_free_map(V, ratio=0.05 if kind == "shell" else 0.01)
    cpu = orc.OracleMixedStretch(V, T, fm, mats, **kw); cpu.reset()
    gpu = mf.MixedStretch(V, T, fm, mats, **kw); gpu.reset()
    ont = onewton.OracleNewton(cpu, V, T, fm, density=density, h=h)
    gpu.newton_init(density=density, h=h)
    gpu.set_option("cg_zero_guess", 1)
    ont.step(outer_steps=1)
    gpu.newton_step(outer_steps=1, precond=2, cg_tol=1e-13, cg_max_iters=5000)
    x = gpu.newton_positions()
    assert vec_rel(x, ont.x) <= 1e-9
    cpu.update(ont.x, h); gpu.update(x, h)
    return gpu, cpu, ont


CASES = [("tet", "Stable-Neohookean", (4, 3, 2)), ("tet", "ARAP", (4, 3, 2)), ("tri2d", "Stable-Neohookean", (8, 6, 0)),
         ("shell", "Stable-Neohookean", (6, 5, 0))]


@pytest.mark.parametrize("kind,model,n", CASES)
def test_kkt_operator_and_rhs_match_oracle(built, kind, model, n):
    gpu, cpu, ont = _pair(kind, model, n)
    sd = osaddle.OracleSaddle(cpu, ont)
    nk = gpu.kkt_size()
    assert nk == sd.n + 2 * sd.m
    rng = np.random.default_rng(5)
    for scale in (1.0, 1e-3):
        v = rng.standard_normal(nk) * scale
        assert vec_rel(gpu.kkt_product(v), sd.kkt_product(v)) <= 1e-10
    b_g, b_o = gpu.kkt_rhs(), sd.kkt_rhs()
    for lo, hi in ((0, sd.n), (sd.n, sd.n + sd.m), (sd.n + sd.m, nk)):
        assert vec_rel(b_g[lo:hi], b_o[lo:hi]) <= 1e-9
    gpu.close()


@pytest.mark.parametrize("kind,model,n", CASES)
def test_dual_condensed_operator_and_rhs_match_oracle(built, kind, model, n):
    gpu, cpu, ont = _pair(kind, model, n)
    sd = osaddle.OracleSaddle(cpu, ont)
    lhs, rhs, _ = sd.dual_system()
    rng = np.random.default_rng(6)
    la = rng.standard_normal(sd.m)
    assert vec_rel(gpu.dual_product(la), lhs @ la) <= 1e-10
    assert vec_rel(gpu.dual_rhs(), rhs) <= 1e-9
    gpu.close()


def test_dual_condensed_cg_and_post_solve_follow_the_eigen_recurrence(built):
    gpu, cpu, ont = _pair("tet", "Stable-Neohookean", (4, 3, 2))
    sd = osaddle.OracleSaddle(cpu, ont)
    lhs, rhs, parts = sd.dual_system()
    its = 40
    la_o, it_o, err_o = opcg.conjugate_gradient(lhs, rhs, np.zeros(sd.m), lambda r: r.copy(), 1e-30, its)
    la_g, dx_g, ds_g, it_g, err_g = gpu.dual_solve_cg(tol=1e-30, max_iters=its)
    assert it_g == it_o == its
    assert vec_rel(la_g, la_o) <= 1e-7
    assert abs(err_g - err_o) <= 1e-6 * err_o
    dx_o, ds_o = sd.dual_post_solve(la_g, parts)
    assert vec_rel(dx_g, dx_o) <= 1e-10 and vec_rel(ds_g, ds_o) <= 1e-10
    assert vec_rel(gpu.get("la"), la_g) == 0.0 and vec_rel(gpu.get("ds"), ds_g) == 0.0
    gpu.close()


def test_kkt_minres_follows_the_recurrence_and_reduces_the_residual(built):
    gpu, cpu, ont = _pair("tet", "Stable-Neohookean", (3, 2, 2))
    sd = osaddle.OracleSaddle(cpu, ont)
    b = sd.kkt_rhs()
    # the Lanczos vectors of this badly scaled KKT matrix lose orthogonality quickly, so two correct implementations
    # of the recurrence drift apart in rounding: 20 iterations are held to 1e-6, 60 to 1e-3 (measured 1e-5)
    for its, tol in ((20, 1e-6), (60, 1e-3)):
        x_o, it_o, err_o = osaddle.minres(sd.kkt_product, b, np.zeros(b.size), 1e-30, its)
        x_g, it_g, err_g = gpu.kkt_solve_minres(tol=1e-30, max_iters=its)
        assert it_g == it_o == its
        assert vec_rel(x_g, x_o) <= tol
        assert abs(err_g - err_o) <= 10 * tol * err_o and err_g < 1.0
    # the true residual of the returned iterate agrees with the recurrence's estimate
    r = b - sd.kkt_product(x_g)
    assert abs(np.linalg.norm(r) / np.linalg.norm(b) - err_g) <= 1e-3 * err_g
    # post_solve: delta() = ds part, lambda() = la part
    assert vec_rel(gpu.get("ds"), x_g[sd.n:sd.n + sd.m]) == 0.0 and vec_rel(gpu.get("la"), x_g[sd.n + sd.m:]) == 0.0
    gpu.close()
