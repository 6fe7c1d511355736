"""GPU: the device-resident Newton time step (mfb_newton_*) against the oracle's restatement of
NewtonOptimizer::step (oracle/newton.py) - energy / decrement / alpha histories within 1e-8 (north_star), final
positions, and a second time step that starts from the BDF1 velocity of the first."""
import numpy as np
import pytest

import mixed_fem_b200 as mf
from mixed_fem_b200 import meshgen
from oracle import newton as onewton
from oracle import orc
from tests.common import LA, MU, compare_histories, vec_rel

pytestmark = pytest.mark.gpu


def _run(dim, model, n, h, steps=2, outer=3, pin=0.01):
    """`outer` = 3 Newton iterations per time step for the multi-step runs: both sides then do the same number of
    iterations.  Letting the loop run into its termination test (decrement ~ newton_tol, energy change ~ 1e-12) makes the
    iteration count of a step a coin flip on rounding noise, and the next step's history inherits the difference at the
    1e-7 level (measured: 1 run in 40).  The convergence tail is covered by the single-step runs with outer = 6."""
    if dim == 3:
        V, T = meshgen.kuhn_grid(*n)
    else:
        V, T = meshgen.tri_grid(n[0], n[1])
    fm = meshgen.pinned_free_map(V, ratio=pin)
    mats = [(model, MU, LA)]
    cpu = orc.OracleMixedStretch(V, T, fm, mats); cpu.reset()
    gpu = mf.MixedStretch(V, T, fm, mats); gpu.reset()
    ont = onewton.OracleNewton(cpu, V, T, fm, density=1000.0, h=h)
    gpu.newton_init(density=1000.0, h=h)
    gpu.set_option("cg_zero_guess", 1)        # the oracle solves exactly; see the note on the guess in mfb_newton.cu
    if dim == 3:
        gpu.set_option("asm_mode", 2)         # deterministic assembly: the history is bit-reproducible from run to run
    for _ in range(steps):
        lo = ont.step(outer_steps=outer, newton_tol=1e-10)
        lg = gpu.newton_step(outer_steps=outer, newton_tol=1e-10, precond=2, cg_tol=1e-13, cg_max_iters=5000)
        compare_histories(lg, lo)
        assert vec_rel(gpu.newton_positions(), ont.x) <= 1e-9
        assert vec_rel(gpu.get("s"), cpu.get("s")) <= 1e-9
        assert not gpu.get("la").any()
    st = gpu.status()
    assert st["nonfinite"] == 0 and st["polar_maxiter"] == 0
    gpu.close()


@pytest.mark.parametrize("model", ["Stable-Neohookean", "ARAP", "Corot"])
def test_newton_history_matches_oracle_tets(built, model):
    _run(3, model, (8, 3, 3), 0.02)
    _run(3, model, (8, 3, 3), 0.02, steps=1, outer=6)


def test_newton_history_matches_oracle_triangles(built):
    """2D mixed-stretch Newton step, Stable-Neohookean triangles, hanging from the left edge (the C1 configuration's
    element type and material on a synthetic sheet)."""
    # h = 0.01: with the default 0.0333 the condensed system of this thin sheet has kappa ~ 1e5 and the CG solution (tol
    # 1e-13) differs from the oracle's direct solve by more than the 1e-8 bar - a linear-solver effect
    _run(2, "Stable-Neohookean", (16, 10, 0), 0.01)
    _run(2, "Stable-Neohookean", (16, 10, 0), 0.01, steps=1, outer=6)


def test_newton_with_reference_default_solver_settings(built):
    """eigen-pcg-diag, itr_tol 1e-4 (config.h:131): histories agree with the oracle running the Eigen-CG restatement with
    the same tolerance only to the CG tolerance, so the bar here is convergence, not 1e-8 parity."""
    V, T = meshgen.kuhn_grid(10, 4, 4)
    fm = meshgen.pinned_free_map(V)
    gpu = mf.MixedStretch(V, T, fm, [("Stable-Neohookean", MU, LA)]); gpu.reset()
    gpu.newton_init(h=0.0333)
    lg = gpu.newton_step(outer_steps=5, precond=1, cg_tol=1e-4, cg_max_iters=500)
    assert lg["n_iters"] >= 2 and np.all(np.isfinite(lg["energy"]))
    assert lg["decrement"][-1] < lg["decrement"][0]
    assert np.all(np.diff(lg["energy"]) <= 1e-9 * np.abs(lg["energy"][:-1]))
    assert lg["cg_iters_total"] > 0
    gpu.close()


def test_newton_history_matches_oracle_shell(built):
    """config C5: TriMesh shell (x-dependent local Jacobian, triangle mass weights on doublearea*thickness/2) through two
    device-resident time steps, cloth parameters of apps/cloth_sim.cpp:64-71."""
    V, T, N = meshgen.shell_grid(12, 8)
    fm = meshgen.pinned_free_map(V, ratio=0.05)
    mu, la = meshgen.enu_to_lame(0.57e6, 0.3)
    mats = [("Stable-Neohookean", mu, la)]
    cpu = orc.OracleMixedStretch(V, T, fm, mats, normals=N, thickness=1e-3); cpu.reset()
    gpu = mf.MixedStretch(V, T, fm, mats, normals=N, thickness=1e-3); gpu.reset()
    ont = onewton.OracleNewton(cpu, V, T, fm, density=100.0, h=0.01)
    gpu.newton_init(density=100.0, h=0.01)
    gpu.set_option("cg_zero_guess", 1)
    for outer in (3, 6):                 # see _run: only the last step runs into the termination test
        lo = ont.step(outer_steps=outer, newton_tol=1e-10)
        lg = gpu.newton_step(outer_steps=outer, newton_tol=1e-10, precond=2, cg_tol=1e-13, cg_max_iters=5000)
        compare_histories(lg, lo)
        assert vec_rel(gpu.newton_positions(), ont.x) <= 1e-9
        assert vec_rel(gpu.get("s"), cpu.get("s")) <= 1e-9
    st = gpu.status()
    assert st["nonfinite"] == 0 and st["polar_maxiter"] == 0
    gpu.close()
