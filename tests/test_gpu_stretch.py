"""GPU: the non-mixed Stretch variable (mfb_stretch_*, mixed-fem_b200/csrc/mfb_stretch.cu) against oracle/stretch.py:
energy, assembled gradient and assembled matrix on the Assembler pattern, 1e-10 relative (north_star)."""
import numpy as np
import pytest
import scipy.sparse as sp

import mixed_fem_b200 as mf
from mixed_fem_b200 import meshgen
from oracle import stretch as ostretch
from tests.common import LA, MU, vec_rel

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n,scale", [((6, 3, 3), 1.0), ((5, 4, 3), 6.0), ((9, 7, 0), 1.0), ((8, 6, 0), 6.0)])
def test_stretch_update_matches_oracle(built, n, scale):
    """scale 6: strongly deformed state, a good part of the element Hessians is indefinite and goes through the clamp;
    n[2] == 0: 2D triangles (config C1's element type and material)"""
    V, T = meshgen.kuhn_grid(*n) if n[2] else meshgen.tri_grid(n[0], n[1])
    fm = meshgen.pinned_free_map(V)
    x = meshgen.deformed_positions(V, 1.0 / n[0], scale=scale)
    h = 0.0333
    gpu = mf.MixedStretch(V, T, fm, [("Stable-Neohookean", MU, LA)])
    cpu = ostretch.OracleStretch(V, T, fm, MU, LA)
    cpu.update(x, h)
    gpu.stretch_update(x, h)
    outer, inner, vals = gpu.stretch_lhs()
    A = sp.csr_matrix((vals, inner, outer), shape=cpu.lhs.shape)
    assert abs(A - cpu.lhs).max() <= 1e-10 * abs(cpu.lhs).max()
    assert vec_rel(gpu.stretch_gradient(), cpu.grad) <= 1e-10
    assert abs(gpu.stretch_energy(x) - cpu.energy(x)) <= 1e-11 * abs(cpu.energy(x))
    if scale > 1.0:     # the clamp was exercised
        F = meshgen.deformation_gradients(V, T, x)
        _, H = ostretch.gradient_hessian(F, MU, LA)
        assert (np.linalg.eigvalsh(H)[:, 0] < -1e-6 * MU).mean() > 0.05
    gpu.close()


def test_stretch_arap_f_form_matches_oracle(built):
    """ARAP F-form (arap.cpp:134-162): polar rotation and dR/dF, no psd fix"""
    V, T = meshgen.kuhn_grid(5, 3, 3)
    fm = meshgen.pinned_free_map(V)
    x = meshgen.deformed_positions(V, 0.2, scale=3.0)
    h = 0.0333
    gpu = mf.MixedStretch(V, T, fm, [("ARAP", MU, LA)])
    cpu = ostretch.OracleStretch(V, T, fm, MU, LA, model="ARAP")
    cpu.update(x, h)
    gpu.stretch_update(x, h)
    outer, inner, vals = gpu.stretch_lhs()
    A = sp.csr_matrix((vals, inner, outer), shape=cpu.lhs.shape)
    assert abs(A - cpu.lhs).max() <= 1e-10 * abs(cpu.lhs).max()
    assert vec_rel(gpu.stretch_gradient(), cpu.grad) <= 1e-10
    assert abs(gpu.stretch_energy(x) - cpu.energy(x)) <= 1e-10 * abs(cpu.energy(x))
    assert gpu.status()["polar_hfail"] == 0
    gpu.close()


@pytest.mark.parametrize("model,n,scale", [("Neohookean", (6, 3, 3), 2.0), ("FixedCorot", (5, 4, 3), 1.0), ("FixedCorot", (5, 4, 3), 5.0),
                                           ("ARAP", (9, 7, 0), 3.0), ("FixedCorot", (9, 7, 0), 1.0), ("FixedCorot", (8, 6, 0), 5.0)])
def test_stretch_remaining_f_forms_match_oracle(built, model, n, scale):
    """the other F-form overloads the reference has: Neohookean 9-vector (neohookean.cpp:93-227), FixedCorot 9- and 4-vector
    (fixed_corotational.cpp:136-358: ARAP part + volume term, simple_psd_fix at 1e-6, and the quirks - no volume term in the
    3D energy, mu in 3D where 2D has 2 mu), ARAP 4-vector (arap.cpp:82-132: closed-form 2x2 polar).  The oracle's closed
    forms are pinned against the reference's expression text (tests/test_oracle_stretch.py)."""
    V, T = meshgen.kuhn_grid(*n) if n[2] else meshgen.tri_grid(n[0], n[1])
    fm = meshgen.pinned_free_map(V)
    x = meshgen.deformed_positions(V, 1.0 / n[0], scale=scale)
    h = 0.0333
    gpu = mf.MixedStretch(V, T, fm, [(model, MU, LA)])
    cpu = ostretch.OracleStretch(V, T, fm, MU, LA, model=model)
    F = meshgen.deformation_gradients(V, T, x)
    assert np.linalg.det(F).min() > 0.05
    cpu.update(x, h)
    gpu.stretch_update(x, h)
    outer, inner, vals = gpu.stretch_lhs()
    A = sp.csr_matrix((vals, inner, outer), shape=cpu.lhs.shape)
    assert abs(A - cpu.lhs).max() <= 1e-10 * abs(cpu.lhs).max()
    assert vec_rel(gpu.stretch_gradient(), cpu.grad) <= 1e-10
    assert abs(gpu.stretch_energy(x) - cpu.energy(x)) <= 1e-10 * abs(cpu.energy(x))
    assert gpu.status()["polar_hfail"] == 0
    gpu.close()


def test_stretch_per_element_materials_match_oracle(built):
    """one F-form material per element (the reference's Stretch variable asks each element's own material_,
    stretch.cpp:36-60): SNH and FixedCorot halves of one mesh, different Lame parameters"""
    V, T = meshgen.kuhn_grid(6, 4, 3)
    fm = meshgen.pinned_free_map(V)
    x = meshgen.deformed_positions(V, 1.0 / 6, scale=3.0)
    mats = [("Stable-Neohookean", MU, LA), ("FixedCorot", 0.5 * MU, 2.0 * LA)]
    mid = (np.arange(T.shape[0]) % 3 == 0).astype(np.int32)
    h = 0.0333
    gpu = mf.MixedStretch(V, T, fm, mats, mat_id=mid)
    cpu = ostretch.OracleStretch(V, T, fm, materials=mats, mat_id=mid)
    cpu.update(x, h)
    gpu.stretch_update(x, h)
    outer, inner, vals = gpu.stretch_lhs()
    A = sp.csr_matrix((vals, inner, outer), shape=cpu.lhs.shape)
    assert abs(A - cpu.lhs).max() <= 1e-10 * abs(cpu.lhs).max()
    assert vec_rel(gpu.stretch_gradient(), cpu.grad) <= 1e-10
    assert abs(gpu.stretch_energy(x) - cpu.energy(x)) <= 1e-10 * abs(cpu.energy(x))
    # and it is not the single-material answer
    one = ostretch.OracleStretch(V, T, fm, MU, LA)
    one.update(x, h)
    assert vec_rel(one.grad, cpu.grad) > 1e-3
    gpu.close()


def test_stretch_rejects_what_it_does_not_implement(built):
    V, T = meshgen.kuhn_grid(2, 2, 2)
    fm = meshgen.pinned_free_map(V)
    gpu = mf.MixedStretch(V, T, fm, [("Corot", MU, LA)])           # Corotational has no F-form overload in the reference
    with pytest.raises(mf.MfbError):
        gpu.stretch_update(V.reshape(-1), 0.1)
    gpu.close()
    V2, T2 = meshgen.tri_grid(3, 3)
    gpu = mf.MixedStretch(V2, T2, meshgen.pinned_free_map(V2), [("Stable-Neohookean", MU, LA)])
    gpu.close()
