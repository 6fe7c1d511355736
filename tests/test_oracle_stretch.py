"""CPU: the oracle's Stretch variable (oracle/stretch.py).  The closed-form Stable Neo-Hookean F-form gradient / Hessian
are checked against central differences of the reference's expanded energy expression (stable_neohookean.cpp:107), the
assembled vector / matrix against differences of the assembled energy."""
import numpy as np

from mixed_fem_b200 import meshgen
from oracle import stretch as ostretch
from tests.common import LA, MU


def test_snh_f_form_closed_forms_match_the_expanded_energy():
    rng = np.random.default_rng(1)
    F = np.eye(3)[None] + 0.3 * rng.standard_normal((20, 3, 3))
    g, H = ostretch.gradient_hessian(F, MU, LA)
    F9 = np.transpose(F, (0, 2, 1)).reshape(-1, 9)
    eps = 1e-6
    for i in range(9):
        d = np.zeros(9); d[i] = eps
        fd = (ostretch.energy_expanded(F9 + d, MU, LA) - ostretch.energy_expanded(F9 - d, MU, LA)) / (2 * eps)
        assert np.allclose(fd, g[:, i], rtol=1e-6, atol=1e-6 * np.abs(g).max())
        gp, _ = ostretch.gradient_hessian(np.transpose((F9 + d).reshape(-1, 3, 3), (0, 2, 1)), MU, LA)
        gm, _ = ostretch.gradient_hessian(np.transpose((F9 - d).reshape(-1, 3, 3), (0, 2, 1)), MU, LA)
        assert np.allclose((gp - gm) / (2 * eps), H[:, :, i], rtol=1e-5, atol=1e-6 * np.abs(H).max())
    assert np.abs(H - np.transpose(H, (0, 2, 1))).max() <= 1e-9 * np.abs(H).max()


def test_simple_psd_fix_clamps_at_zero():
    rng = np.random.default_rng(2)
    A = rng.standard_normal((5, 9, 9)); A = A + np.transpose(A, (0, 2, 1))
    B = ostretch.simple_psd_fix(A, 0.0)
    w = np.linalg.eigvalsh(B)
    assert w.min() >= -1e-12 and np.allclose(np.linalg.eigvalsh(A).clip(min=0), w, atol=1e-10)


def test_assembled_gradient_is_the_derivative_of_the_energy():
    V, T = meshgen.kuhn_grid(3, 2, 2)
    fm = meshgen.pinned_free_map(V)
    st = ostretch.OracleStretch(V, T, fm, MU, LA)
    x = meshgen.deformed_positions(V, 1.0 / 3)
    h = 0.1
    st.update(x, h)
    free = np.nonzero(fm >= 0)[0]
    rng = np.random.default_rng(0)
    for v in rng.choice(free, 6, replace=False):
        for l in range(3):
            d = np.zeros_like(x); d[3 * v + l] = 1e-6
            fd = (st.energy(x + d) - st.energy(x - d)) / 2e-6 * h * h
            assert abs(fd - st.grad[3 * fm[v] + l]) <= 1e-5 * np.abs(st.grad).max()
    assert abs(st.lhs - st.lhs.T).max() <= 1e-10 * abs(st.lhs).max()


def test_snh_f_form_2d_matches_the_reference_expressions():
    """2D: closed forms against the reference's OWN expanded gradient / Hessian entries (stable_neohookean.cpp:283-321,
    short enough to transcribe a few of) and against differences of its energy"""
    rng = np.random.default_rng(4)
    F = np.eye(2)[None] + 0.3 * rng.standard_normal((10, 2, 2))
    g, H = ostretch.gradient_hessian(F, MU, LA)
    F4 = np.transpose(F, (0, 2, 1)).reshape(-1, 4)
    F1_1, F2_1, F1_2, F2_2 = F4.T
    t = -F1_1 * F2_2 + F1_2 * F2_1 + 1.0
    assert np.allclose(g[:, 0], F1_1 * MU - F2_2 * MU - F2_2 * LA * t, rtol=1e-13)
    assert np.allclose(g[:, 1], F1_2 * MU + F2_1 * MU + F1_2 * LA * t, rtol=1e-13)
    assert np.allclose(H[:, 0, 3], -LA - MU + F1_1 * F2_2 * LA * 2.0 - F1_2 * F2_1 * LA, rtol=1e-13)
    assert np.allclose(H[:, 1, 2], LA + MU - F1_1 * F2_2 * LA + F1_2 * F2_1 * LA * 2.0, rtol=1e-13)
    assert np.allclose(H[:, 2, 3], -F1_1 * F2_1 * LA, rtol=1e-13) and np.allclose(H[:, 1, 1], MU + F1_2 ** 2 * LA, rtol=1e-13)
    eps = 1e-6
    for i in range(4):
        d = np.zeros(4); d[i] = eps
        fd = (ostretch.energy_expanded_2d(F4 + d, MU, LA) - ostretch.energy_expanded_2d(F4 - d, MU, LA)) / (2 * eps)
        assert np.allclose(fd, g[:, i], rtol=1e-6, atol=1e-6 * np.abs(g).max())


def test_assembled_gradient_2d():
    V, T = meshgen.tri_grid(4, 3)
    fm = meshgen.pinned_free_map(V)
    st = ostretch.OracleStretch(V, T, fm, MU, LA)
    x = meshgen.deformed_positions(V, 0.25)
    st.update(x, 0.1)
    for v in np.nonzero(fm >= 0)[0][:5]:
        for l in range(2):
            d = np.zeros_like(x); d[2 * v + l] = 1e-6
            fd = (st.energy(x + d) - st.energy(x - d)) / 2e-6 * 0.01
            assert abs(fd - st.grad[2 * fm[v] + l]) <= 1e-5 * np.abs(st.grad).max()


def test_arap_f_form_closed_dRdF_matches_finite_differences():
    rng = np.random.default_rng(5)
    import scipy.linalg as sl
    F = np.stack([sl.expm(np.cross(np.eye(3), rng.standard_normal(3) * 0.7)) @ (np.diag([1.2, 0.9, 1.05]) + 0.15 * rng.standard_normal((3, 3)))
                  for _ in range(12)])
    g, H, R = ostretch.arap_gradient_hessian(F, MU)
    assert np.abs(R - np.stack([sl.polar(f)[0] for f in F])).max() <= 1e-12
    assert np.abs(H - np.transpose(H, (0, 2, 1))).max() <= 1e-9 * MU
    F9 = np.transpose(F, (0, 2, 1)).reshape(-1, 9)
    eps = 1e-6
    for i in range(9):
        d = np.zeros(9); d[i] = eps
        Fp = np.transpose((F9 + d).reshape(-1, 3, 3), (0, 2, 1)); Fm = np.transpose((F9 - d).reshape(-1, 3, 3), (0, 2, 1))
        fd_e = (ostretch.arap_energy(Fp, MU) - ostretch.arap_energy(Fm, MU)) / (2 * eps)
        assert np.allclose(fd_e, g[:, i], rtol=1e-6, atol=1e-7 * MU)
        gp, _, _ = ostretch.arap_gradient_hessian(Fp, MU); gm, _, _ = ostretch.arap_gradient_hessian(Fm, MU)
        assert np.allclose((gp - gm) / (2 * eps), H[:, :, i], rtol=1e-5, atol=1e-7 * MU)
