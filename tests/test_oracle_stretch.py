"""CPU: the oracle's Stretch variable (oracle/stretch.py).  The closed-form Stable Neo-Hookean F-form gradient / Hessian
are checked against central differences of the reference's expanded energy expression (stable_neohookean.cpp:107), the
assembled vector / matrix against differences of the assembled energy."""
import numpy as np
import pytest

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


def test_fform_closed_forms_match_the_reference_expressions():
    """tests/golden/fform_golden.json holds numbers evaluated from the reference's OWN generated F-form code
    (make_fform_golden.py): Stable-Neohookean 9- and 4-vector gradient / Hessian, Neohookean 9-vector energy / gradient /
    Hessian (before the psd fix), and the volume term FixedCorot adds in 3D.  The oracle's closed forms reproduce them."""
    import json
    import os
    from oracle import stretch as ost
    cases = json.load(open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "fform_golden.json")))["cases"]
    assert len(cases) >= 40
    seen = set()
    for c in cases:
        n = c["n"]; d = 3 if n == 9 else 2
        F = np.array(c["F"]).reshape(d, d).T[None]                 # column-major vec -> matrix
        mu, la = c["mu"], c["la"]
        scale = max(mu, la)
        if c["model"] == "Stable-Neohookean":
            g, H = ost.gradient_hessian(F, mu, la)
            assert np.abs(g[0] - c["g_assign"]).max() <= 1e-12 * scale
            assert np.abs(H[0] - np.array(c["H_assign"])).max() <= 1e-12 * scale
            Fv = np.array(c["F"])[None]
            e = ost.energy_expanded(Fv, mu, la) if d == 3 else ost.energy_expanded_2d(Fv, mu, la)
            assert abs(e[0] - c["energy"]) <= 1e-12 * scale
        elif c["model"] == "Neohookean":
            g, H = ost.neohookean_gradient_hessian(F, mu, la)
            assert np.abs(g[0] - c["g_assign"]).max() <= 1e-11 * scale
            Href = np.array(c["H_assign"])
            assert np.abs(H[0] - Href).max() <= 1e-11 * max(scale, np.abs(Href).max())
            assert abs(ost.neohookean_energy(F, mu, la)[0] - c["energy"]) <= 1e-12 * scale
        else:
            gv, Hv = ost.volume_term(F, la)
            assert np.abs(gv[0] - c["g_add"]).max() <= 1e-12 * scale
            assert np.abs(Hv[0] - np.array(c["H_add"])).max() <= 1e-12 * scale
        seen.add((c["model"], n))
    assert seen == {("Stable-Neohookean", 9), ("Stable-Neohookean", 4), ("Neohookean", 9), ("FixedCorot", 9)}


@pytest.mark.parametrize("model,dim", [("Neohookean", 3), ("FixedCorot", 3), ("ARAP", 2), ("FixedCorot", 2)])
def test_fform_gradients_are_derivatives_of_the_energies(model, dim):
    """central differences: g = d psi / dF and (before the psd fix) H = dg / dF for the F-forms that are not pure expression
    text - the ARAP parts (rotation and its derivative) and, for FixedCorot 2D, the whole form.  FixedCorot 3D: the
    reference's energy lacks the volume term its gradient has (fixed_corotational.cpp:215-223), so g is checked against
    psi + la/2 (J - 1)^2."""
    from oracle import stretch as ost
    rng = np.random.default_rng(2)
    mu, la = 3.0, 7.0
    F = np.eye(dim)[None] + 0.2 * rng.standard_normal((6, dim, dim))
    M = dim * dim

    def psi(Fm):
        e, _, _ = ost.model_energy_gradient_hessian(model, Fm, mu, la)
        if model == "FixedCorot" and dim == 3:
            e = e + 0.5 * la * (np.linalg.det(Fm) - 1.0) ** 2
        return e
    _, g, _ = ost.model_energy_gradient_hessian(model, F, mu, la)
    eps = 1e-6
    for k in range(M):
        dF = np.zeros((dim, dim)); dF[k % dim, k // dim] = eps
        fd = (psi(F + dF[None]) - psi(F - dF[None])) / (2 * eps)
        assert np.abs(fd - g[:, k]).max() <= 1e-7 * (mu + la)
