"""CPU checks of the oracle's shell path (TriMesh + MixedStretch<3>, src/mesh/tri_mesh.cpp): nothing in the reference
pins it (cloth_sim.cpp is dead code), so the restatement is checked against independent derivations:
rest-state deformation gradient, finite differences of F(x) against the updated local Jacobian, and the assembled matrix
against an explicit dense J^T W K W J product."""
import numpy as np

from mixed_fem_b200 import meshgen
from oracle import orc
from tests.common import LA, MU


def _shell(n=(6, 4), thickness=1e-3):
    V, T, N = meshgen.shell_grid(*n)
    fm = meshgen.pinned_free_map(V, ratio=0.05)
    var = orc.OracleMixedStretch(V, T, fm, [("Stable-Neohookean", MU, LA)], normals=N, thickness=thickness)
    var.reset()
    return V, T, N, fm, var


def _F(var, x):
    """deformation gradients through the oracle: evaluate_constraint leaves S(F(x)) in grad-like form; use update + S/R"""
    var.update(x, 0.01)
    R = var.get("R").reshape(-1, 3, 3).transpose(0, 2, 1)
    S6 = var.get("S").reshape(-1, 6)
    S = np.zeros((S6.shape[0], 3, 3))
    S[:, 0, 0], S[:, 1, 1], S[:, 2, 2] = S6[:, 0], S6[:, 1], S6[:, 2]
    S[:, 0, 1] = S[:, 1, 0] = S6[:, 3]; S[:, 0, 2] = S[:, 2, 0] = S6[:, 4]; S[:, 1, 2] = S[:, 2, 1] = S6[:, 5]
    return R @ S


def _F_direct(V, T, N, x):
    """F = sum_k x_k (x) dphi_k + n(x) (x) N with dphi from the pseudo-inverse (linear_tri3d_dphi_dX.cpp:5-17)"""
    X = V[T]
    Tm = np.stack([X[:, 1] - X[:, 0], X[:, 2] - X[:, 0]], axis=2)             # (nE,3,2)
    pinv = np.linalg.inv(np.transpose(Tm, (0, 2, 1)) @ Tm) @ np.transpose(Tm, (0, 2, 1))   # (nE,2,3)
    dphi = np.concatenate([-pinv.sum(axis=1, keepdims=True), pinv], axis=1)   # (nE,3,3)
    xe = x.reshape(-1, 3)[T]
    F = np.einsum("ekr,ekc->erc", xe, dphi)
    n = np.cross(xe[:, 1] - xe[:, 0], xe[:, 2] - xe[:, 0]); n /= np.linalg.norm(n, axis=1, keepdims=True)
    return F + n[:, :, None] * N[:, None, :]


def test_rest_state_is_identity_and_volumes():
    V, T, N, fm, var = _shell()
    F = _F(var, V.reshape(-1))
    assert np.abs(F - np.eye(3)).max() < 1e-12          # tangent part + n (x) n = I at rest
    area = 0.5 * np.linalg.norm(np.cross(V[T[:, 1]] - V[T[:, 0]], V[T[:, 2]] - V[T[:, 0]]), axis=1)
    assert np.allclose(var.get("vols"), area * 1e-3, rtol=1e-13)


def test_deformation_gradient_matches_direct_formula():
    V, T, N, fm, var = _shell()
    rng = np.random.default_rng(0)
    x = (V @ meshgen._affine().T + 0.02 * rng.standard_normal(V.shape)).reshape(-1)
    # R sym(R^T F) reproduces F only to the polar tolerance (1e-6, newton_procrustes.h:108): structural check
    assert np.abs(_F(var, x) - _F_direct(V, T, N, x)).max() < 5e-6


def test_rhs_and_matrix_use_the_updated_jacobian():
    """rhs = -PJW (dSdF gl) and A = PJW K (PJW)^T with J = dF/dx at the CURRENT x (update_jacobian): J is checked by central
    differences of the direct F formula, the matrix by the dense product."""
    V, T, N, fm, var = _shell((4, 3))
    rng = np.random.default_rng(1)
    x = (V @ meshgen._affine().T + 0.02 * rng.standard_normal(V.shape)).reshape(-1)
    var.update(x, 0.02)
    nE, nV = T.shape[0], V.shape[0]
    J = np.zeros((9 * nE, 3 * nV))
    eps = 1e-6
    for i in range(3 * nV):
        xp, xm = x.copy(), x.copy(); xp[i] += eps; xm[i] -= eps
        d = (_F_direct(V, T, N, xp) - _F_direct(V, T, N, xm)) / (2 * eps)       # (nE, r, c)
        J[:, i] = np.transpose(d, (0, 2, 1)).reshape(-1)                         # column-major per element
    free = np.nonzero(fm >= 0)[0]
    order = np.argsort(fm[free]); free = free[order]
    fd = (3 * free[:, None] + np.arange(3)[None, :]).reshape(-1)
    vols = var.get("vols")
    W = np.repeat(vols, 9)
    PJW = (J[:, fd].T) * W[None, :]
    dSdF = var.get("dSdF").reshape(nE, 6, 9)                                     # column-major 9 x 6 per element
    H = var.get("H").reshape(nE, 6, 6)
    K = np.zeros((9 * nE, 9 * nE))
    for e in range(nE):
        D = dSdF[e].T                                                            # 9 x 6
        K[9 * e:9 * e + 9, 9 * e:9 * e + 9] = D @ H[e].T @ D.T
    A_dense = PJW @ K @ PJW.T
    outer, inner, vals = var.csr()
    import scipy.sparse as sp
    A = sp.csr_matrix((vals, inner, outer), shape=(fd.size, fd.size)).toarray()
    assert np.linalg.norm(A - A_dense) <= 1e-6 * np.linalg.norm(A_dense)        # FD Jacobian: 1e-6 is the FD accuracy
    gl_like = var.rhs()
    gl = var.get("gl").reshape(nE, 6)
    t = np.einsum("enm,en->em", dSdF, gl).reshape(-1)
    assert np.linalg.norm(gl_like + PJW @ t) <= 1e-6 * np.linalg.norm(gl_like)
