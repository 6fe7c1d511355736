"""ORACLE — TEST INFRASTRUCTURE ONLY (imported by tests/ only; never by the product).

numpy restatement of the reference's NON-mixed variable for tetrahedra and 2D triangles with the Stable Neo-Hookean
F-form:

  Stretch<DIM>::energy / update / rhs / gradient ..... src/variables/stretch.cpp:16-68
  StableNeohookean::energy/gradient/hessian(Vector9d)  src/energies/stable_neohookean.cpp:93-232 (Vector4d: :274-324)
  sim::simple_psd_fix(H, 0.0) ........................ deps/Bartels/src/simple_psd_fix.cpp:5-22
  local Jacobian Jloc = dphi^T (x) I_3 ............... src/mesh/tet_mesh.cpp:43-75
  Assembler / VecAssembler ........................... src/utils/sparse_utils.h:11-120 (sums of element blocks at free dofs)

The reference's gradient and Hessian are symbolic-toolbox expansions (81 polynomials); `energy_expanded` transcribes the
energy expression of stable_neohookean.cpp:107 term by term, and tests/test_oracle_stretch.py checks the closed forms
used here against its finite differences.  PARITY UNPINNED otherwise (no fixtures in the reference).
"""
import numpy as np
import scipy.sparse as sp

from mixed_fem_b200 import meshgen


def energy_expanded(F9, mu, la):
    """stable_neohookean.cpp:93-107 verbatim (F9 column-major: F(0)=F11, F(1)=F21, F(2)=F31, F(3)=F12, ...)"""
    F1_1, F2_1, F3_1, F1_2, F2_2, F3_2, F1_3, F2_3, F3_3 = [F9[..., i] for i in range(9)]
    t = -F1_1 * F2_2 * F3_3 + F1_1 * F2_3 * F3_2 + F1_2 * F2_1 * F3_3 - F1_2 * F2_3 * F3_1 - F1_3 * F2_1 * F3_2 + F1_3 * F2_2 * F3_1 + 1.0
    return mu * t + (la * t ** 2) / 2.0 + (mu * (F1_1 ** 2 + F1_2 ** 2 + F1_3 ** 2 + F2_1 ** 2 + F2_2 ** 2 + F2_3 ** 2 + F3_1 ** 2 + F3_2 ** 2 + F3_3 ** 2 - 3.0)) / 2.0


_EPS = np.zeros((3, 3, 3))
for _i, _j, _k in ((0, 1, 2), (1, 2, 0), (2, 0, 1)):
    _EPS[_i, _j, _k] = 1.0
    _EPS[_j, _i, _k] = -1.0


_EPS2 = np.array([[0.0, 1.0], [-1.0, 0.0]])


def energy_expanded_2d(F4, mu, la):
    """stable_neohookean.cpp:274-282 verbatim (F4 column-major: F11, F21, F12, F22)"""
    F1_1, F2_1, F1_2, F2_2 = [F4[..., i] for i in range(4)]
    t = -F1_1 * F2_2 + F1_2 * F2_1 + 1.0
    return mu * t + (la * t ** 2) / 2.0 + (mu * (F1_1 * F1_1 + F1_2 * F1_2 + F2_1 * F2_1 + F2_2 * F2_2 - 2.0)) / 2.0


def gradient_hessian(F, mu, la):
    """F (nE,d,d) -> g (nE,d*d), H (nE,d*d,d*d), index d*c + r (column-major vec F), BEFORE the psd fix"""
    nE, d = F.shape[0], F.shape[1]
    M = d * d
    J = np.linalg.det(F)
    cof = J[:, None, None] * np.transpose(np.linalg.inv(F), (0, 2, 1))        # dJ/dF
    vec = lambda A: np.transpose(A, (0, 2, 1)).reshape(nE, M)
    k1 = la * (J - 1.0) - mu
    g = mu * vec(F) + k1[:, None] * vec(cof)
    if d == 3:      # d2J/dF(r,c)dF(s,c2) = eps(r,s,m) eps(c,c2,n) F(m,n)
        d2 = np.einsum("rsm,cdn,emn->ecrds", _EPS, _EPS, F).reshape(nE, M, M)      # [3c+r, 3c2+s]
    else:           # J = F00 F11 - F01 F10
        d2 = np.broadcast_to(np.einsum("rs,cd->crds", _EPS2, _EPS2).reshape(1, M, M), (nE, M, M))
    H = mu * np.eye(M)[None] + la * np.einsum("ei,ej->eij", vec(cof), vec(cof)) + k1[:, None, None] * d2
    return g, H


def polar_rotation(F):
    """R of F = R S (S symmetric, R in SO(3)): what iARAP::rotation returns (src/svd/iARAP.cpp), via numpy's SVD"""
    U, sig, Vt = np.linalg.svd(F)
    flip = np.linalg.det(U @ Vt) < 0
    U = U.copy()
    U[flip, :, -1] *= -1.0
    return U @ Vt


def arap_gradient_hessian(F, mu):
    """src/energies/arap.cpp:143-162: g = mu (F - R), H = mu (I - dR/dF), NO psd fix.  dR/dF in closed form:
    column (r,c): w = (tr(Y) I - Y)^-1 (r_c x e_r), Y = R F^T, dR(:,b) = w x r_b  (checked against finite differences
    of polar_rotation in tests/test_oracle_stretch.py)"""
    nE = F.shape[0]
    R = polar_rotation(F)
    Y = R @ np.transpose(F, (0, 2, 1))
    Hp = np.trace(Y, axis1=1, axis2=2)[:, None, None] * np.eye(3)[None] - 0.5 * (Y + np.transpose(Y, (0, 2, 1)))
    Hpi = np.linalg.inv(Hp)
    dRdF = np.zeros((nE, 9, 9))
    for c in range(3):
        for r in range(3):
            er = np.zeros(3); er[r] = 1.0
            a = np.cross(R[:, :, c], er[None, :])
            w = np.einsum("eij,ej->ei", Hpi, a)
            for b in range(3):
                dRdF[:, 3 * b:3 * b + 3, 3 * c + r] = np.cross(w, R[:, :, b])
    vec = lambda A: np.transpose(A, (0, 2, 1)).reshape(nE, 9)
    g = mu * (vec(F) - vec(R))
    H = mu * (np.eye(9)[None] - dRdF)
    return g, H, R


def arap_energy(F, mu):
    """arap.cpp:134-141: mu/2 (|F|^2 - 2 tr S + 3)"""
    R = polar_rotation(F)
    trS = np.einsum("eij,eij->e", R, F)
    return 0.5 * mu * (np.einsum("eij,eij->e", F, F) - 2.0 * trS + 3.0)


def simple_psd_fix(H, tol=0.0):
    w, V = np.linalg.eigh(H)
    w = np.where(w < tol, tol, w)
    return np.einsum("eik,ek,ejk->eij", V, w, V)


class OracleStretch:
    def __init__(self, V, T, free_map, mu, la, model="Stable-Neohookean"):
        self.V, self.T, self.fm = np.asarray(V, float), np.asarray(T), np.asarray(free_map)
        self.mu, self.la, self.model = mu, la, model
        self.d = d = self.V.shape[1]
        X = self.V[self.T]
        Dm = np.transpose(X[:, 1:, :] - X[:, :1, :], (0, 2, 1))
        inv = np.linalg.inv(Dm)
        self.dphi = np.concatenate([-inv.sum(axis=1, keepdims=True), inv], axis=1)      # (nE,d+1,d), linear_*_dphi_dX.cpp
        # |igl::volume| for tets (tet_mesh.cpp:31-34), igl::doublearea for 2D triangles (tri2d_mesh.cpp:32-34)
        self.vol = np.abs(np.linalg.det(Dm)) / 6.0 if d == 3 else np.linalg.det(Dm)
        self.nF = int(self.fm.max()) + 1
        # Jloc (nE, d*d, d*(d+1)): dF(d c + r) / dx(d k + r2) = dphi(k,c) delta(r,r2)
        nE = self.T.shape[0]
        self.Jloc = np.einsum("ekc,rs->ecrks", self.dphi, np.eye(d)).reshape(nE, d * d, d * (d + 1))

    def energy(self, x):
        d = self.d
        F = meshgen.deformation_gradients(self.V, self.T, np.asarray(x, float))
        Fv = np.transpose(F, (0, 2, 1)).reshape(-1, d * d)
        if self.model == "ARAP":
            return float(np.sum(arap_energy(F, self.mu) * self.vol))
        psi = energy_expanded(Fv, self.mu, self.la) if d == 3 else energy_expanded_2d(Fv, self.mu, self.la)
        return float(np.sum(psi * self.vol))

    def update(self, x, h):
        d = self.d
        nd = d * (d + 1)
        F = meshgen.deformation_gradients(self.V, self.T, np.asarray(x, float))
        if self.model == "ARAP":
            g, H, _ = arap_gradient_hessian(F, self.mu)
        else:
            g, H = gradient_hessian(F, self.mu, self.la)
            H = simple_psd_fix(H, 0.0)
        w = self.vol * h * h
        He = np.einsum("eia,eij,ejb->eab", self.Jloc, H, self.Jloc) * w[:, None, None]      # (nE,nd,nd)
        ge = np.einsum("eia,ei->ea", self.Jloc, g) * w[:, None]                              # (nE,nd)
        dofs = (d * self.fm[self.T][:, :, None] + np.arange(d)[None, None, :]).reshape(-1, nd)
        free = (self.fm[self.T] >= 0)[:, :, None].repeat(d, axis=2).reshape(-1, nd)
        rows = np.broadcast_to(dofs[:, :, None], He.shape)
        cols = np.broadcast_to(dofs[:, None, :], He.shape)
        keep = free[:, :, None] & free[:, None, :]
        n = d * self.nF
        self.lhs = sp.coo_matrix((He[keep], (rows[keep], cols[keep])), shape=(n, n)).tocsr()
        self.grad = np.bincount(dofs[free], weights=ge[free], minlength=n)

    def rhs(self):
        return -self.grad
