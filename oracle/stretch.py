"""ORACLE — TEST INFRASTRUCTURE ONLY (imported by tests/ only; never by the product).

numpy restatement of the reference's NON-mixed variable for tetrahedra with the Stable Neo-Hookean F-form:

  Stretch<DIM>::energy / update / rhs / gradient ..... src/variables/stretch.cpp:16-68
  StableNeohookean::energy/gradient/hessian(Vector9d)  src/energies/stable_neohookean.cpp:93-232
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


def gradient_hessian(F, mu, la):
    """F (nE,3,3) -> g (nE,9), H (nE,9,9), index 3c + r (column-major vec F), BEFORE the psd fix"""
    nE = F.shape[0]
    J = np.linalg.det(F)
    cof = J[:, None, None] * np.transpose(np.linalg.inv(F), (0, 2, 1))        # dJ/dF
    vec = lambda A: np.transpose(A, (0, 2, 1)).reshape(nE, 9)
    k1 = la * (J - 1.0) - mu
    g = mu * vec(F) + k1[:, None] * vec(cof)
    # d2J/dF(r,c)dF(s,c2) = eps(r,s,m) eps(c,c2,n) F(m,n)
    d2 = np.einsum("rsm,cdn,emn->ecrds", _EPS, _EPS, F).reshape(nE, 9, 9)       # [3c+r, 3d+s]
    H = mu * np.eye(9)[None] + la * np.einsum("ei,ej->eij", vec(cof), vec(cof)) + k1[:, None, None] * d2
    return g, H


def simple_psd_fix(H, tol=0.0):
    w, V = np.linalg.eigh(H)
    w = np.where(w < tol, tol, w)
    return np.einsum("eik,ek,ejk->eij", V, w, V)


class OracleStretch:
    def __init__(self, V, T, free_map, mu, la):
        self.V, self.T, self.fm = np.asarray(V, float), np.asarray(T), np.asarray(free_map)
        self.mu, self.la = mu, la
        self.dphi = meshgen.tet_dphi(self.V, self.T)                              # (nE,4,3)
        X = self.V[self.T]
        self.vol = np.abs(np.linalg.det(X[:, 1:, :] - X[:, :1, :])) / 6.0          # |igl::volume| (tet_mesh.cpp:31-34)
        self.nF = int(self.fm.max()) + 1
        # Jloc (nE, 9, 12): dF(3c+r) / dx(3k+r2) = dphi(k,c) delta(r,r2)
        nE = self.T.shape[0]
        self.Jloc = np.einsum("ekc,rs->ecrks", self.dphi, np.eye(3)).reshape(nE, 9, 12)

    def energy(self, x):
        F = meshgen.deformation_gradients(self.V, self.T, np.asarray(x, float))
        F9 = np.transpose(F, (0, 2, 1)).reshape(-1, 9)
        return float(np.sum(energy_expanded(F9, self.mu, self.la) * self.vol))

    def update(self, x, h):
        F = meshgen.deformation_gradients(self.V, self.T, np.asarray(x, float))
        g, H = gradient_hessian(F, self.mu, self.la)
        H = simple_psd_fix(H, 0.0)
        w = self.vol * h * h
        He = np.einsum("eia,eij,ejb->eab", self.Jloc, H, self.Jloc) * w[:, None, None]      # (nE,12,12)
        ge = np.einsum("eia,ei->ea", self.Jloc, g) * w[:, None]                              # (nE,12)
        dofs = (3 * self.fm[self.T][:, :, None] + np.arange(3)[None, None, :]).reshape(-1, 12)
        free = (self.fm[self.T] >= 0)[:, :, None].repeat(3, axis=2).reshape(-1, 12)
        rows = np.broadcast_to(dofs[:, :, None], He.shape)
        cols = np.broadcast_to(dofs[:, None, :], He.shape)
        keep = free[:, :, None] & free[:, None, :]
        n = 3 * self.nF
        self.lhs = sp.coo_matrix((He[keep], (rows[keep], cols[keep])), shape=(n, n)).tocsr()
        self.grad = np.bincount(dofs[free], weights=ge[free], minlength=n)

    def rhs(self):
        return -self.grad
