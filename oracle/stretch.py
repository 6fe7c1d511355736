"""ORACLE — TEST INFRASTRUCTURE ONLY (imported by tests/ only; never by the product).

numpy restatement of the reference's NON-mixed variable for tetrahedra and 2D triangles with the F-form overloads of
Stable Neo-Hookean, Neo-Hookean (3D), ARAP and FixedCorot (closed forms; tests/test_oracle_stretch.py pins them against
numbers evaluated from the reference's generated expression text, tests/golden/fform_golden.json):

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


def _cof_d2J(F):
    """(J, vec cof F (nE,M), d2J/dF2 (nE,M,M)) with index d*c + r (column-major vec F)"""
    nE, d = F.shape[0], F.shape[1]
    M = d * d
    J = np.linalg.det(F)
    cof = J[:, None, None] * np.transpose(np.linalg.inv(F), (0, 2, 1))
    vcof = np.transpose(cof, (0, 2, 1)).reshape(nE, M)
    if d == 3:
        d2 = np.einsum("rsm,cdn,emn->ecrds", _EPS, _EPS, F).reshape(nE, M, M)
    else:
        d2 = np.broadcast_to(np.einsum("rs,cd->crds", _EPS2, _EPS2).reshape(1, M, M), (nE, M, M))
    return J, vcof, d2


def neohookean_energy(F, mu, la):
    """neohookean.cpp:93-106: mu/2 (|F|^2 - 3) - mu log J + la/2 (log J)^2"""
    J = np.linalg.det(F)
    return 0.5 * mu * (np.einsum("eij,eij->e", F, F) - 3.0) - mu * np.log(J) + 0.5 * la * np.log(J) ** 2


def neohookean_gradient_hessian(F, mu, la):
    """closed form of neohookean.cpp:108-227 (81 symbolic-toolbox expressions), BEFORE sim::simple_psd_fix(H) (tol 1e-6):
    g = mu F + (la log J - mu)/J cof F ; H = mu I + (la (1 - log J) + mu)/J^2 cof cof^T + (la log J - mu)/J d2J/dF2"""
    nE = F.shape[0]
    J, vcof, d2 = _cof_d2J(F)
    lj = np.log(J)
    k1 = (la * lj - mu) / J
    k2 = (la * (1.0 - lj) + mu) / (J * J)
    vec = lambda A: np.transpose(A, (0, 2, 1)).reshape(nE, 9)
    g = mu * vec(F) + k1[:, None] * vcof
    H = mu * np.eye(9)[None] + k2[:, None, None] * np.einsum("ei,ej->eij", vcof, vcof) + k1[:, None, None] * d2
    return g, H


def volume_term(F, la):
    """the term FixedCorot adds to the ARAP part (fixed_corotational.cpp:233-254, :266-357; 2D: :160-169, :196-209):
    g += la (J - 1) cof F ; H += la (cof cof^T + (J - 1) d2J/dF2)"""
    J, vcof, d2 = _cof_d2J(F)
    return la * (J - 1.0)[:, None] * vcof, la * (np.einsum("ei,ej->eij", vcof, vcof) + (J - 1.0)[:, None, None] * d2)


def arap2d_parts(F):
    """R = U V^T with the flip fix of mfem::svd<double,2> (svd_eigen.h:7-31) and dR/dF = 2/(s0 + s1) q0 q0^T,
    q0 = vec(U [0 -1; 1 0] V^T)/sqrt 2 (arap.cpp:112-132): returns (R (nE,2,2), dRdF (nE,4,4), s0 + s1)"""
    U, sig, Vt = np.linalg.svd(F)
    flip = np.linalg.det(U @ Vt) < 0
    U = U.copy(); sig = sig.copy()
    U[flip, :, -1] *= -1.0
    sig[flip, -1] *= -1.0
    R = U @ Vt
    Sk = np.array([[0.0, -1.0], [1.0, 0.0]])
    Q0 = (U @ Sk[None] @ Vt) / np.sqrt(2.0)
    q0 = np.transpose(Q0, (0, 2, 1)).reshape(-1, 4)
    l0 = 2.0 / (sig[:, 0] + sig[:, 1])
    return R, l0[:, None, None] * np.einsum("ei,ej->eij", q0, q0), sig[:, 0] + sig[:, 1]


def model_energy_gradient_hessian(model, F, mu, la):
    """(psi (nE,), g (nE,M), H (nE,M,M) AFTER the model's psd fix) of the F-form overloads, d = 2 or 3:
    Stable-Neohookean (fix at 0), Neohookean (3D; fix at 1e-6), ARAP (no fix), FixedCorot (fix at 1e-6).
    FixedCorot quirks reproduced: the 3D energy has no volume term (fixed_corotational.cpp:215-223) and carries mu/2, the 2D
    one mu and la/2 (J-1)^2 (:136-151); the 3D gradient / Hessian use mu, the 2D ones 2 mu (:225-232 vs :153-172)."""
    nE, d = F.shape[0], F.shape[1]
    M = d * d
    vec = lambda A: np.transpose(A, (0, 2, 1)).reshape(nE, M)
    Fv = vec(F)
    if model == "Stable-Neohookean":
        psi = energy_expanded(Fv, mu, la) if d == 3 else energy_expanded_2d(Fv, mu, la)
        g, H = gradient_hessian(F, mu, la)
        return psi, g, simple_psd_fix(H, 0.0)
    if model == "Neohookean":
        assert d == 3
        g, H = neohookean_gradient_hessian(F, mu, la)
        return neohookean_energy(F, mu, la), g, simple_psd_fix(H, 1e-6)
    if d == 3:
        gA, HA, R = arap_gradient_hessian(F, mu)
        psiA = arap_energy(F, mu)
        if model == "ARAP":
            return psiA, gA, HA
        gv, Hv = volume_term(F, la)
        return psiA, gA + gv, simple_psd_fix(HA + Hv, 1e-6)
    R, dRdF, _ = arap2d_parts(F)
    diff2 = np.einsum("eij,eij->e", F - R, F - R)
    if model == "ARAP":
        return 0.5 * mu * diff2, mu * (Fv - vec(R)), mu * (np.eye(4)[None] - dRdF)
    assert model == "FixedCorot"
    J = np.linalg.det(F)
    gv, Hv = volume_term(F, la)
    return mu * diff2 + 0.5 * la * (J - 1.0) ** 2, 2.0 * mu * (Fv - vec(R)) + gv, simple_psd_fix(2.0 * mu * (np.eye(4)[None] - dRdF) + Hv, 1e-6)


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
    def __init__(self, V, T, free_map, mu=None, la=None, model="Stable-Neohookean", materials=None, mat_id=None):
        """one material (mu, la, model) or `materials` = [(model, mu, la), ...] with a per-element index `mat_id`
        (Mesh::elements_[i].material_, mesh.cpp:21-23: Stretch::update asks each element's own material, stretch.cpp:41-46)"""
        self.V, self.T, self.fm = np.asarray(V, float), np.asarray(T), np.asarray(free_map)
        self.materials = [(model, mu, la)] if materials is None else [tuple(m[:3]) for m in materials]
        self.mat_id = np.zeros(self.T.shape[0], dtype=np.int64) if mat_id is None else np.asarray(mat_id, dtype=np.int64)
        self.mu, self.la, self.model = self.materials[0][1], self.materials[0][2], self.materials[0][0]
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
        psi, _, _ = self._per_element(F)
        return float(np.sum(psi * self.vol))

    def _per_element(self, F):
        nE, d = F.shape[0], F.shape[1]
        M = d * d
        psi, g, H = np.zeros(nE), np.zeros((nE, M)), np.zeros((nE, M, M))
        for k, (model, mu, la) in enumerate(self.materials):
            sel = self.mat_id == k
            if sel.any():
                psi[sel], g[sel], H[sel] = model_energy_gradient_hessian(model, F[sel], mu, la)
        return psi, g, H

    def update(self, x, h):
        d = self.d
        nd = d * (d + 1)
        F = meshgen.deformation_gradients(self.V, self.T, np.asarray(x, float))
        _, g, H = self._per_element(F)
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
