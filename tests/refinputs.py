"""Helpers for the tests that run on the REFERENCE'S OWN input files (copies under tests/golden/ref_inputs, made by
tests/golden/make_ref_inputs.py) and on its numeric matrix dump (scripts/mats/*.mkt + tests/golden/mkt_state.npz)."""
import os

import numpy as np
import scipy.sparse as sp

from mixed_fem_b200 import meshgen, sceneio

HERE = os.path.dirname(os.path.abspath(__file__))
RI = os.path.join(HERE, "golden", "ref_inputs")
SCENES_ROOT = os.path.join(RI, "scenes")        # '../models/...' of the scene files resolves from here


def scene(rel, dim, **kw):
    return sceneio.load_scene(os.path.join(RI, rel), dim, root=SCENES_ROOT, **kw)


def seeded_state(V, T, x, mu, seed=20221101, back=0.05, normals=None):
    """R / s / la of a 'previous' iterate for the deformed positions x: the polar factors of F at x pulled `back` of
    the way towards the rest shape plus a small jitter, multipliers iid U(-1,1) mu 1e-3 (the recipe of
    meshgen.synthetic_state, SURVEY.md section 8d, on a real mesh)."""
    rng = np.random.default_rng(seed)
    d = V.shape[1]
    X0 = V.reshape(-1)
    edge = np.linalg.norm(V[T[:, 1]] - V[T[:, 0]], axis=1).mean()
    xprev = x + back * (X0 - x) + rng.uniform(-0.01, 0.01, size=x.shape) * edge
    if normals is None:
        F = meshgen.deformation_gradients(V, T, xprev)
    else:
        F = shell_deformation_gradients(V, T, normals, xprev)
    R, S = meshgen.exact_polar(F)
    s = meshgen.sym_to_vec(S)
    la = rng.uniform(-1.0, 1.0, size=s.shape) * mu * 1e-3
    Rcm = np.transpose(R, (0, 2, 1)).reshape(T.shape[0], -1)
    return {"x": x, "R": np.ascontiguousarray(Rcm).reshape(-1), "s": s.reshape(-1), "la": la.reshape(-1)}


def shell_deformation_gradients(V, T, N, x):
    """F = J0 x + n(x) (x) N of TriMesh::deformation_gradient (src/mesh/tri_mesh.cpp:104-126), numpy"""
    X = V[T]
    xe = x.reshape(-1, 3)[T]
    Tm = np.stack([X[:, 1] - X[:, 0], X[:, 2] - X[:, 0]], axis=2)          # (nE,3,2)
    G = np.linalg.inv(np.transpose(Tm, (0, 2, 1)) @ Tm) @ np.transpose(Tm, (0, 2, 1))   # (nE,2,3): rows 1..2 of dphi
    dphi = np.concatenate([-G.sum(axis=1, keepdims=True), G], axis=1)      # (nE,3,3)
    F = np.einsum("ekr,ekc->erc", xe, dphi)
    n = np.cross(xe[:, 1] - xe[:, 0], xe[:, 2] - xe[:, 0])
    n /= np.linalg.norm(n, axis=1, keepdims=True)
    return F + np.einsum("er,ec->erc", n, N)


def mkt(name):
    r, c, v, sh = sceneio.read_mkt(os.path.join(RI, "scripts", "mats", name + ".mkt"))
    return sp.coo_matrix((v, (r, c)), shape=sh).tocsr()


def mkt_state():
    z = np.load(os.path.join(HERE, "golden", "mkt_state.npz"))
    mu, la = meshgen.enu_to_lame(float(z["E"]), float(z["nu"]))
    return {"V": z["V"], "T": z["T"], "fm": z["free_map"], "x": z["x"], "s": z["s"], "h": float(z["h"]),
            "materials": [(str(z["model"]), mu, la)], "density": float(z["density"]), "fit_residuals": z["fit_residuals"]}


def check_against_mkt_dump(var, st, jac_triplets=None, tol=1e-10):
    """var: a MixedStretch (oracle or CUDA mirror) after set('s'), update(x, h) on the mkt state.  Compares with the
    matrices the REFERENCE wrote: lhs_K = hessian() = blockdiag(Hloc_), lhs_Gs = jacobian_mixed() = blockdiag(vol Sym),
    lhs_Gx = jacobian_x(), rhs_gs = -gradient_mixed(), rhs_gl = -gradient_dual().  Returns the relative errors."""
    nE = st["T"].shape[0]
    K, Gs, Gx = mkt("lhs_K"), mkt("lhs_Gs"), mkt("lhs_Gx")
    el = np.arange(nE)
    Hloc = var.get("Hloc").reshape(nE, 3, 3)          # column-major per element, symmetric
    Kref = np.stack([np.stack([np.asarray(K[3 * el + r, 3 * el + c]).ravel() for c in range(3)], axis=1) for r in range(3)], axis=1)
    errs = {"K": float(np.abs(Hloc - Kref).max() / np.abs(Kref).max())}
    vols = var.get("vols")
    gs_diag = (vols[:, None] * np.array([1.0, 1.0, 2.0])[None, :]).reshape(-1)
    errs["Gs"] = float(np.abs(gs_diag - Gs.diagonal()).max() / np.abs(Gs.diagonal()).max())
    assert Gs.nnz == 9 * nE and abs(Gs - sp.diags(Gs.diagonal())).max() == 0.0     # dense 3x3 blocks with explicit zeros
    errs["rhs_gs"] = float(np.abs(-var.get("grad") - mkt("rhs_gs").toarray().ravel()).max() / np.abs(mkt("rhs_gs").toarray()).max())
    rgl = mkt("rhs_gl").toarray().ravel()
    # gradient_dual() ~ 1e-5 (a nearly converged step): compare on the scale of Sym vol |s|
    errs["rhs_gl"] = float(np.abs(-var.get("grad_la") - rgl).max() / np.abs(gs_diag).max())
    if jac_triplets is None:
        n = Gx.shape[0]
        cols = []
        for j in range(3):                            # the product does not mix elements: column j of every block at once
            e = np.zeros(3 * nE); e[j::3] = 1.0
            cols.append(e)
        B = sp.lil_matrix(Gx.shape)
        for j in range(3 * nE):
            e = np.zeros(3 * nE); e[j] = 1.0
            B[:, j] = var.product("jacobian_x", e, np.zeros(n), transposed=True).reshape(-1, 1)
        B = B.tocsr()
    else:
        r, c, v = jac_triplets
        B = sp.coo_matrix((v, (r, c)), shape=Gx.shape).tocsr()
    errs["Gx"] = float(abs(B - Gx).max() / abs(Gx).max())
    pat_ref = set(zip(*Gx.nonzero()))
    Bz = B.copy(); Bz.eliminate_zeros()
    errs["Gx_pattern_extra"] = len(set(zip(*Bz.nonzero())) - pat_ref)
    bad = {k: v for k, v in errs.items() if not v <= tol}
    assert not bad, (bad, errs)
    return errs
