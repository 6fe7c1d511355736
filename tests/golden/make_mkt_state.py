"""Recovers the INPUT STATE behind the reference's only numeric artefact, the 2D KKT dump scripts/mats/*.mkt
(written by the commented saveMarket lines of src/optimizers/newton_optimizer.cpp:142-156 / linear_system.h:174-209;
the dump does not record x, s, the material or the mesh), so that the dump becomes an END-TO-END golden vector for
the 2D mixed-stretch path.  Found by this script (every step over-determined, residuals printed and stored):

  mesh ........ models/obj/square_xy_145.obj (145 V / 256 triangles; doublearea 0.03125 = the diagonal of lhs_Gs)
  BC .......... both end groups pinned (ratio 0.1, axis 0: the stretch / bend / twist scripts) - the sparsity of
                lhs_Gx equals jacobian_x() of exactly this free_map, and lhs_M equals P M P^T bit for bit
  material .... FixedCorot, E = 1e6, nu = 0.45 (the MaterialConfig defaults), h = 0.0333, density 1000:
                lhs_K (= Hloc_ = vol h^2 d2psi/ds2) and rhs_gs (= -vol h^2 dpsi/ds) of ALL 256 elements are reproduced
                by one s_e (3 unknowns for 9 numbers per element) to 4e-16 relative
  S(F_e) ...... = s_e - rhs_gl_e / (Sym vol)      (rhs_gl = -gradient_dual() = -Sym (S(F) - s) vol)
  rotation .... one angle per element (1 unknown for up to 18 numbers) from its block of lhs_Gx
                (= -P J^T W blockdiag(dSdF_)): 1e-15 relative
  x ........... least-squares integration of F_e = R(theta_e) S_e over the mesh (exactly compatible: residual 1e-15)

Output: tests/golden/mkt_state.npz (V, T, free_map, x, s, material constants, fit residuals).  The tests then run
update(x) on the oracle / the CUDA path and compare Hloc_, jacobian_x(), jacobian_mixed(), gradient_mixed(),
gradient_dual() and P M P^T with the dumped matrices.

Run here (needs scipy and the oracle):  python tests/golden/make_mkt_state.py
"""
import os
import sys

import numpy as np
import scipy.optimize as so
import scipy.sparse as sp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import mixed_fem_b200  # noqa: E402,F401
from mixed_fem_b200 import meshgen, sceneio  # noqa: E402
from oracle import orc  # noqa: E402

RI = os.path.join(HERE, "ref_inputs")
H, VOL, E, NU, RHO = 0.0333, 0.03125, 1e6, 0.45, 1000.0
SYM = np.array([1.0, 1.0, 2.0])


def load(name):
    r, c, v, sh = sceneio.read_mkt(os.path.join(RI, "scripts/mats/%s.mkt" % name))
    return sp.coo_matrix((v, (r, c)), shape=sh).tocsr()


def main():
    K, Gx = load("lhs_K"), load("lhs_Gx").tocsc()
    rgs = load("rhs_gs").toarray().reshape(-1, 3)
    rgl = load("rhs_gl").toarray().reshape(-1, 3)
    V, T = sceneio.read_obj(os.path.join(RI, "models/obj/square_xy_145.obj"))
    V, T = sceneio.remove_unreferenced(V, T)
    V = np.ascontiguousarray(V[:, :2])
    fm = sceneio.free_map_from_fixed(sceneio.fixed_vertices(V, "stretch", 0.1, 0))
    nE = T.shape[0]
    mu, la = meshgen.enu_to_lame(E, NU)
    iu = np.triu_indices(3)
    X = V[T]
    Dm = np.transpose(X[:, 1:, :] - X[:, :1, :], (0, 2, 1))
    inv = np.linalg.inv(Dm)
    dphi = np.concatenate([-inv.sum(axis=1, keepdims=True), inv], axis=1)
    s_fit = np.zeros((nE, 3)); SF = np.zeros((nE, 3)); theta = np.full(nE, np.nan)
    res_s = np.zeros(nE); res_g = np.zeros(nE)
    for e in range(nE):
        Ke = K[3 * e:3 * e + 3, 3 * e:3 * e + 3].toarray()

        def rs(s):
            _, g, Hm = orc.material("FixedCorot", mu, la, s)
            return np.concatenate([(H * H * VOL * Hm)[iu] - Ke[iu], -H * H * VOL * g - rgs[e]])
        r = so.least_squares(rs, [1.0, 1.0, 0.0], xtol=1e-15, ftol=1e-15, gtol=1e-15)
        s_fit[e] = r.x
        res_s[e] = np.abs(r.fun).max() / np.abs(Ke).max()
        SF[e] = s_fit[e] - rgl[e] / (SYM * VOL)
        Sm = np.array([[SF[e, 0], SF[e, 2]], [SF[e, 2], SF[e, 1]]])
        rows = []
        for k in range(3):
            f = fm[T[e, k]]
            if f >= 0:
                for l in range(2):
                    rows.append(Gx[2 * f + l, 3 * e:3 * e + 3].toarray().ravel())
        if not rows:
            continue
        data = np.array(rows)

        def gx_model(th):
            c, sn = np.cos(th), np.sin(th)
            Fm = np.array([[c, -sn], [sn, c]]) @ Sm
            _, _, dsdF = orc.polar_svd2(Fm)
            dSdF = dsdF.T * SYM[None, :]
            out = []
            for k in range(3):
                if fm[T[e, k]] < 0:
                    continue
                for l in range(2):
                    Jt = np.zeros(4)
                    for c_ in range(2):
                        Jt[2 * c_ + l] = dphi[e, k, c_]
                    out.append(-VOL * (Jt @ dSdF))
            return np.array(out)
        best = None
        for t0 in np.linspace(-np.pi, np.pi, 9):
            r2 = so.least_squares(lambda t: (gx_model(t[0]) - data).ravel(), [t0], xtol=1e-15, ftol=1e-15, gtol=1e-15)
            if best is None or r2.cost < best.cost:
                best = r2
        theta[e] = np.arctan2(np.sin(best.x[0]), np.cos(best.x[0]))
        res_g[e] = np.abs(best.fun).max() / np.abs(data).max()
    # integrate F_e = R(theta_e) S_e: rows of  (x_k - x_0) = F_e (X_k - X_0), k = 1,2, for the elements with a known angle
    known = np.nonzero(~np.isnan(theta))[0]
    rows, cols, vals, rhs = [], [], [], []
    q = 0
    for e in known:
        c, sn = np.cos(theta[e]), np.sin(theta[e])
        Fm = np.array([[c, -sn], [sn, c]]) @ np.array([[SF[e, 0], SF[e, 2]], [SF[e, 2], SF[e, 1]]])
        for k in (1, 2):
            d = Fm @ (V[T[e, k]] - V[T[e, 0]])
            for l in range(2):
                rows += [q, q]; cols += [2 * T[e, k] + l, 2 * T[e, 0] + l]; vals += [1.0, -1.0]; rhs.append(d[l]); q += 1
    # gauge: translation fixed by the centroid of the rest shape
    for l in range(2):
        for v in range(V.shape[0]):
            rows.append(q); cols.append(2 * v + l); vals.append(1.0 / V.shape[0])
        rhs.append(V[:, l].mean()); q += 1
    A = sp.coo_matrix((vals, (rows, cols)), shape=(q, 2 * V.shape[0])).tocsr()
    b = np.array(rhs)
    x = np.linalg.lstsq(A.toarray(), b, rcond=None)[0]
    res_x = float(np.abs(A @ x - b).max())
    # the 16 elements without a free vertex were not used: their S(F) from x must still match the dump
    Fall = meshgen.deformation_gradients(V, T, x)
    _, Sall = meshgen.exact_polar(Fall)
    res_S = float(np.abs(meshgen.sym_to_vec(Sall) - SF).max())
    print("fit residuals: K/rhs_gs %.2e  Gx %.2e  integration of x %.2e  S(F(x)) vs dump, all elements %.2e" %
          (res_s.max(), np.nanmax(res_g), res_x, res_S))
    assert res_s.max() < 1e-13 and np.nanmax(res_g) < 1e-12 and res_x < 1e-12 and res_S < 1e-11
    np.savez_compressed(os.path.join(HERE, "mkt_state.npz"), V=V, T=T.astype(np.int32), free_map=fm, x=x, s=s_fit.reshape(-1),
                        theta=theta, h=H, E=E, nu=NU, density=RHO, model="FixedCorot",
                        fit_residuals=np.array([res_s.max(), np.nanmax(res_g), res_x, res_S]))


if __name__ == "__main__":
    main()
