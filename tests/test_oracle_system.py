"""Mesh-level CPU checks of the oracle: assembler pattern/values against scipy.sparse, the Schur-complement rhs()
and solve() against an explicit KKT elimination (mirrors scripts/subspace_solver.m:36-60 and
src/linear_solvers/linear_system.h:120-239), and F = J x against numpy."""
import numpy as np
import pytest
import scipy.sparse as sp

import mixed_fem_b200 as mf
from mixed_fem_b200 import meshgen
from oracle import orc

MU, LA = meshgen.enu_to_lame(1e6, 0.45)


def make(nx=4, ny=3, nz=2, model="Stable-Neohookean", dim=3):
    if dim == 3:
        V, T = meshgen.kuhn_grid(nx, ny, nz)
    else:
        V, T = meshgen.tri_grid(nx, ny)
    fm = meshgen.pinned_free_map(V, ratio=0.01)
    st = meshgen.synthetic_state(V, T, 1.0 / nx, MU)
    o = orc.OracleMixedStretch(V, T, fm, [(model, MU, LA)])
    o.reset()
    if dim == 3:
        o.set("R", st["R"])
    o.set("s", st["s"]); o.set("la", st["la"])
    return V, T, fm, st, o


def test_free_map_and_grid():
    V, T = meshgen.kuhn_grid(3, 2, 2)
    assert T.shape == (72, 4) and V.shape == (36, 3)
    fm = meshgen.pinned_free_map(V)
    assert (fm[V[:, 0] == 0] == -1).all()
    free = fm[fm >= 0]
    assert np.array_equal(free, np.arange(free.size))     # ascending in vertex order (mesh.cpp:95-101)


@pytest.mark.parametrize("dim", [3, 2])
def test_assembler_pattern_matches_scipy(dim):
    V, T, fm, st, o = make(dim=dim)
    outer, inner, _ = o.csr(values=False)
    npe = dim + 1
    rows, cols = [], []
    for e in range(T.shape[0]):
        for j in range(npe):
            for k in range(npe):
                gi, gj = fm[T[e, j]], fm[T[e, k]]
                if gi < 0 or gj < 0:
                    continue
                for a in range(dim):
                    for c in range(dim):
                        rows.append(dim * gi + a); cols.append(dim * gj + c)
    n = dim * (fm.max() + 1)
    A = sp.csr_matrix((np.ones(len(rows)), (rows, cols)), shape=(n, n))
    A.sum_duplicates(); A.sort_indices()
    assert np.array_equal(outer, A.indptr)
    assert np.array_equal(inner, A.indices)


@pytest.mark.parametrize("dim,model", [(3, "Stable-Neohookean"), (3, "ARAP"), (2, "Stable-Neohookean")])
def test_assembled_matrix_and_scatter_map(dim, model):
    V, T, fm, st, o = make(dim=dim, model=model)
    o.update(st["x"], 0.0333)
    outer, inner, vals = o.csr()
    npe, ld = dim + 1, dim * (dim + 1)
    Aloc = o.get("Aloc").reshape(T.shape[0], ld, ld).transpose(0, 2, 1)   # column-major blocks -> [e][r][c]
    smap, stride = o.scatter_map()
    acc = np.zeros_like(vals)
    for e in range(T.shape[0]):
        for j in range(npe):
            for k in range(npe):
                if smap[e, j, k] < 0:
                    assert fm[T[e, j]] < 0 or fm[T[e, k]] < 0
                    continue
                for a in range(dim):
                    for c in range(dim):
                        acc[smap[e, j, k] + a * stride[e, j] + c] += Aloc[e, dim * j + a, dim * k + c]
    assert np.linalg.norm(acc - vals) <= 1e-13 * np.linalg.norm(vals)
    # symmetric
    n = outer.size - 1
    A = sp.csr_matrix((vals, inner, outer), shape=(n, n))
    assert abs(A - A.T).max() <= 1e-12 * abs(A).max()


def _explicit_operators(o, V, T, fm, dim):
    """Gx = -PJW blockdiag(dSdF), K = blockdiag(Hloc), Gs = blockdiag(vol Sym) as dense matrices."""
    nE, N, M = T.shape[0], o.N, o.M
    nf = dim * o.nfree
    dSdF = o.get("dSdF").reshape(nE, N, M).transpose(0, 2, 1)     # [e][m][n]
    dphi = o.get("dphi").reshape(nE, dim + 1, dim)
    vol = o.get("vols")
    PJW = np.zeros((nf, M * nE))
    for e in range(nE):
        for k in range(dim + 1):
            f = fm[T[e, k]]
            if f < 0:
                continue
            for c in range(dim):
                for r in range(dim):
                    PJW[dim * f + r, M * e + dim * c + r] += vol[e] * dphi[e, k, c]
    D = np.zeros((M * nE, N * nE))
    for e in range(nE):
        D[M * e:M * e + M, N * e:N * e + N] = dSdF[e]
    return PJW, D, vol


@pytest.mark.parametrize("dim", [3, 2])
def test_rhs_and_solve_are_the_schur_complement(dim):
    V, T, fm, st, o = make(nx=3, ny=2, nz=2, dim=dim)
    o.update(st["x"], 0.0333)
    rhs = o.rhs()
    nE, N = T.shape[0], o.N
    PJW, D, vol = _explicit_operators(o, V, T, fm, dim)
    sym = np.array([1, 1, 1, 2, 2, 2.0]) if dim == 3 else np.array([1, 1, 2.0])
    H_ = o.get("H").reshape(nE, N, N)
    g = o.get("g").reshape(nE, N)
    S = o.get("S").reshape(nE, N)
    s = o.get("s").reshape(nE, N)
    # lhs == PJW D blockdiag(H_) D^T PJW^T
    Hb = np.zeros((N * nE, N * nE))
    for e in range(nE):
        Hb[N * e:N * e + N, N * e:N * e + N] = H_[e]
    outer, inner, vals = o.csr()
    n = outer.size - 1
    A = sp.csr_matrix((vals, inner, outer), shape=(n, n)).toarray()
    Aref = PJW @ D @ Hb @ D.T @ PJW.T
    assert np.linalg.norm(A - Aref) <= 1e-12 * np.linalg.norm(Aref)
    # rhs == -PJW D gl,  gl = vol H_ Sym (S - s) + Syminv g
    gl = np.stack([vol[e] * H_[e] @ (sym * (S[e] - s[e])) + g[e] / sym for e in range(nE)])
    assert np.linalg.norm(rhs + PJW @ D @ gl.reshape(-1)) <= 1e-12 * np.linalg.norm(rhs)
    # solve: la = gl + H_ D^T PJW^T dx ; ds = -Hinv (g - Sym la)
    dx = np.random.default_rng(0).standard_normal(n) * 1e-3
    o.solve(dx)
    la = gl.reshape(-1) + Hb @ (D.T @ (PJW.T @ dx))
    assert np.linalg.norm(o.get("la") - la) <= 1e-12 * np.linalg.norm(la)
    Hinv = o.get("Hinv").reshape(nE, N, N)
    ds = np.stack([-Hinv[e] @ (g[e] - sym * la.reshape(nE, N)[e]) for e in range(nE)]).reshape(-1)
    assert np.linalg.norm(o.get("ds") - ds) <= 1e-12 * np.linalg.norm(ds)
    # grad_x = PJW D la_old ; grad = vol g ; grad_la = Sym (S - s) vol
    assert np.linalg.norm(o.get("grad_x") - PJW @ D @ st["la"]) <= 1e-12 * np.linalg.norm(o.get("grad_x"))
    assert np.allclose(o.get("grad").reshape(nE, N), vol[:, None] * g, rtol=1e-14)
    assert np.allclose(o.get("grad_la").reshape(nE, N), sym * (S - s) * vol[:, None], rtol=1e-12, atol=1e-18)


def test_deformation_gradient_and_stretch_against_numpy():
    V, T, fm, st, o = make()
    o.update_rotations(st["x"])
    F = meshgen.deformation_gradients(V, T, st["x"])
    R, S = meshgen.exact_polar(F)
    Ro = o.get("R").reshape(-1, 3, 3).transpose(0, 2, 1)
    assert np.abs(Ro - R).max() < 1e-6            # tol of newton_procrustes
    assert np.abs(o.get("S").reshape(-1, 6) - meshgen.sym_to_vec(S)).max() < 1e-6
    assert o.iters().max() <= 4 and o.status() == 0


def test_energy_and_constraint_value():
    V, T, fm, st, o = make()
    o.update(st["x"], 0.0333)
    vol = o.get("vols")
    s = st["s"].reshape(-1, 6)
    e = sum(orc.material("Stable-Neohookean", MU, LA, s[i])[0] * vol[i] for i in range(T.shape[0]))
    assert abs(o.energy(st["x"], st["s"]) - e) <= 1e-12 * abs(e)
    S = o.get("S").reshape(-1, 6)
    sym = np.array([1, 1, 1, 2, 2, 2.0])
    c = float(np.sum(st["la"].reshape(-1, 6) * sym * (S - s) * vol[:, None]))
    assert abs(o.constraint_value(st["x"], st["s"]) - c) <= 1e-10 * abs(c)
    assert np.allclose(o.evaluate_constraint(st["x"]), o.get("grad_la"), rtol=0, atol=0)


def test_2d_volume_is_signed_double_area():
    V, T = meshgen.tri_grid(2, 2)
    fm = meshgen.pinned_free_map(V)
    o = orc.OracleMixedStretch(V, T, fm, [("ARAP", MU, LA)])
    assert np.allclose(o.get("vols"), 0.25)        # two triangles per 0.5 x 0.5 square: area 0.125, doubled
    o2 = orc.OracleMixedStretch(V, T[:, [0, 2, 1]], fm, [("ARAP", MU, LA)])
    assert np.allclose(o2.get("vols"), -0.25)      # clockwise -> negative (tri2d_mesh.cpp:32-34 takes it as is)
