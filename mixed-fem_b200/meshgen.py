"""Deterministic synthetic inputs shared by the CUDA path, the tests and the benchmark.

Kuhn box grids (6 tets per cube along the (0,0,0)-(1,1,1) diagonal), the pinned-face boundary rule that
mirrors `init_boundary_groups` group 0 (src/boundary_conditions/boundary_condition.cpp:5-21), the reference's
free-DOF numbering (src/mesh/mesh.cpp:95-101) and the seeded deformed state.  numpy only: no device code here.
"""
import itertools
import numpy as np

SEED = 20221101


def enu_to_lame(E, nu):
    """src/config.h:13-16"""
    mu = E / (2.0 * (1.0 + nu))
    la = (E * nu) / ((1.0 + nu) * (1.0 - 2.0 * nu))
    return mu, la


def kuhn_grid(nx, ny, nz, h=None):
    """Vertices (nV,3) float64 and tets (nE,4) int32 of an nx*ny*nz box of cubes with edge h (default 1/nx).

    vertex id = i + (nx+1)*(j + (ny+1)*k); elements are cube-major (x fastest), then Kuhn tet 0..5; vertex order
    inside each tet is fixed so that igl::volume (deps/libigl/include/igl/volume.cpp:29) is positive.
    """
    if h is None:
        h = 1.0 / nx
    i, j, k = np.meshgrid(np.arange(nx + 1), np.arange(ny + 1), np.arange(nz + 1), indexing="ij")
    vid = (i + (nx + 1) * (j + (ny + 1) * k))
    V = np.empty(((nx + 1) * (ny + 1) * (nz + 1), 3))
    V[vid.ravel(), 0] = i.ravel() * h
    V[vid.ravel(), 1] = j.ravel() * h
    V[vid.ravel(), 2] = k.ravel() * h
    ci, cj, ck = np.meshgrid(np.arange(nx), np.arange(ny), np.arange(nz), indexing="ij")
    # cube-major, x fastest
    order = np.argsort((ci + nx * (cj + ny * ck)).ravel(), kind="stable")
    ci, cj, ck = ci.ravel()[order], cj.ravel()[order], ck.ravel()[order]
    base = ci + (nx + 1) * (cj + (ny + 1) * ck)
    step = np.array([1, nx + 1, (nx + 1) * (ny + 1)])
    T = np.empty((base.size, 6, 4), dtype=np.int64)
    for t, perm in enumerate(itertools.permutations(range(3))):
        v0 = base
        v1 = v0 + step[perm[0]]
        v2 = v1 + step[perm[1]]
        v3 = v2 + step[perm[2]]
        # parity of the permutation decides orientation; swap two vertices for odd ones
        inv = sum(1 for a in range(3) for b in range(a + 1, 3) if perm[a] > perm[b])
        if inv % 2 == 0:
            T[:, t, :] = np.stack([v0, v1, v2, v3], axis=1)
        else:
            T[:, t, :] = np.stack([v0, v2, v1, v3], axis=1)
    T = T.reshape(-1, 4).astype(np.int32)
    # make igl::volume positive for every tet (it is uniform per parity class; check one and flip if needed)
    a, b, c, d = (V[T[:6, q]] for q in range(4))
    vol = -np.einsum("ij,ij->i", a - d, np.cross(b - d, c - d)) / 6.0
    for t in range(6):
        if vol[t] < 0:
            T[t::6, 1], T[t::6, 2] = T[t::6, 2].copy(), T[t::6, 1].copy()
    return V, T


def tri_grid(nx, ny, h=None):
    """2D triangle grid (two triangles per square), for the Tri2DMesh path."""
    if h is None:
        h = 1.0 / nx
    i, j = np.meshgrid(np.arange(nx + 1), np.arange(ny + 1), indexing="ij")
    vid = i + (nx + 1) * j
    V = np.empty(((nx + 1) * (ny + 1), 2))
    V[vid.ravel(), 0] = i.ravel() * h
    V[vid.ravel(), 1] = j.ravel() * h
    ci, cj = np.meshgrid(np.arange(nx), np.arange(ny), indexing="ij")
    order = np.argsort((ci + nx * cj).ravel(), kind="stable")
    ci, cj = ci.ravel()[order], cj.ravel()[order]
    b = ci + (nx + 1) * cj
    t0 = np.stack([b, b + 1, b + 1 + (nx + 1)], axis=1)
    t1 = np.stack([b, b + 1 + (nx + 1), b + (nx + 1)], axis=1)
    T = np.stack([t0, t1], axis=1).reshape(-1, 3).astype(np.int32)
    return V, T


def shell_grid(nx, ny, h=None, bump=0.15):
    """Cloth-like sheet for the TriMesh / shell path (config C5): the tri_grid in the x-z plane of 3D space with a
    smooth out-of-plane bump so face normals vary.  Returns V (nV,3), T (nE,3), per-face unit rest normals N (nE,3)."""
    V2, T = tri_grid(nx, ny, h)
    L = max(V2[:, 0].max(), 1e-300)
    y = bump * L * np.sin(np.pi * V2[:, 0] / L) * np.cos(0.5 * np.pi * V2[:, 1] / L)
    V = np.stack([V2[:, 0], y, V2[:, 1]], axis=1)
    return V, T, face_normals(V, T)


def face_normals(V, T):
    e1 = V[T[:, 1]] - V[T[:, 0]]
    e2 = V[T[:, 2]] - V[T[:, 0]]
    n = np.cross(e1, e2)
    return n / np.linalg.norm(n, axis=1, keepdims=True)


def pinned_free_map(V, ratio=0.01, axis=0):
    """free_map (nV,) int32: -1 for vertices with coord < min + ratio*range along `axis`, else ascending free id."""
    lo, hi = V[:, axis].min(), V[:, axis].max()
    fixed = V[:, axis] < lo + ratio * (hi - lo)
    fm = np.full(V.shape[0], -1, dtype=np.int32)
    fm[~fixed] = np.arange(int((~fixed).sum()), dtype=np.int32)
    return fm


def _affine():
    ax = np.array([1.0, 1.0, 1.0]) / np.sqrt(3.0)
    th = 0.3
    K = np.array([[0, -ax[2], ax[1]], [ax[2], 0, -ax[0]], [-ax[1], ax[0], 0]])
    Q = np.eye(3) + np.sin(th) * K + (1 - np.cos(th)) * (K @ K)
    return Q @ np.diag([1.2, 0.9, 1.05])


def deformed_positions(V, h, seed=SEED, scale=1.0):
    """x (nV*d,) interleaved: affine map (rotation 0.3 rad about (1,1,1) composed with stretch diag(1.2,.9,1.05))
    plus iid U(-0.02,0.02)*h jitter scaled by `scale`.  The same rng stream gives the same jitter for any scale."""
    rng = np.random.default_rng(seed)
    d = V.shape[1]
    A = _affine()[:d, :d]
    delta = rng.uniform(-0.02, 0.02, size=V.shape) * h
    return (V @ A.T + scale * delta).reshape(-1)


def tet_dphi(V, T):
    """(nE,4,3) shape-function gradients: rows 1..3 = Dm^-1, row 0 = -column sums (linear_tet_dphi_dX.cpp:5-17)."""
    X = V[T]
    Dm = np.transpose(X[:, 1:, :] - X[:, :1, :], (0, 2, 1))
    inv = np.linalg.inv(Dm)
    return np.concatenate([-inv.sum(axis=1, keepdims=True), inv], axis=1)


def deformation_gradients(V, T, x):
    """(nE,d,d) F = sum_k x_k (x) dphi_k."""
    d = V.shape[1]
    if d == 3:
        dphi = tet_dphi(V, T)
    else:
        X = V[T]
        Dm = np.transpose(X[:, 1:, :] - X[:, :1, :], (0, 2, 1))
        inv = np.linalg.inv(Dm)
        dphi = np.concatenate([-inv.sum(axis=1, keepdims=True), inv], axis=1)
    xe = x.reshape(-1, d)[T]
    return np.einsum("ekr,ekc->erc", xe, dphi)


def exact_polar(F):
    """Batched polar factors via SVD: returns R (nE,d,d), S (nE,d,d)."""
    U, sig, Vt = np.linalg.svd(F)
    det = np.linalg.det(U @ Vt)
    U[:, :, -1] *= np.sign(det)[:, None]
    R = U @ Vt
    S = np.transpose(R, (0, 2, 1)) @ F
    S = 0.5 * (S + np.transpose(S, (0, 2, 1)))
    return R, S


def sym_to_vec(S):
    d = S.shape[-1]
    if d == 3:
        return np.stack([S[:, 0, 0], S[:, 1, 1], S[:, 2, 2], S[:, 0, 1], S[:, 0, 2], S[:, 1, 2]], axis=1)
    return np.stack([S[:, 0, 0], S[:, 1, 1], S[:, 0, 1]], axis=1)


def synthetic_state(V, T, h, mu, seed=SEED):
    """The §8(d) state: x (current), and R / s / la from the 'previous' state (half the jitter).

    Returns dict with x (nV*d), R (nE*d*d, column-major per element), s (nE*N), la (nE*N)."""
    rng = np.random.default_rng(seed)
    d = V.shape[1]
    A = _affine()[:d, :d]
    delta = rng.uniform(-0.02, 0.02, size=V.shape) * h
    x = (V @ A.T + delta).reshape(-1)
    xprev = (V @ A.T + 0.5 * delta).reshape(-1)
    R, S = exact_polar(deformation_gradients(V, T, xprev))
    s = sym_to_vec(S)
    la = rng.uniform(-1.0, 1.0, size=s.shape) * mu * 1e-3
    Rcm = np.transpose(R, (0, 2, 1)).reshape(T.shape[0], -1)  # column-major per element
    return {"x": x, "R": np.ascontiguousarray(Rcm).reshape(-1), "s": s.reshape(-1), "la": la.reshape(-1), "xprev": xprev}


CONFIGS = {
    # name: (nx, ny, nz)  — BASELINE.md §3
    "C2": (64, 16, 16),      # 98 304 tets
    "C3": (55, 55, 55),      # 998 250 tets
    "C4": (119, 119, 119),   # 10 110 954 tets
}
