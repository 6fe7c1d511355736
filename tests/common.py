"""Shared helpers for the parity tests: build the same seeded problem on the oracle (CPU) and on the CUDA path."""
import numpy as np

import mixed_fem_b200 as mf
from mixed_fem_b200 import meshgen
from oracle import orc

MU, LA = meshgen.enu_to_lame(1e6, 0.45)
DT = 0.0333
TOL = 1e-10          # north_star: element gradients, Hessians and assembled matrices within 1e-10 relative


def mesh(dim=3, n=(6, 4, 3), shuffle=False, jitter=0.0, seed=0, pin=0.01):
    if dim == 3:
        V, T = meshgen.kuhn_grid(*n)
    else:
        V, T = meshgen.tri_grid(n[0], n[1])
    rng = np.random.default_rng(seed)
    h = 1.0 / n[0]
    if jitter:
        V = V + rng.uniform(-jitter, jitter, size=V.shape) * h
    if shuffle:   # irregular vertex numbering and element order, like an unstructured mesh
        perm = rng.permutation(V.shape[0])
        inv = np.empty_like(perm); inv[perm] = np.arange(perm.size)
        V = V[perm]
        T = inv[T].astype(np.int32)
        T = T[rng.permutation(T.shape[0])]
    fm = meshgen.pinned_free_map(V, ratio=pin) if pin is not None else np.arange(V.shape[0], dtype=np.int32)
    return V, T, fm, h


def pair(V, T, fm, materials, mat_id=None, h=None, state=True):
    """(gpu, cpu, st): both reset, then given the same seeded R / s / la."""
    gpu = mf.MixedStretch(V, T, fm, materials, mat_id=mat_id)
    cpu = orc.OracleMixedStretch(V, T, fm, materials, mat_id=mat_id)
    st = meshgen.synthetic_state(V, T, h if h else 1.0, materials[0][1])
    for o in (gpu, cpu):
        o.reset()
        if state:
            if V.shape[1] == 3:
                o.set("R", st["R"])
            o.set("s", st["s"]); o.set("la", st["la"])
    return gpu, cpu, st


def block_rel(a, b, width):
    """max over elements of |a_e - b_e|_F / |b_e|_F (norm-wise per element block)."""
    a = np.asarray(a).reshape(-1, width); b = np.asarray(b).reshape(-1, width)
    nb = np.linalg.norm(b, axis=1)
    na = np.linalg.norm(a - b, axis=1)
    scale = np.where(nb > 0, nb, 1.0)
    return float((na / scale).max())


def vec_rel(a, b):
    nb = np.linalg.norm(b)
    return float(np.linalg.norm(np.asarray(a) - np.asarray(b)) / (nb if nb > 0 else 1.0))


def compare_histories(a, b, rtol=1e-8, min_len=2):
    """Newton histories (north_star: within 1e-8).  Energies and decrements are compared up to and including the first
    iteration k whose relative energy change is below 1e-8, the line-search alpha before k: from there on f(alpha) < f(0)
    compares numbers that differ in the last bits, so the accepted alpha - and with it everything after - is a coin flip
    between two correct implementations."""
    n = min(a["n_iters"], b["n_iters"])
    res = np.asarray(b["energy_res"][:n])
    # ... or whose step is below 1e-7: the energy change such a step can make (~ decrement^2) is under the rounding noise
    # of the energy sum, so the line search there is the same coin flip (seen on scenes/2D/trampoline.json, iteration 2)
    settled = np.nonzero((res <= 1e-8) | (np.asarray(b["decrement"][:n]) <= 1e-7))[0]
    k = int(settled[0]) if settled.size else n - 1
    if not settled.size:
        assert a["n_iters"] == b["n_iters"], (a["n_iters"], b["n_iters"])
    ea, eb = np.asarray(a["energy"][:k + 1]), np.asarray(b["energy"][:k + 1])
    assert np.all(np.abs(ea - eb) <= rtol * np.abs(eb)), (ea, eb)
    da, db = np.asarray(a["decrement"][:k + 1]), np.asarray(b["decrement"][:k + 1])
    assert np.all(np.abs(da - db) <= rtol * np.abs(db) + 1e-12), (da, db)
    assert np.array_equal(np.asarray(a["alpha"][:k]), np.asarray(b["alpha"][:k]))
    assert k >= min_len, "history too short to mean anything"
