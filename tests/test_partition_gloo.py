"""N>1 path on CPU: element partition + interface all-reduce over gloo (world_size 2 and 3), with the oracle standing in
for each rank's local update.  Checks that the slab decomposition + interface sum reproduces the single-domain grad_x_,
rhs_ and matvec, i.e. A = sum_r P_r^T A_r P_r."""
import os
import socket
import sys

import numpy as np
import pytest
import scipy.sparse as sp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, q):
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    import mixed_fem_b200  # noqa: F401
    from mixed_fem_b200 import meshgen, partition
    from oracle import orc
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    orc.lib().orc_set_num_threads(1)
    V, T = meshgen.kuhn_grid(6, 3, 3)
    fm = meshgen.pinned_free_map(V, ratio=0.01)
    mu, la = meshgen.enu_to_lame(1e6, 0.45)
    st = meshgen.synthetic_state(V, T, 1.0 / 6, mu)
    loc, parts, shared = partition.make_local(V, T, fm, world, rank)
    o = orc.OracleMixedStretch(loc.V, loc.T, loc.free_map, [("Stable-Neohookean", mu, la)])
    o.reset()
    R = st["R"].reshape(-1, 9)[loc.elems].reshape(-1)
    o.set("R", R); o.set("s", loc.restrict_element_vector(st["s"], 6)); o.set("la", loc.restrict_element_vector(st["la"], 6))
    o.update(loc.restrict_vertex_vector(st["x"], 3), 0.0333)

    def allreduce(buf):
        t = torch.from_numpy(buf)
        dist.all_reduce(t)

    gx = partition.interface_sum_numpy(loc, o.get("grad_x").copy(), 3, allreduce)
    rhs = partition.interface_sum_numpy(loc, o.rhs().copy(), 3, allreduce)
    # distributed matvec y = A p with the sub-assembled local matrices
    nFg = int(fm.max()) + 1
    p_glob = np.random.default_rng(1).standard_normal(3 * nFg)
    p_loc = p_glob.reshape(-1, 3)[loc.free_l2g].reshape(-1)
    outer, inner, vals = o.csr()
    A = sp.csr_matrix((vals, inner, outer), shape=(outer.size - 1, outer.size - 1))
    y = partition.interface_sum_numpy(loc, A @ p_loc, 3, allreduce)
    q.put((rank, loc.free_l2g, gx, rhs, y, loc.elems.size, loc.if_local.size))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_slab_partition_interface_sum_matches_single_domain(world):
    import torch.multiprocessing as mp
    from mixed_fem_b200 import meshgen
    from oracle import orc
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    V, T = meshgen.kuhn_grid(6, 3, 3)
    fm = meshgen.pinned_free_map(V, ratio=0.01)
    mu, la = meshgen.enu_to_lame(1e6, 0.45)
    st = meshgen.synthetic_state(V, T, 1.0 / 6, mu)
    o = orc.OracleMixedStretch(V, T, fm, [("Stable-Neohookean", mu, la)])
    o.reset()
    o.set("R", st["R"]); o.set("s", st["s"]); o.set("la", st["la"])
    o.update(st["x"], 0.0333)
    gx_g, rhs_g = o.get("grad_x").reshape(-1, 3), o.rhs().reshape(-1, 3)
    outer, inner, vals = o.csr()
    A = sp.csr_matrix((vals, inner, outer), shape=(outer.size - 1, outer.size - 1))
    y_g = (A @ np.random.default_rng(1).standard_normal(A.shape[0])).reshape(-1, 3)
    assert sum(r[5] for r in res) == T.shape[0]
    assert all(r[6] > 0 for r in res)
    covered = np.zeros(gx_g.shape[0], dtype=bool)
    for rank, l2g, gx, rhs, y, _, _ in res:
        covered[l2g] = True
        for a, b in ((gx, gx_g), (rhs, rhs_g), (y, y_g)):
            a = a.reshape(-1, 3)
            # interior vertices of a slab hold the full value; interface vertices hold it after the sum
            full = np.ones(l2g.size, dtype=bool)
            assert np.linalg.norm(a[full] - b[l2g][full]) <= 1e-12 * np.linalg.norm(b)
    assert covered.all()


def test_partition_is_a_partition():
    from mixed_fem_b200 import meshgen, partition
    V, T = meshgen.kuhn_grid(8, 3, 2)
    fm = meshgen.pinned_free_map(V)
    for n in (1, 2, 4, 8):
        parts = partition.partition_elements(V, T, n)
        allel = np.concatenate(parts)
        assert np.array_equal(np.sort(allel), np.arange(T.shape[0]))
        assert max(len(p) for p in parts) - min(len(p) for p in parts) <= 1
        shared = partition.shared_free_vertices(T, fm, parts)
        assert (n == 1) == (shared.size == 0)
        for r in range(n):
            loc = partition.LocalProblem(V, T, fm, parts[r], shared)
            assert np.array_equal(loc.vmap[loc.T], T[parts[r]])
            free = loc.free_map[loc.free_map >= 0]
            assert np.array_equal(free, np.arange(free.size))
            assert np.all(np.diff(loc.free_l2g) > 0)
            assert np.array_equal(shared[loc.if_pos], loc.free_l2g[loc.if_local])


def _cg_worker(rank, world, port, q):
    """the distributed CG of mixed-fem_b200/csrc/mfb_pcg.cu restated in numpy on top of the SAME host-side interface
    description the library consumes (LocalProblem.if_local / if_pos / dot_weight): sub-assembled operator, interface
    sum of the product and of the diagonal blocks, owner-weighted dot products summed over ranks."""
    import torch
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    import mixed_fem_b200  # noqa: F401
    from mixed_fem_b200 import meshgen, partition
    from oracle import orc
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(1)
    orc.lib().orc_set_num_threads(1)
    V, T = meshgen.kuhn_grid(6, 3, 3)
    fm = meshgen.pinned_free_map(V, ratio=0.01)
    mu, la = meshgen.enu_to_lame(1e6, 0.45)
    st = meshgen.synthetic_state(V, T, 1.0 / 6, mu)
    loc, parts, shared = partition.make_local(V, T, fm, world, rank)
    o = orc.OracleMixedStretch(loc.V, loc.T, loc.free_map, [("Stable-Neohookean", mu, la)])
    o.reset()
    o.set("R", st["R"].reshape(-1, 9)[loc.elems].reshape(-1))
    o.set("s", loc.restrict_element_vector(st["s"], 6)); o.set("la", loc.restrict_element_vector(st["la"], 6))
    o.update(loc.restrict_vertex_vector(st["x"], 3), 0.0333)
    outer, inner, vals = o.csr()
    n = outer.size - 1
    A = sp.csr_matrix((vals, inner, outer), shape=(n, n))

    def allreduce(buf):
        dist.all_reduce(torch.from_numpy(buf))

    def gsum(v):
        t = torch.tensor([v], dtype=torch.float64)
        dist.all_reduce(t)
        return float(t[0])

    w3 = np.repeat(loc.dot_weight, 3)
    wdot = lambda a, b: gsum(float((w3 * a) @ b))
    nFg = int(fm.max()) + 1
    b = np.random.default_rng(2).standard_normal(3 * nFg).reshape(-1, 3)[loc.free_l2g].reshape(-1)
    # Jacobi preconditioner from the summed diagonal
    d = partition.interface_sum_numpy(loc, A.diagonal().copy(), 3, allreduce)
    x = np.zeros(n)
    r = b.copy()
    thr = (1e-10) ** 2 * wdot(b, b)
    p = r / d
    absnew = wdot(r, p)
    it = 0
    while it < 2000:
        t = partition.interface_sum_numpy(loc, A @ p, 3, allreduce)
        # p.(A p) from the UNSUMMED local products would be sum_r p_r.(A_r p_r); with the summed t it needs the weights
        alpha = absnew / wdot(p, t)
        x += alpha * p
        r -= alpha * t
        z = r / d
        rr, rz = wdot(r, r), wdot(r, z)
        if rr < thr:
            break
        p = z + (rz / absnew) * p
        absnew = rz
        it += 1
    q.put((rank, loc.free_l2g, x, it, float(loc.dot_weight.sum())))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_distributed_cg_with_owner_weights_matches_global_solve(world):
    import scipy.sparse.linalg as spla
    import torch.multiprocessing as mp
    from mixed_fem_b200 import meshgen
    from oracle import orc
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_cg_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=300) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    V, T = meshgen.kuhn_grid(6, 3, 3)
    fm = meshgen.pinned_free_map(V, ratio=0.01)
    mu, la = meshgen.enu_to_lame(1e6, 0.45)
    st = meshgen.synthetic_state(V, T, 1.0 / 6, mu)
    o = orc.OracleMixedStretch(V, T, fm, [("Stable-Neohookean", mu, la)])
    o.reset()
    o.set("R", st["R"]); o.set("s", st["s"]); o.set("la", st["la"])
    o.update(st["x"], 0.0333)
    outer, inner, vals = o.csr()
    A = sp.csr_matrix((vals, inner, outer), shape=(outer.size - 1, outer.size - 1))
    b = np.random.default_rng(2).standard_normal(A.shape[0])
    xg = spla.spsolve(A.tocsc(), b).reshape(-1, 3)
    # every global free vertex is counted exactly once by the dot-product weights
    assert abs(sum(r[4] for r in res) - xg.shape[0]) < 1e-9
    its = {r[3] for r in res}
    assert len(its) == 1 and 0 < its.pop() < 2000          # same control flow on every rank
    for rank, l2g, x, it, _ in res:
        assert np.linalg.norm(x.reshape(-1, 3) - xg[l2g]) <= 1e-7 * np.linalg.norm(xg)


def test_partitioner_rejects_bad_arguments_and_handles_the_smallest_cases(built):
    """mfb_partition_create (host code behind the C ABI): argument errors come back as messages; as many parts as elements is
    the limit (every part one element), one more is an error rather than an empty rank"""
    from mixed_fem_b200 import meshgen, partition
    V, T = meshgen.kuhn_grid(1, 1, 1)            # 6 tets
    fm = meshgen.pinned_free_map(V, ratio=0.3)
    for rank in range(6):
        loc, _, _ = partition.make_local(V, T, fm, 6, rank)
        assert np.asarray(loc.T).shape == (1, 4)
    cases = {
        "more parts": dict(nparts=7, rank=0),
        "bad arguments": dict(nparts=2, rank=2),
        "bad arguments ": dict(nparts=65, rank=0),
    }
    for key, kw in cases.items():
        with pytest.raises(RuntimeError) as ei:
            partition.make_local(V, T, fm, kw["nparts"], kw["rank"])
        assert key.strip() in str(ei.value), str(ei.value)
    bad = T.copy(); bad[3, 1] = V.shape[0]
    with pytest.raises(RuntimeError) as ei:
        partition.make_local(V, bad, fm, 2, 0)
    assert "vertex id out of range" in str(ei.value)
    free = np.flatnonzero(fm >= 0)
    assert free.size >= 2
    f2 = fm.copy(); f2[free[1]] = f2[free[0]]
    with pytest.raises(RuntimeError) as ei:
        partition.make_local(V, T, f2, 2, 0)
    assert "appears twice" in str(ei.value) or "skips ids" in str(ei.value)
    f2 = fm.copy(); f2[0] = -7
    with pytest.raises(RuntimeError) as ei:
        partition.make_local(V, T, f2, 2, 0)
    assert "< -1" in str(ei.value)
