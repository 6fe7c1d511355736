"""Element partitioning of a mesh across ranks (one process per GPU) and the interface-vertex exchange.

The reference is a single-address-space OpenMP code; this is the new multi-GPU layer of §8(e):
elements are split into contiguous slabs (sorted by centroid along the longest axis), every rank builds a LOCAL
MixedStretch on its elements with locally renumbered vertices, and the only coupling is at vertices shared between
slabs: the per-vertex vectors (grad_x_, rhs_, and matvec results of the sub-assembled matrix) are summed over ranks
with ONE all-reduce on a packed interface buffer (NCCL over NVLink on GPUs, gloo in the CPU tests).  The assembled
matrix is never exchanged: A = sum_r P_r^T A_r P_r is applied as local SpMV + interface sum.
"""
import numpy as np


def partition_elements(V, T, nparts):
    """list of element-index arrays (ascending), slabs of equal element count along the longest axis."""
    if nparts == 1:
        return [np.arange(T.shape[0])]
    ext = V.max(axis=0) - V.min(axis=0)
    axis = int(np.argmax(ext))
    cen = V[T][:, :, axis].mean(axis=1)
    order = np.argsort(cen, kind="stable")
    chunks = np.array_split(order, nparts)
    return [np.sort(c) for c in chunks]


class LocalProblem:
    """The piece of the global problem one rank owns."""

    def __init__(self, V, T, free_map, elems, shared_free_ids, shared_owner=None, rank=0):
        elems = np.asarray(elems)
        Tg = T[elems]
        self.vmap = np.unique(Tg)                         # local vertex -> global vertex (ascending)
        g2l = np.full(V.shape[0], -1, dtype=np.int64)
        g2l[self.vmap] = np.arange(self.vmap.size)
        self.T = g2l[Tg].astype(np.int32)
        self.V = np.ascontiguousarray(V[self.vmap])
        gfree = np.asarray(free_map)[self.vmap]           # global free id per local vertex
        fixed = gfree < 0
        self.free_map = np.full(self.vmap.size, -1, dtype=np.int32)
        self.free_map[~fixed] = np.arange(int((~fixed).sum()), dtype=np.int32)
        self.free_l2g = gfree[~fixed].astype(np.int64)    # local free id -> global free id (ascending)
        self.elems = elems
        # interface: positions (in the packed buffer of ALL shared free vertices) of this rank's shared vertices
        pos = np.searchsorted(shared_free_ids, self.free_l2g)
        pos[pos >= shared_free_ids.size] = 0
        hit = shared_free_ids[pos] == self.free_l2g if shared_free_ids.size else np.zeros(self.free_l2g.size, bool)
        self.if_local = np.nonzero(hit)[0].astype(np.int64)     # local free ids on the interface
        self.if_pos = pos[hit].astype(np.int64)                 # their slots in the packed buffer
        self.n_shared = int(shared_free_ids.size)
        # global dot products count every vertex once: weight 1 for interior vertices and for interface vertices this
        # rank owns (owner = lowest rank touching the vertex), 0 for copies owned elsewhere
        self.dot_weight = np.ones(self.free_l2g.size)
        if shared_owner is not None and self.if_local.size:
            self.dot_weight[self.if_local] = (np.asarray(shared_owner)[self.if_pos] == rank).astype(np.float64)

    def restrict_vertex_vector(self, xg, dim):
        """global vertex vector (dim*nV) -> local (dim*nVl)"""
        return np.ascontiguousarray(xg.reshape(-1, dim)[self.vmap]).reshape(-1)

    def restrict_element_vector(self, vg, n):
        return np.ascontiguousarray(vg.reshape(-1, n)[self.elems]).reshape(-1)


def shared_free_vertices(T, free_map, parts, with_owner=False):
    """ascending global free ids of the vertices touched by more than one part (and, optionally, the lowest rank
    touching each: its owner)"""
    nV = len(free_map)
    count = np.zeros(nV, dtype=np.int32)
    owner = np.full(nV, -1, dtype=np.int32)
    for r, p in enumerate(parts):
        u = np.unique(T[p])
        count[u] += 1
        first = u[owner[u] < 0]
        owner[first] = r
    fm = np.asarray(free_map)
    sh = np.nonzero((count > 1) & (fm >= 0))[0]
    order = np.argsort(fm[sh], kind="stable")
    ids = fm[sh][order].astype(np.int64)
    if with_owner:
        return ids, owner[sh][order]
    return ids


def make_local(V, T, free_map, nparts, rank):
    parts = partition_elements(V, T, nparts)
    shared, owner = shared_free_vertices(T, free_map, parts, with_owner=True)
    return LocalProblem(V, T, free_map, parts[rank], shared, owner, rank), parts, shared


def interface_sum_numpy(local, vec, dim, allreduce):
    """vec (dim*nFl) <- sum over ranks at interface vertices.  `allreduce(buf)` sums a numpy buffer in place."""
    buf = np.zeros((local.n_shared, dim))
    v = vec.reshape(-1, dim)
    buf[local.if_pos] = v[local.if_local]
    allreduce(buf)
    v[local.if_local] = buf[local.if_pos]
    return vec


class TorchInterface:
    """Device-side pack -> all_reduce -> unpack of interface entries for vectors living in HBM (torch is plumbing:
    it owns the packed buffer and calls NCCL)."""

    def __init__(self, local, dim, device):
        import torch
        self.torch = torch
        self.dim = dim
        self.idx_local = torch.as_tensor(local.if_local, device=device)
        self.idx_pos = torch.as_tensor(local.if_pos, device=device)
        self.buf = torch.zeros((local.n_shared, dim * 2), dtype=torch.float64, device=device)

    def sum2(self, a, b):
        """a, b: (nFl, dim) float64 tensors viewing grad_x_ and rhs_; one all-reduce for both."""
        import torch.distributed as dist
        d = self.dim
        self.buf.zero_()
        self.buf[self.idx_pos, :d] = a[self.idx_local]
        self.buf[self.idx_pos, d:] = b[self.idx_local]
        dist.all_reduce(self.buf)
        a[self.idx_local] = self.buf[self.idx_pos, :d]
        b[self.idx_local] = self.buf[self.idx_pos, d:]
