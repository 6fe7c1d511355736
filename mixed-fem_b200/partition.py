"""Element partitioning of a mesh across ranks (one process per GPU) and the interface-vertex exchange.

The reference is a single-address-space OpenMP code; this is the new multi-GPU layer of §8(e):
elements are split into contiguous slabs (sorted by centroid along the longest axis), every rank builds a LOCAL
MixedStretch on its elements with locally renumbered vertices, and the only coupling is at vertices shared between
slabs: the per-vertex vectors (grad_x_, rhs_, and matvec results of the sub-assembled matrix) are summed over ranks
over ranks (NCCL over NVLink on GPUs: point-to-point with the neighbouring parts; gloo all-reduce in the CPU tests).
The assembled matrix is never exchanged: A = sum_r P_r^T A_r P_r is applied as local SpMV + interface sum.

The partitioner itself lives in the library (mfb_partition_create, csrc/mfb_partition.cu: host code behind the C ABI);
`make_local` wraps it.  The numpy functions below restate the slab rule and are what the CPU tests check it against.
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


def make_local_numpy(V, T, free_map, nparts, rank):
    """numpy restatement of the slab partition (kept as the cross-check of the library's partitioner in the tests)"""
    parts = partition_elements(V, T, nparts)
    shared, owner = shared_free_vertices(T, free_map, parts, with_owner=True)
    return LocalProblem(V, T, free_map, parts[rank], shared, owner, rank), parts, shared


class CLocalProblem:
    """This rank's part from the library's partitioner (mfb_partition_create, host code behind the C ABI): same
    attributes as LocalProblem plus the per-neighbour lists and the interface-first element order."""

    def __init__(self, V, T, free_map, nparts, rank, method=0):
        import ctypes as C
        from . import lib
        L = lib()
        V = np.ascontiguousarray(V, dtype=np.float64)
        T = np.ascontiguousarray(T, dtype=np.int32)
        fm = np.ascontiguousarray(free_map, dtype=np.int32)
        L.mfb_partition_create.argtypes = [C.POINTER(C.c_void_p), C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_int32),
                                           C.POINTER(C.c_double), C.POINTER(C.c_int32), C.c_int32, C.c_int32, C.c_int32]
        L.mfb_partition_destroy.argtypes = [C.c_void_p]
        L.mfb_partition_destroy.restype = None
        L.mfb_partition_sizes.argtypes = [C.c_void_p, C.POINTER(C.c_int32)]
        L.mfb_partition_get.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_void_p), C.POINTER(C.c_int64)]
        self._L = L
        self.h = C.c_void_p()
        rc = L.mfb_partition_create(C.byref(self.h), V.shape[1], T.shape[1], V.shape[0], T.shape[0], T.ctypes.data_as(C.POINTER(C.c_int32)),
                                    V.ctypes.data_as(C.POINTER(C.c_double)), fm.ctypes.data_as(C.POINTER(C.c_int32)), nparts, rank, method)
        if rc != 0:
            raise RuntimeError("mfb_partition_create failed: %d %s" % (rc, L.mfb_last_error_string().decode()))
        sizes = (C.c_int32 * 8)()
        L.mfb_partition_sizes(self.h, sizes)
        self.n_if_elems, self.n_shared = int(sizes[3]), int(sizes[4])
        d, npe = V.shape[1], T.shape[1]

        def get(what, dtype):
            ptr, cnt = C.c_void_p(), C.c_int64()
            L.mfb_partition_get(self.h, what, C.byref(ptr), C.byref(cnt))
            if cnt.value == 0:
                return np.zeros(0, dtype=dtype)
            buf = (C.c_char * (cnt.value * np.dtype(dtype).itemsize)).from_address(ptr.value)
            return np.frombuffer(buf, dtype=dtype).copy()
        self.elems = get(0, np.int32).astype(np.int64)
        self.vmap = get(1, np.int32).astype(np.int64)
        self.T = get(2, np.int32).reshape(-1, npe)
        self.V = get(3, np.float64).reshape(-1, d)
        self.free_map = get(4, np.int32)
        self.free_l2g = get(5, np.int64)
        self.if_local = get(6, np.int32).astype(np.int64)
        self.if_pos = get(7, np.int32).astype(np.int64)
        self.dot_weight = get(8, np.float64)
        self.nbr_rank = get(9, np.int32)
        self.nbr_count = get(10, np.int32)
        self.nbr_local = get(11, np.int32)
        self.elem_part = get(12, np.int32)
        self.rank, self.nparts = rank, nparts

    restrict_vertex_vector = LocalProblem.restrict_vertex_vector
    restrict_element_vector = LocalProblem.restrict_element_vector

    def close(self):
        if getattr(self, "h", None):
            self._L.mfb_partition_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def make_local(V, T, free_map, nparts, rank, method=0):
    """(this rank's part, list of element-index arrays of all parts, ascending global free ids of the shared vertices)
    from the library's partitioner; method 0 = slabs along the longest axis, 1 = recursive coordinate bisection"""
    loc = CLocalProblem(V, T, free_map, nparts, rank, method)
    parts = [np.nonzero(loc.elem_part == q)[0] for q in range(nparts)]
    fm = np.asarray(free_map)
    shared = shared_free_vertices(np.asarray(T), fm, parts) if nparts > 1 else np.zeros(0, dtype=np.int64)
    return loc, parts, shared


def interface_sum_numpy(local, vec, dim, allreduce):
    """vec (dim*nFl) <- sum over ranks at interface vertices.  `allreduce(buf)` sums a numpy buffer in place."""
    buf = np.zeros((local.n_shared, dim))
    v = vec.reshape(-1, dim)
    buf[local.if_pos] = v[local.if_local]
    allreduce(buf)
    v[local.if_local] = buf[local.if_pos]
    return vec
