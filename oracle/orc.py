"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes loader for oracle/liborc.so, the CPU restatement of the reference's mixed-stretch path
(oracle/mfem_oracle.cpp).  Import this only from tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs; the product package never imports it.
"""
import ctypes as C
import os
import subprocess
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
MODELS = {"Stable-Neohookean": 0, "Neohookean": 1, "Corot": 2, "FixedCorot": 3, "ARAP": 4, "Fung": 5}


class _Mat(C.Structure):
    _fields_ = [("model", C.c_int32), ("mu", C.c_double), ("la", C.c_double)]


def build(march=None, out=None, force=False):
    """Compile liborc.so (or `out`) with the recipe in oracle/Makefile."""
    target = out or os.path.join(_HERE, "liborc.so")
    srcs = [os.path.join(_HERE, f) for f in ("mfem_oracle.cpp", "orc_linalg.hpp", "orc_materials.hpp", "orc_polar.hpp")]
    if not force and os.path.exists(target) and all(os.path.getmtime(target) >= os.path.getmtime(s) for s in srcs):
        return target
    cmd = ["make", "-C", _HERE, "-B", "OUT=" + target]
    if march:
        cmd.append("ORC_MARCH=" + march)
    subprocess.run(cmd, check=True, capture_output=True)
    return target


_lib = None


def lib(path=None):
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or build()
    L = C.CDLL(p)
    dp = C.POINTER(C.c_double)
    ip = C.POINTER(C.c_int32)
    L.orc_create.restype = C.c_void_p
    L.orc_create.argtypes = [C.c_int, C.c_int, C.c_int, ip, dp, ip, ip, C.c_int, C.POINTER(_Mat)]
    L.orc_create_shell.restype = C.c_void_p
    L.orc_create_shell.argtypes = [C.c_int, C.c_int, ip, dp, ip, dp, C.c_double, ip, C.c_int, C.POINTER(_Mat)]
    L.orc_destroy.argtypes = [C.c_void_p]
    L.orc_reset.argtypes = [C.c_void_p]
    L.orc_update.argtypes = [C.c_void_p, dp, C.c_double]
    L.orc_update_rotations.argtypes = [C.c_void_p, dp]
    L.orc_rhs.argtypes = [C.c_void_p, dp]
    L.orc_solve.argtypes = [C.c_void_p, dp]
    L.orc_energy.restype = C.c_double
    L.orc_energy.argtypes = [C.c_void_p, dp, dp]
    L.orc_constraint_value.restype = C.c_double
    L.orc_constraint_value.argtypes = [C.c_void_p, dp, dp]
    L.orc_evaluate_constraint.argtypes = [C.c_void_p, dp, dp]
    L.orc_product.argtypes = [C.c_void_p, C.c_int, dp, dp, C.c_int]
    L.orc_nfree.argtypes = [C.c_void_p]
    L.orc_status.argtypes = [C.c_void_p]
    L.orc_nnz.restype = C.c_int64
    L.orc_nnz.argtypes = [C.c_void_p]
    L.orc_csr.argtypes = [C.c_void_p, ip, ip, dp]
    L.orc_scatter_map.argtypes = [C.c_void_p, ip, ip]
    L.orc_get.restype = C.c_int64
    L.orc_get.argtypes = [C.c_void_p, C.c_char_p, dp]
    L.orc_set.argtypes = [C.c_void_p, C.c_char_p, dp]
    L.orc_get_iters.argtypes = [C.c_void_p, ip]
    L.orc_timers.argtypes = [C.c_void_p, dp, C.c_int]
    L.orc_set_num_threads.argtypes = [C.c_int]
    L.orc_newton_procrustes.argtypes = [dp, dp, C.c_int, dp, C.c_double, C.c_int, ip]
    L.orc_polar_svd3.argtypes = [dp, dp, dp, C.c_int, dp, ip]
    L.orc_polar_svd2.argtypes = [dp, dp, dp, dp]
    L.orc_rodrigues.argtypes = [dp, dp]
    L.orc_colpiv_solve3.argtypes = [dp, dp, dp]
    L.orc_sym_eig.argtypes = [C.c_int, dp, dp, dp]
    L.orc_psd_fix_invert.argtypes = [C.c_int, dp, dp, dp, C.c_double]
    L.orc_svd2.argtypes = [dp, dp, dp, dp, C.c_int]
    L.orc_material_energy.restype = C.c_double
    L.orc_material_energy.argtypes = [C.c_int, C.c_double, C.c_double, C.c_int, dp]
    L.orc_material_gradient.argtypes = [C.c_int, C.c_double, C.c_double, C.c_int, dp, dp]
    L.orc_material_hessian.argtypes = [C.c_int, C.c_double, C.c_double, C.c_int, dp, dp]
    if path is None:
        _lib = L
    return L


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


TIMER_KEYS = ("rotations", "derivatives", "Aloc", "assemble", "rhs", "solve", "energy", "constraint")


class OracleMixedStretch:
    """CPU MixedStretch<DIM> (src/variables/mixed_stretch.h:12-164) over plain numpy arrays."""

    def __init__(self, V, T, free_map, materials, mat_id=None, libpath=None, normals=None, thickness=1e-3):
        """normals given (nE x 3) => shell: TriMesh + MixedStretch<3> (src/mesh/tri_mesh.cpp), T is nE x 3, V is nV x 3."""
        self.L = lib(libpath)
        V = _f64(V)
        T = np.ascontiguousarray(T, dtype=np.int32)
        self.shell = normals is not None
        self.dim = V.shape[1]
        self.nV, self.nE = V.shape[0], T.shape[0]
        self.N = 6 if self.dim == 3 else 3
        self.M = self.dim * self.dim
        fm = np.ascontiguousarray(free_map, dtype=np.int32)
        mats = (_Mat * len(materials))()
        for i, (model, mu, la) in enumerate(materials):
            mats[i].model = MODELS[model] if isinstance(model, str) else int(model)
            mats[i].mu, mats[i].la = float(mu), float(la)
        mid = None if mat_id is None else np.ascontiguousarray(mat_id, dtype=np.int32)
        self._keep = (V, T, fm, mid)
        if self.shell:
            nrm = _f64(normals)
            self._keep += (nrm,)
            self.h = self.L.orc_create_shell(self.nV, self.nE, _ip(T), _dp(V), _ip(fm), _dp(nrm), float(thickness),
                                             _ip(mid) if mid is not None else None, len(materials), mats)
        else:
            self.h = self.L.orc_create(self.dim, self.nV, self.nE, _ip(T), _dp(V), _ip(fm),
                                       _ip(mid) if mid is not None else None, len(materials), mats)
        self.nfree = self.L.orc_nfree(self.h)

    def __del__(self):
        try:
            self.L.orc_destroy(self.h)
        except Exception:
            pass

    def reset(self):
        self.L.orc_reset(self.h)

    def update(self, x, dt):
        x = _f64(x)
        self.L.orc_update(self.h, _dp(x), float(dt))

    def update_rotations(self, x):
        x = _f64(x)
        self.L.orc_update_rotations(self.h, _dp(x))

    def rhs(self):
        out = np.empty(self.dim * self.nfree)
        self.L.orc_rhs(self.h, _dp(out))
        return out

    def solve(self, dx):
        dx = _f64(dx)
        self.L.orc_solve(self.h, _dp(dx))

    def energy(self, x, s):
        x, s = _f64(x), _f64(s)
        return self.L.orc_energy(self.h, _dp(x), _dp(s))

    def constraint_value(self, x, s):
        x, s = _f64(x), _f64(s)
        return self.L.orc_constraint_value(self.h, _dp(x), _dp(s))

    def evaluate_constraint(self, x):
        x = _f64(x)
        out = np.empty(self.N * self.nE)
        self.L.orc_evaluate_constraint(self.h, _dp(x), _dp(out))
        return out

    def product(self, which, x, out, transposed=False):
        idx = {"hessian": 0, "hessian_inv": 1, "jacobian_mixed": 2, "jacobian_x": 3}[which]
        x = _f64(x)
        out = _f64(out).copy()
        self.L.orc_product(self.h, idx, _dp(x), _dp(out), int(transposed))
        return out

    def get(self, name):
        n = self.L.orc_get(self.h, name.encode(), None)
        if n < 0:
            raise KeyError(name)
        out = np.empty(n)
        self.L.orc_get(self.h, name.encode(), _dp(out))
        return out

    def set(self, name, val):
        val = _f64(val)
        if self.L.orc_set(self.h, name.encode(), _dp(val)) != 0:
            raise KeyError(name)

    def iters(self):
        out = np.empty(self.nE, dtype=np.int32)
        self.L.orc_get_iters(self.h, _ip(out))
        return out

    def status(self):
        return self.L.orc_status(self.h)

    def csr(self, values=True):
        nnz = self.L.orc_nnz(self.h)
        outer = np.empty(self.dim * self.nfree + 1, dtype=np.int32)
        inner = np.empty(nnz, dtype=np.int32)
        vals = np.empty(nnz) if values else None
        self.L.orc_csr(self.h, _ip(outer), _ip(inner), _dp(vals) if values else None)
        return outer, inner, vals

    def scatter_map(self):
        npe = 3 if self.shell else self.dim + 1
        m = np.empty((self.nE, npe, npe), dtype=np.int32)
        st = np.empty((self.nE, npe), dtype=np.int32)
        self.L.orc_scatter_map(self.h, _ip(m), _ip(st))
        return m, st

    def timers(self, reset=False):
        ms = np.zeros(8)
        self.L.orc_timers(self.h, _dp(ms), int(reset))
        return dict(zip(TIMER_KEYS, ms.tolist()))


# ---- low-level probes ----
def newton_procrustes(R, F, compute_gradients=True, tol=1e-6, max_iter=100):
    L = lib()
    R = _f64(np.asarray(R).T).copy()  # column-major storage
    Fc = _f64(np.asarray(F).T).copy()
    d = np.zeros(81)
    st = np.zeros(1, dtype=np.int32)
    it = L.orc_newton_procrustes(_dp(R), _dp(Fc), int(compute_gradients), _dp(d), tol, max_iter, _ip(st))
    return R.reshape(3, 3).T.copy(), d.reshape(9, 9).T.copy(), it, int(st[0])


def polar_svd3(R, F, compute_gradients=True):
    L = lib()
    R = _f64(np.asarray(R).T).copy()
    Fc = _f64(np.asarray(F).T).copy()
    s = np.zeros(6)
    d = np.zeros(54)
    st = np.zeros(1, dtype=np.int32)
    it = L.orc_polar_svd3(_dp(R), _dp(s), _dp(Fc), int(compute_gradients), _dp(d), _ip(st))
    return R.reshape(3, 3).T.copy(), s, d.reshape(9, 6).T.copy(), it, int(st[0])  # dsdF as 6x9


def polar_svd2(F):
    L = lib()
    R = np.zeros(4)
    Fc = _f64(np.asarray(F).T).copy()
    s = np.zeros(3)
    d = np.zeros(12)
    L.orc_polar_svd2(_dp(R), _dp(s), _dp(Fc), _dp(d))
    return R.reshape(2, 2).T.copy(), s, d.reshape(4, 3).T.copy()  # dsdF as 3x4


def rodrigues(omega):
    R = np.zeros(9)
    om = _f64(omega)
    lib().orc_rodrigues(_dp(R), _dp(om))
    return R.reshape(3, 3).T.copy()


def colpiv_solve3(A, b):
    A = _f64(np.asarray(A).T).copy()
    b = _f64(b)
    x = np.zeros(3)
    lib().orc_colpiv_solve3(_dp(A), _dp(b), _dp(x))
    return x


def sym_eig(A):
    A = np.asarray(A, dtype=np.float64)
    n = A.shape[0]
    Ac = _f64(A.T).copy()
    ev = np.zeros(n)
    V = np.zeros(n * n)
    lib().orc_sym_eig(n, _dp(Ac), _dp(ev), _dp(V))
    return ev, V.reshape(n, n).T.copy()


def psd_fix_invert(A, tol=1e-8):
    A = np.asarray(A, dtype=np.float64)
    n = A.shape[0]
    Ac = _f64(A.T).copy()
    Ai = np.zeros(n * n)
    ev = np.zeros(n)
    lib().orc_psd_fix_invert(n, _dp(Ac), _dp(Ai), _dp(ev), tol)
    return Ac.reshape(n, n).T.copy(), Ai.reshape(n, n).T.copy(), ev


def svd2(F, flipfix=True):
    Fc = _f64(np.asarray(F).T).copy()
    sig = np.zeros(2)
    U = np.zeros(4)
    V = np.zeros(4)
    lib().orc_svd2(_dp(Fc), _dp(sig), _dp(U), _dp(V), int(flipfix))
    return U.reshape(2, 2).T.copy(), sig, V.reshape(2, 2).T.copy()


def material(model, mu, la, s):
    """returns (psi, grad, hess) of the S-form material at s (len 6 or 3)."""
    L = lib()
    m = MODELS[model] if isinstance(model, str) else int(model)
    s = _f64(s)
    n = s.size
    g = np.zeros(n)
    H = np.zeros(n * n)
    e = L.orc_material_energy(m, mu, la, n, _dp(s))
    L.orc_material_gradient(m, mu, la, n, _dp(s), _dp(g))
    L.orc_material_hessian(m, mu, la, n, _dp(s), _dp(H))
    return e, g, H.reshape(n, n).T.copy()
