"""mixed-fem_b200 — B200-native (sm_100a, FP64) mixed-stretch hot path of evouga/mixed-fem.

This package is a thin ctypes binding over the C-ABI shared library `libmfem_b200.so` (include/mfem_b200.h)
plus `MixedStretch`, a host-side mirror of `mfem::MixedStretch<DIM>` (src/variables/mixed_stretch.h:12-164)
with the reference's method names.  There is NO CPU fallback: if the library or a CUDA device is missing the
calls raise.  The package directory name contains a hyphen; import it as `mixed_fem_b200` (repo-root shim).
"""
import ctypes as C
import os
import subprocess

import numpy as np

from . import meshgen  # noqa: F401

_HERE = os.path.dirname(os.path.abspath(__file__))
# MFB_LIB: a differently configured build of the same library (kernel experiments, tools/build_variant.sh)
LIB_PATH = os.environ.get("MFB_LIB") or os.path.join(_HERE, "libmfem_b200.so")

MODELS = {"Stable-Neohookean": 0, "Neohookean": 1, "Corot": 2, "FixedCorot": 3, "ARAP": 4, "Fung": 5}
FIELDS = {n: i for i, n in enumerate(
    ["s", "ds", "la", "R", "S", "g", "H", "Hinv", "Hloc", "dSdF", "evals", "gl", "grad", "grad_la", "grad_x", "rhs",
     "vols", "iters"])}

EXPORTS = [
    "mfb_create", "mfb_destroy", "mfb_set_stream", "mfb_last_error_string", "mfb_version", "mfb_dims",
    "mfb_csr_pattern", "mfb_scatter_map", "mfb_reset", "mfb_update", "mfb_update_dev", "mfb_rhs", "mfb_rhs_dev",
    "mfb_solve", "mfb_solve_dev", "mfb_post_solve", "mfb_energy", "mfb_constraint_value", "mfb_energy_dev",
    "mfb_constraint_value_dev", "mfb_evaluate_constraint", "mfb_lhs_values", "mfb_lhs_values_dev", "mfb_product",
    "mfb_field_size", "mfb_get", "mfb_set", "mfb_field_dev", "mfb_last_status", "mfb_spmv_dev", "mfb_pcg_dev",
    "mfb_num_kernel_launches", "mfb_set_option", "mfb_kernel_time_ms", "mfb_phase_name", "mfb_phase_times",
    "mfb_newton_init", "mfb_newton_init_density", "mfb_newton_step", "mfb_newton_positions", "mfb_newton_positions_dev", "mfb_newton_free",
    "mfb_jacobian_x_nnz", "mfb_jacobian_x_coo", "mfb_lhs_bsr",
    "mfb_kkt_size", "mfb_kkt_product", "mfb_kkt_rhs", "mfb_kkt_layout_dev", "mfb_kkt_product_dev", "mfb_kkt_solve_minres",
    "mfb_dual_product", "mfb_dual_product_dev", "mfb_dual_rhs", "mfb_dual_solve_cg",
    "mfb_stretch_update", "mfb_stretch_update_dev", "mfb_stretch_lhs_values", "mfb_stretch_gradient", "mfb_stretch_dev", "mfb_stretch_energy",
    "mfb_comm_unique_id", "mfb_comm_init", "mfb_comm_set_interface", "mfb_comm_set_partition", "mfb_comm_peer_state", "mfb_interface_sum_dev", "mfb_allreduce_dev",
    "mfb_partition_create", "mfb_partition_destroy", "mfb_partition_sizes", "mfb_partition_get",
    "mfb_energy_noalloc_dev", "mfb_noalloc_max_elements",
]


class MfbError(RuntimeError):
    pass


class _Material(C.Structure):
    _fields_ = [("model", C.c_int32), ("mu", C.c_double), ("la", C.c_double), ("density", C.c_double)]


class _MeshDesc(C.Structure):
    _fields_ = [("elem_kind", C.c_int32), ("nV", C.c_int32), ("nE", C.c_int32),
                ("T", C.POINTER(C.c_int32)), ("Vref", C.POINTER(C.c_double)), ("free_map", C.POINTER(C.c_int32)),
                ("mat_id", C.POINTER(C.c_int32)), ("n_mat", C.c_int32), ("materials", C.POINTER(_Material)),
                ("device", C.c_int32), ("normals", C.POINTER(C.c_double)), ("thickness", C.c_double)]


class _Status(C.Structure):
    _fields_ = [("polar_maxiter", C.c_int64), ("polar_hfail", C.c_int64), ("nonfinite", C.c_int64),
                ("polar_iters", C.c_int64), ("eig_sweeps", C.c_int64), ("psd_slow", C.c_int64)]


NEWTON_LOG_MAX = 64


class _NewtonOpts(C.Structure):
    _fields_ = [("outer_steps", C.c_int32), ("ls_iters", C.c_int32), ("newton_tol", C.c_double), ("precond", C.c_int32),
                ("cg_tol", C.c_double), ("cg_max_iters", C.c_int32)]


class _NewtonLog(C.Structure):
    _fields_ = [("n_iters", C.c_int32), ("cg_iters_total", C.c_int64),
                ("energy", C.c_double * NEWTON_LOG_MAX), ("energy_res", C.c_double * NEWTON_LOG_MAX),
                ("decrement", C.c_double * NEWTON_LOG_MAX), ("alpha", C.c_double * NEWTON_LOG_MAX),
                ("cg_error", C.c_double * NEWTON_LOG_MAX),
                ("cg_iters", C.c_int32 * NEWTON_LOG_MAX), ("ls_iters", C.c_int32 * NEWTON_LOG_MAX),
                ("fem_gradient_norm", C.c_double * NEWTON_LOG_MAX)]


def build(force=False, verbose=False):
    """Compile csrc/*.cu into libmfem_b200.so for sm_100a (nvcc cross-compiles without a GPU)."""
    srcdir = os.path.join(_HERE, "csrc")
    srcs = [os.path.join(srcdir, f) for f in os.listdir(srcdir) if f.endswith((".cu", ".cuh"))]
    srcs.append(os.path.join(_HERE, "..", "include", "mfem_b200.h"))
    if not force and os.path.exists(LIB_PATH) and all(os.path.getmtime(LIB_PATH) >= os.path.getmtime(s) for s in srcs):
        return LIB_PATH
    r = subprocess.run(["make", "-C", srcdir, "-j4"], capture_output=True, text=True)
    if r.returncode != 0:
        raise MfbError("nvcc build failed:\n" + r.stdout[-4000:] + r.stderr[-4000:])
    if verbose:
        print(r.stdout[-2000:])
    return LIB_PATH


_lib = None


def lib():
    """Load the C-ABI library (fails loudly when it is missing: no CPU path exists)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise MfbError("libmfem_b200.so not built: run `python -c 'import __graft_entry__ as g; g.build()'`")
    L = C.CDLL(LIB_PATH)
    vp, dp, ip = C.c_void_p, C.POINTER(C.c_double), C.POINTER(C.c_int32)
    L.mfb_create.argtypes = [C.POINTER(vp), C.POINTER(_MeshDesc)]
    L.mfb_destroy.argtypes = [vp]
    L.mfb_destroy.restype = None
    L.mfb_set_stream.argtypes = [vp, vp]
    L.mfb_last_error_string.restype = C.c_char_p
    L.mfb_version.restype = C.c_char_p
    L.mfb_dims.argtypes = [vp, ip, ip, ip, ip, ip, C.POINTER(C.c_int64)]
    L.mfb_csr_pattern.argtypes = [vp, C.POINTER(ip), C.POINTER(ip), C.POINTER(C.c_int64), ip]
    L.mfb_scatter_map.argtypes = [vp, C.POINTER(ip), C.POINTER(ip)]
    L.mfb_reset.argtypes = [vp, dp]
    L.mfb_update.argtypes = [vp, dp, C.c_double]
    L.mfb_lhs_bsr.argtypes = [vp, vp, ip, ip, dp]
    L.mfb_jacobian_x_nnz.argtypes = [vp]
    L.mfb_jacobian_x_nnz.restype = C.c_int64
    L.mfb_jacobian_x_coo.argtypes = [vp, ip, ip, dp]
    L.mfb_comm_unique_id.argtypes = [C.c_char_p]
    L.mfb_comm_init.argtypes = [vp, C.c_char_p, C.c_int, C.c_int]
    L.mfb_comm_set_interface.argtypes = [vp, ip, ip, C.c_int32, C.c_int32, dp]
    L.mfb_interface_sum_dev.argtypes = [vp, vp, C.c_int, vp, C.c_int]
    L.mfb_allreduce_dev.argtypes = [vp, vp, C.c_int, C.c_int]
    L.mfb_update_dev.argtypes = [vp, vp, C.c_double]
    L.mfb_rhs.argtypes = [vp, dp]
    L.mfb_rhs_dev.argtypes = [vp, C.POINTER(vp)]
    L.mfb_solve.argtypes = [vp, dp]
    L.mfb_solve_dev.argtypes = [vp, vp]
    L.mfb_post_solve.argtypes = [vp]
    L.mfb_energy.argtypes = [vp, dp, dp, dp]
    L.mfb_constraint_value.argtypes = [vp, dp, dp, dp]
    L.mfb_energy_dev.argtypes = [vp, vp, dp]
    L.mfb_energy_noalloc_dev.argtypes = [vp, vp, vp]
    L.mfb_noalloc_max_elements.restype = C.c_int64
    L.mfb_constraint_value_dev.argtypes = [vp, vp, vp, dp]
    L.mfb_evaluate_constraint.argtypes = [vp, dp, dp]
    L.mfb_lhs_values.argtypes = [vp, dp]
    L.mfb_lhs_values_dev.argtypes = [vp, C.POINTER(vp)]
    L.mfb_product.argtypes = [vp, C.c_int, dp, dp, C.c_int]
    L.mfb_field_size.argtypes = [vp, C.c_int]
    L.mfb_field_size.restype = C.c_int64
    L.mfb_get.argtypes = [vp, C.c_int, dp]
    L.mfb_set.argtypes = [vp, C.c_int, dp]
    L.mfb_field_dev.argtypes = [vp, C.c_int, C.POINTER(vp), C.POINTER(C.c_int64), ip]
    L.mfb_last_status.argtypes = [vp, C.POINTER(_Status)]
    L.mfb_spmv_dev.argtypes = [vp, vp, vp, vp]
    L.mfb_pcg_dev.argtypes = [vp, vp, vp, vp, C.c_int, C.c_double, C.c_int, ip, dp]
    L.mfb_num_kernel_launches.argtypes = [vp, C.POINTER(C.c_int64), C.c_int]
    L.mfb_set_option.argtypes = [vp, C.c_char_p, C.c_int]
    L.mfb_kernel_time_ms.argtypes = [vp, dp, ip]
    L.mfb_phase_name.argtypes = [C.c_int]
    L.mfb_phase_name.restype = C.c_char_p
    L.mfb_phase_times.argtypes = [vp, dp, ip, C.c_int]
    L.mfb_newton_init.argtypes = [vp, ip, C.c_double, C.c_double, dp, dp, dp]
    L.mfb_newton_init_density.argtypes = [vp, ip, dp, C.c_double, dp, dp, dp]
    L.mfb_newton_step.argtypes = [vp, C.POINTER(_NewtonOpts), C.POINTER(_NewtonLog)]
    L.mfb_newton_positions.argtypes = [vp, dp]
    L.mfb_newton_positions_dev.argtypes = [vp, C.POINTER(vp), C.POINTER(vp)]
    L.mfb_newton_free.argtypes = [vp]
    L.mfb_newton_free.restype = None
    i64p = C.POINTER(C.c_int64)
    L.mfb_kkt_size.argtypes = [vp]
    L.mfb_kkt_size.restype = C.c_int64
    L.mfb_kkt_product.argtypes = [vp, dp, dp]
    L.mfb_kkt_rhs.argtypes = [vp, dp]
    L.mfb_kkt_layout_dev.argtypes = [vp, i64p, i64p, i64p, i64p]
    L.mfb_kkt_product_dev.argtypes = [vp, vp, vp]
    L.mfb_kkt_solve_minres.argtypes = [vp, dp, C.c_double, C.c_int, ip, dp]
    L.mfb_dual_product.argtypes = [vp, dp, dp]
    L.mfb_dual_product_dev.argtypes = [vp, vp, vp]
    L.mfb_dual_rhs.argtypes = [vp, dp]
    L.mfb_dual_solve_cg.argtypes = [vp, dp, C.c_double, C.c_int, ip, dp, dp, dp]
    L.mfb_stretch_update.argtypes = [vp, dp, C.c_double]
    L.mfb_stretch_update_dev.argtypes = [vp, vp, C.c_double]
    L.mfb_stretch_lhs_values.argtypes = [vp, dp]
    L.mfb_stretch_gradient.argtypes = [vp, dp]
    L.mfb_stretch_dev.argtypes = [vp, C.POINTER(vp), C.POINTER(vp)]
    L.mfb_stretch_energy.argtypes = [vp, dp, dp]
    _lib = L
    return L


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _ip(a):
    return a.ctypes.data_as(C.POINTER(C.c_int32))


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class MixedStretch:
    """Host-side mirror of mfem::MixedStretch<DIM> over the C ABI (same method names and argument meaning).

    materials: list of (model_name, mu, la[, density]); mat_id: per-element index or None.
    """

    name = "mixed-stretch"   # src/variables/mixed_stretch.h:21-23

    def __init__(self, V, T, free_map, materials, mat_id=None, device=0, normals=None, thickness=1e-3):
        """normals (nE x 3) given => shell: TriMesh + MixedStretch<3> (src/mesh/tri_mesh.cpp), T is nE x 3, V is nV x 3."""
        self.L = lib()
        V = _f64(V)
        T = np.ascontiguousarray(T, dtype=np.int32)
        fm = np.ascontiguousarray(free_map, dtype=np.int32)
        self.dim = V.shape[1]
        self.nV, self.nE = V.shape[0], T.shape[0]
        self.N = 6 if self.dim == 3 else 3
        self.M = self.dim * self.dim
        mats = (_Material * len(materials))()
        for i, m in enumerate(materials):           # (model, mu, la) or (model, mu, la, density)
            model, mu, la = m[0], m[1], m[2]
            mats[i].model = MODELS[model] if isinstance(model, str) else int(model)
            mats[i].mu, mats[i].la = float(mu), float(la)
            mats[i].density = float(m[3]) if len(m) > 3 else 1000.0
        mid = None if mat_id is None else np.ascontiguousarray(mat_id, dtype=np.int32)
        d = _MeshDesc()
        self.shell = normals is not None
        d.elem_kind = 2 if self.shell else (0 if self.dim == 3 else 1)
        nrm = _f64(normals) if self.shell else None
        d.normals = _dp(nrm) if self.shell else None
        d.thickness = float(thickness)
        d.nV, d.nE = self.nV, self.nE
        d.T, d.Vref, d.free_map = _ip(T), _dp(V), _ip(fm)
        d.mat_id = _ip(mid) if mid is not None else None
        d.n_mat, d.materials, d.device = len(materials), mats, device
        self.device = device
        self.h = C.c_void_p()
        self._check(self.L.mfb_create(C.byref(self.h), C.byref(d)))
        nF, nnz = C.c_int32(), C.c_int64()
        self.L.mfb_dims(self.h, None, None, None, C.byref(nF), None, C.byref(nnz))
        self.nfree, self.nnz = nF.value, nnz.value
        self._free_map = fm

    # ---- device-resident Newton time step (NewtonOptimizer<DIM>::step, newton_optimizer.cpp:14-139) ----
    def newton_init(self, density=1000.0, h=0.0333, force=None, x0=None, v0=None):
        """Displacement::reset + BDF1 + mass matrix (displacement.cpp:73-93); force = body-force d-vector (default gravity).
        density: a scalar, None (each element takes its material's density) or one value per element (the density of the
        element's OBJECT in a multi-object scene, sceneio.Scene.density_e)."""
        f = None if force is None else _f64(force)
        x0 = None if x0 is None else _f64(x0)
        v0 = None if v0 is None else _f64(v0)
        tail = (float(h), _dp(f) if f is not None else None, _dp(x0) if x0 is not None else None, _dp(v0) if v0 is not None else None)
        if density is not None and np.ndim(density) > 0:
            rho = _f64(density)
            if rho.size != self.nE:
                raise MfbError("density: expected one value per element")
            self._check(self.L.mfb_newton_init_density(self.h, _ip(self._free_map), _dp(rho), *tail))
        else:
            self._check(self.L.mfb_newton_init(self.h, _ip(self._free_map), 0.0 if density is None else float(density), *tail))

    def newton_step(self, outer_steps=5, ls_iters=20, newton_tol=1e-10, precond=1, cg_tol=1e-4, cg_max_iters=500):
        """One time step; returns the per-iteration columns the reference prints (newton_optimizer.cpp:92-97)."""
        o = _NewtonOpts(outer_steps, ls_iters, newton_tol, precond, cg_tol, cg_max_iters)
        log = _NewtonLog()
        self._check(self.L.mfb_newton_step(self.h, C.byref(o), C.byref(log)))
        n = min(log.n_iters, NEWTON_LOG_MAX)
        return {"n_iters": log.n_iters, "cg_iters_total": log.cg_iters_total,
                "energy": np.array(log.energy[:n]), "energy_res": np.array(log.energy_res[:n]),
                "decrement": np.array(log.decrement[:n]), "alpha": np.array(log.alpha[:n]),
                "cg_iters": np.array(log.cg_iters[:n]), "cg_error": np.array(log.cg_error[:n]), "ls_iters": np.array(log.ls_iters[:n]),
                "fem_gradient_norm": np.array(log.fem_gradient_norm[:n])}

    def newton_positions_dev(self):
        """(device pointer of x, d*nV ; device pointer of the PMP^T values on the Assembler pattern)"""
        xp, mp = C.c_void_p(), C.c_void_p()
        self._check(self.L.mfb_newton_positions_dev(self.h, C.byref(xp), C.byref(mp)))
        return xp.value, mp.value

    def newton_positions(self):
        out = np.empty(self.dim * self.nV)
        self._check(self.L.mfb_newton_positions(self.h, _dp(out)))
        return out

    def _check(self, rc):
        if rc != 0:
            raise MfbError("mfem_b200 error %d: %s" % (rc, self.L.mfb_last_error_string().decode()))

    def close(self):
        if getattr(self, "h", None):
            self.L.mfb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- MixedVariable interface ----
    def reset(self, x=None):
        if x is None:
            self._check(self.L.mfb_reset(self.h, None))
        else:
            x = _f64(x)
            self._check(self.L.mfb_reset(self.h, _dp(x)))

    def update(self, x, dt):
        x = _f64(x)
        self._check(self.L.mfb_update(self.h, _dp(x), float(dt)))

    def update_dev(self, x_ptr, dt):
        self._check(self.L.mfb_update_dev(self.h, C.c_void_p(x_ptr), float(dt)))

    def rhs(self, out=None):
        out = np.empty(self.dim * self.nfree) if out is None else out
        self._check(self.L.mfb_rhs(self.h, _dp(out)))
        return out

    def rhs_dev(self):
        p = C.c_void_p()
        self._check(self.L.mfb_rhs_dev(self.h, C.byref(p)))
        return p.value

    def solve(self, dx):
        dx = _f64(dx)
        self._check(self.L.mfb_solve(self.h, _dp(dx)))

    def solve_dev(self, dx_ptr):
        self._check(self.L.mfb_solve_dev(self.h, C.c_void_p(dx_ptr)))

    def post_solve(self):
        self._check(self.L.mfb_post_solve(self.h))

    def energy(self, x, s):
        s = _f64(s)
        e = C.c_double()
        self._check(self.L.mfb_energy(self.h, None, _dp(s), C.byref(e)))
        return e.value

    def constraint_value(self, x, s):
        x, s = _f64(x), _f64(s)
        e = C.c_double()
        self._check(self.L.mfb_constraint_value(self.h, _dp(x), _dp(s), C.byref(e)))
        return e.value

    def energy_dev(self, s_ptr=None):
        e = C.c_double()
        self._check(self.L.mfb_energy_dev(self.h, C.c_void_p(s_ptr) if s_ptr else None, C.byref(e)))
        return e.value

    def energy_noalloc_dev(self, e_ptr, s_ptr=None):
        """allocation-free energy(): result written to the device double at e_ptr, stream-ordered, no host sync (no counterpart
        in the reference; see include/mfem_b200.h)"""
        self._check(self.L.mfb_energy_noalloc_dev(self.h, C.c_void_p(s_ptr) if s_ptr else None, C.c_void_p(e_ptr)))

    def constraint_value_dev(self, x_ptr, s_ptr=None):
        e = C.c_double()
        self._check(self.L.mfb_constraint_value_dev(self.h, C.c_void_p(x_ptr), C.c_void_p(s_ptr) if s_ptr else None, C.byref(e)))
        return e.value

    def evaluate_constraint(self, x):
        x = _f64(x)
        out = np.empty(self.N * self.nE)
        self._check(self.L.mfb_evaluate_constraint(self.h, _dp(x), _dp(out)))
        return out

    def gradient(self):
        return self.get("grad_x")

    def gradient_mixed(self):
        return self.get("grad")

    def gradient_dual(self):
        return self.get("grad_la")

    def value(self):
        return self.get("s")

    def delta(self):
        return self.get("ds")

    def lambda_(self):
        return self.get("la")

    def max_stresses(self):
        return self.get("evals")

    def size(self):
        return self.N * self.nE

    def size_dual(self):
        return self.N * self.nE

    def lhs(self):
        """(outer, inner, values) of the RowMajor CSR matrix A_ (mixed_stretch.h:37-39)."""
        outer, inner = self.csr_pattern()
        vals = np.empty(self.nnz)
        self._check(self.L.mfb_lhs_values(self.h, _dp(vals)))
        return outer, inner, vals

    def lhs_bsr(self, M_ptr=None):
        """(ptr, col, val[nb, d, d]) block-CSR of A (+ M): the amgcl static_matrix layout (amgcl_solver.h:40-75)."""
        d = self.dim
        nb = self.nnz // (d * d)
        ptr = np.empty(self.nfree + 1, dtype=np.int32); col = np.empty(max(nb, 1), dtype=np.int32); val = np.empty(max(nb, 1) * d * d)
        self._check(self.L.mfb_lhs_bsr(self.h, C.c_void_p(M_ptr) if M_ptr else None, _ip(ptr), _ip(col), _dp(val)))
        return ptr, col[:nb], val[:nb * d * d].reshape(nb, d, d)

    def lhs_values_dev(self):
        p = C.c_void_p()
        self._check(self.L.mfb_lhs_values_dev(self.h, C.byref(p)))
        return p.value

    def product(self, which, x, out, transposed=False):
        idx = {"hessian": 0, "hessian_inv": 1, "jacobian_mixed": 2, "jacobian_x": 3}[which]
        x = _f64(x)
        out = _f64(out).copy()
        self._check(self.L.mfb_product(self.h, idx, _dp(x), _dp(out), int(transposed)))
        return out

    # ---- assembler products ----
    def csr_pattern(self):
        po, pi = C.POINTER(C.c_int32)(), C.POINTER(C.c_int32)()
        nnz, nrows = C.c_int64(), C.c_int32()
        self._check(self.L.mfb_csr_pattern(self.h, C.byref(po), C.byref(pi), C.byref(nnz), C.byref(nrows)))
        outer = np.ctypeslib.as_array(po, shape=(nrows.value + 1,)).copy()
        inner = np.ctypeslib.as_array(pi, shape=(max(nnz.value, 1),)).copy()[:nnz.value]
        return outer, inner

    def scatter_map(self):
        pm, ps = C.POINTER(C.c_int32)(), C.POINTER(C.c_int32)()
        self._check(self.L.mfb_scatter_map(self.h, C.byref(pm), C.byref(ps)))
        npe = 3 if self.shell else self.dim + 1
        m = np.ctypeslib.as_array(pm, shape=(self.nE, npe, npe)).copy()
        st = np.ctypeslib.as_array(ps, shape=(self.nE, npe)).copy()
        return m, st

    # ---- state ----
    def get(self, name):
        f = FIELDS[name]
        n = self.L.mfb_field_size(self.h, f)
        out = np.empty(n)
        self._check(self.L.mfb_get(self.h, f, _dp(out)))
        return out

    def set(self, name, val):
        val = _f64(val)
        self._check(self.L.mfb_set(self.h, FIELDS[name], _dp(val)))

    def field_dev(self, name):
        p, ld, nc = C.c_void_p(), C.c_int64(), C.c_int32()
        self._check(self.L.mfb_field_dev(self.h, FIELDS[name], C.byref(p), C.byref(ld), C.byref(nc)))
        return p.value, ld.value, nc.value

    def status(self):
        st = _Status()
        self._check(self.L.mfb_last_status(self.h, C.byref(st)))
        return {k: getattr(st, k) for k, _ in _Status._fields_}

    def set_stream(self, stream_ptr):
        self._check(self.L.mfb_set_stream(self.h, C.c_void_p(stream_ptr)))

    def set_option(self, name, value):
        self._check(self.L.mfb_set_option(self.h, name.encode(), int(value)))

    def kernel_time_ms(self):
        t, n = C.c_double(), C.c_int32()
        self._check(self.L.mfb_kernel_time_ms(self.h, C.byref(t), C.byref(n)))
        return t.value, n.value

    def phase_times(self, reset=True):
        """{key: (total ms, count)} under the reference's Timer keys (recorded while option "profile" is 1)"""
        n = 7
        ms = np.zeros(n); cnt = np.zeros(n, dtype=np.int32)
        self._check(self.L.mfb_phase_times(self.h, _dp(ms), _ip(cnt), int(reset)))
        return {self.L.mfb_phase_name(k).decode(): (float(ms[k]), int(cnt[k])) for k in range(n)}

    def launches(self, reset=False):
        n = C.c_int64()
        self._check(self.L.mfb_num_kernel_launches(self.h, C.byref(n), int(reset)))
        return n.value

    def jacobian_x(self):
        """explicit -PJW_ * blockdiag(dSdF_) (mixed_stretch.cpp:385-390): (rows, cols, vals) triplets with the pinned rows
        dropped; shape (d*nF, N*nE)."""
        n = int(self.L.mfb_jacobian_x_nnz(self.h))
        r = np.empty(n, dtype=np.int32); c = np.empty(n, dtype=np.int32); v = np.empty(n)
        self._check(self.L.mfb_jacobian_x_coo(self.h, _ip(r), _ip(c), _dp(v)))
        keep = r >= 0
        return r[keep], c[keep], v[keep]

    # ---- the non-mixed Stretch variable (src/variables/stretch.cpp) on the same mesh and pattern ----
    def stretch_update(self, x, h):
        x = _f64(x)
        self._check(self.L.mfb_stretch_update(self.h, _dp(x), h))

    def stretch_lhs(self):
        """(outer, inner, values) of Stretch::lhs() on the Assembler pattern"""
        outer, inner = self.csr_pattern()
        vals = np.empty(self.nnz)
        self._check(self.L.mfb_stretch_lhs_values(self.h, _dp(vals)))
        return outer, inner, vals

    def stretch_gradient(self):
        out = np.empty(self.dim * self.nfree)
        self._check(self.L.mfb_stretch_gradient(self.h, _dp(out)))
        return out

    def stretch_energy(self, x):
        x = _f64(x)
        e = C.c_double(0)
        self._check(self.L.mfb_stretch_energy(self.h, _dp(x), C.byref(e)))
        return e.value

    # ---- saddle-point systems (SystemMatrixIndefinite / DualCondensedSystem, linear_system.h:64-239); need newton_init ----
    def kkt_size(self):
        return int(self.L.mfb_kkt_size(self.h))

    def kkt_product(self, v):
        """BlockMatrix product (block_matrix.h:84-124) on [dx ; ds ; la] in the reference's ordering."""
        v = _f64(v)
        out = np.empty(self.kkt_size())
        self._check(self.L.mfb_kkt_product(self.h, _dp(v), _dp(out)))
        return out

    def kkt_rhs(self):
        out = np.empty(self.kkt_size())
        self._check(self.L.mfb_kkt_rhs(self.h, _dp(out)))
        return out

    def kkt_solve_minres(self, guess=None, tol=1e-10, max_iters=10000):
        """pre_solve + MINRES + post_solve; returns (solution [dx ; ds ; la], iterations, relative residual)."""
        sol = np.zeros(self.kkt_size()) if guess is None else _f64(guess).copy()
        it = C.c_int32(0); err = C.c_double(0)
        self._check(self.L.mfb_kkt_solve_minres(self.h, _dp(sol), tol, max_iters, C.byref(it), C.byref(err)))
        return sol, it.value, err.value

    def dual_product(self, la):
        la = _f64(la)
        out = np.empty(self.N * self.nE)
        self._check(self.L.mfb_dual_product(self.h, _dp(la), _dp(out)))
        return out

    def dual_rhs(self):
        out = np.empty(self.N * self.nE)
        self._check(self.L.mfb_dual_rhs(self.h, _dp(out)))
        return out

    def dual_solve_cg(self, guess=None, tol=1e-10, max_iters=10000):
        """DualCondensedSystem pre_solve + CG + post_solve; returns (la, dx, ds, iterations, relative residual)."""
        la = np.zeros(self.N * self.nE) if guess is None else _f64(guess).copy()
        dx = np.empty(self.dim * self.nfree); ds = np.empty(self.N * self.nE)
        it = C.c_int32(0); err = C.c_double(0)
        self._check(self.L.mfb_dual_solve_cg(self.h, _dp(la), tol, max_iters, C.byref(it), C.byref(err), _dp(dx), _dp(ds)))
        return la, dx, ds, it.value, err.value

    # ---- multi-GPU (one process per GPU; `local` is this rank's partition.LocalProblem) ----
    def comm_init(self, local):
        """NCCL communicator inside the library (id broadcast through torch.distributed: plumbing) + the interface
        description.  Afterwards pcg_dev / newton_step are collective calls."""
        import torch
        import torch.distributed as dist
        rank, world = dist.get_rank(), dist.get_world_size()
        idbuf = C.create_string_buffer(128)
        if rank == 0:
            self._check(self.L.mfb_comm_unique_id(idbuf))
        t = torch.frombuffer(bytearray(idbuf.raw), dtype=torch.uint8).clone()
        if dist.get_backend() == "nccl":
            t = t.cuda(self.device)
        dist.broadcast(t, 0)
        raw = bytes(t.cpu().numpy().tobytes())
        self._check(self.L.mfb_comm_init(self.h, raw, rank, world))
        if getattr(local, "h", None):          # the library's own partition object: neighbour lists, interface-first elements
            self.L.mfb_comm_set_partition.argtypes = [C.c_void_p, C.c_void_p]
            self._check(self.L.mfb_comm_set_partition(self.h, local.h))
            return
        il = np.ascontiguousarray(local.if_local, dtype=np.int32)
        ipos = np.ascontiguousarray(local.if_pos, dtype=np.int32)
        w = _f64(local.dot_weight)
        self._check(self.L.mfb_comm_set_interface(self.h, _ip(il), _ip(ipos), int(il.size), int(local.n_shared), _dp(w)))

    def peer_state(self):
        """1: the CG exchanges through peer mailboxes (CUDA IPC), -1: unavailable on this node (NCCL), 0: not tried yet"""
        self.L.mfb_comm_peer_state.argtypes = [C.c_void_p]
        self.L.mfb_comm_peer_state.restype = C.c_int
        return int(self.L.mfb_comm_peer_state(self.h))

    def interface_sum_dev(self, a_ptr, ncomp_a, b_ptr=None, ncomp_b=0):
        self._check(self.L.mfb_interface_sum_dev(self.h, C.c_void_p(a_ptr), ncomp_a, C.c_void_p(b_ptr) if b_ptr else None, ncomp_b))

    # ---- linear solver side (device pointers) ----
    def spmv_dev(self, x_ptr, y_ptr, M_ptr=None):
        self._check(self.L.mfb_spmv_dev(self.h, C.c_void_p(M_ptr) if M_ptr else None, C.c_void_p(x_ptr), C.c_void_p(y_ptr)))

    def pcg_dev(self, b_ptr, x_ptr, M_ptr=None, precond=1, tol=1e-4, max_iters=500):
        it, rr = C.c_int32(), C.c_double()
        self._check(self.L.mfb_pcg_dev(self.h, C.c_void_p(M_ptr) if M_ptr else None, C.c_void_p(b_ptr), C.c_void_p(x_ptr),
                                       precond, tol, max_iters, C.byref(it), C.byref(rr)))
        return it.value, rr.value
