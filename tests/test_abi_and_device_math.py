"""CPU-side checks of the product: the C-ABI library builds for sm_100a, loads and exports every symbol declared in
include/mfem_b200.h (no compute calls without a GPU), fails loudly without a device, and the per-element device math
(mixed-fem_b200/csrc/mfb_math.cuh, compiled here as plain C++ by a TEST-ONLY harness) agrees with the oracle."""
import ctypes as C
import json
import os
import re
import subprocess

import numpy as np
import pytest
import scipy.linalg as sl

import mixed_fem_b200 as mf
from oracle import orc

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
dp = C.POINTER(C.c_double)


def P(a):
    return a.ctypes.data_as(dp)


def test_header_symbols_are_exported(built):
    hdr = open(os.path.join(ROOT, "include", "mfem_b200.h")).read()
    declared = sorted(set(re.findall(r"\b(mfb_[a-z_0-9]+)\s*\(", hdr)))
    assert len(declared) >= 30
    L = C.CDLL(mf.LIB_PATH)
    for name in declared:
        assert hasattr(L, name), name
    assert sorted(mf.EXPORTS) == declared


def test_library_is_sm100a_only(built):
    out = subprocess.run(["cuobjdump", "-lelf", mf.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_no_cpu_fallback_without_device(built):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from mixed_fem_b200 import meshgen
    V, T = meshgen.kuhn_grid(2, 2, 2)
    fm = meshgen.pinned_free_map(V)
    with pytest.raises(mf.MfbError) as ei:
        mf.MixedStretch(V, T, fm, [("ARAP", 1.0, 1.0)])
    assert "no CUDA device" in str(ei.value) or "error -" in str(ei.value)


def _create_outcome(V, T, fm, mats, **kw):
    """'ok' (a device is present and the context was built), 'nodev' (valid descriptor, no device) or the error text"""
    try:
        ms = mf.MixedStretch(V, T, fm, mats, **kw)
    except mf.MfbError as e:
        return "nodev" if "no CUDA device" in str(e) else str(e)
    ms.close()
    return "ok"


def test_create_validates_the_descriptor_before_touching_the_device(built):
    """mfb_create checks, on the host and before any CUDA call, everything a kernel would otherwise index with or divide by:
    vertex ids, material ids, the free map (ids 0..nF-1 exactly once, -1 = pinned), finite coordinates, non-degenerate
    rest shapes.  Runs without a GPU: a VALID descriptor gets as far as the no-device error."""
    from mixed_fem_b200 import meshgen
    mats = [("Stable-Neohookean", 1.0, 2.0), ("ARAP", 1.0, 1.0)]
    for dim in (3, 2):
        V, T = meshgen.kuhn_grid(3, 2, 2) if dim == 3 else meshgen.tri_grid(4, 3)
        fm = meshgen.pinned_free_map(V)
        assert _create_outcome(V, T, fm, mats[:1]) in ("ok", "nodev")
        assert _create_outcome(V, T, np.full(V.shape[0], -1, np.int32), mats[:1]) in ("ok", "nodev")      # everything pinned is a mesh too
        bad = T.copy(); bad[1, 2] = V.shape[0]
        assert "vertex out of range" in _create_outcome(V, bad, fm, mats[:1])
        bad = T.copy(); bad[0, 0] = -1
        assert "vertex out of range" in _create_outcome(V, bad, fm, mats[:1])
        mid = np.zeros(T.shape[0], np.int32); mid[-1] = 2
        assert "mat_id out of range" in _create_outcome(V, T, fm, mats, mat_id=mid)
        f2 = fm.copy(); free = np.flatnonzero(f2 >= 0); f2[free[1]] = f2[free[0]]
        assert "appears twice" in _create_outcome(V, T, f2, mats[:1]) or "skips ids" in _create_outcome(V, T, f2, mats[:1])
        f2 = fm.copy(); f2[free[-1]] += 3
        assert "skips ids" in _create_outcome(V, T, f2, mats[:1])
        f2 = fm.copy(); f2[0] = -2
        assert "< -1" in _create_outcome(V, T, f2, mats[:1])
        V2 = V.copy(); V2[3, 0] = np.nan
        assert "non-finite" in _create_outcome(V2, T, fm, mats[:1])
        V2 = V.copy(); V2[T[2, 1]] = V2[T[2, 0]]              # two vertices of element 2 coincide
        msg = _create_outcome(V2, T, fm, mats[:1])
        assert "degenerate" in msg and "element" in msg
    with pytest.raises(mf.MfbError) as ei:
        mf.MixedStretch(V[:0], T[:0], fm[:0], mats[:1])
    assert "invalid descriptor" in str(ei.value)
    assert "unknown material model" in _create_outcome(V, T, fm, [(9, 1.0, 1.0)])
    # shells: a sliver with collinear vertices
    Vs = np.array([[0, 0, 0], [1, 0, 0], [2, 0, 0], [0, 1, 0]], float)
    Ts = np.array([[0, 1, 3], [0, 1, 2]], np.int32)
    msg = _create_outcome(Vs, Ts, np.arange(4, dtype=np.int32), mats[:1], normals=np.tile([0.0, 0.0, 1.0], (2, 1)), thickness=1e-3)
    assert "element 1 is degenerate" in msg


def test_reference_input_meshes_pass_the_descriptor_checks(built):
    """the reference's own meshes (tests/golden/ref_inputs) have no degenerate element and consistent free maps under the
    rules of mfb_create - i.e. the checks above reject nothing the reference could run"""
    from tests import refinputs as ri
    from mixed_fem_b200 import sceneio
    for rel in ("models/mesh/beam_20k.mesh", "models/mesh/boat.mesh", "models/mesh/coarse_bunny.mesh", "models/mesh/gummy_bear.mesh"):
        path = os.path.join(ri.RI, rel)
        if not (os.path.exists(path) or os.path.exists(path + ".gz")):
            continue
        V, T, _ = sceneio.read_mesh(path)
        V, T = sceneio.remove_unreferenced(V, T)[:2]
        fm = np.arange(V.shape[0], dtype=np.int32)
        assert _create_outcome(V, T, fm, [("Stable-Neohookean", 1.0, 2.0)]) in ("ok", "nodev"), rel
    sc = ri.scene("scenes/2D/trampoline.json", 2)
    assert _create_outcome(sc.V, sc.T, sc.free_map, sc.materials, mat_id=sc.mat_id) in ("ok", "nodev")


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "mixed-fem_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".hpp", ".cpp")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "liborc" not in txt and "from oracle" not in txt and "import oracle" not in txt, f


@pytest.fixture(scope="module")
def harness():
    bdir = os.path.join(ROOT, "tests", "harness", "_build")
    os.makedirs(bdir, exist_ok=True)
    so = os.path.join(bdir, "libmfbhost.so")
    src = os.path.join(ROOT, "tests", "harness", "mfb_math_host.cpp")
    hdr = os.path.join(ROOT, "mixed-fem_b200", "csrc", "mfb_math.cuh")
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
        subprocess.run([cxx, "-O2", "-std=c++17", "-fPIC", "-shared", "-o", so, src], check=True)
    return C.CDLL(so)


def test_device_materials_match_reference_generated_code(harness):
    G = json.load(open(os.path.join(ROOT, "tests", "golden", "materials_golden.json")))
    for c in G["cases"]:
        n = c["n"]
        s = np.array(c["s"]); g = np.zeros(n); H = np.zeros(n * n); e = np.zeros(1)
        fn = harness.hh_material6 if n == 6 else harness.hh_material3
        fn(orc.MODELS[c["model"]], C.c_double(c["mu"]), C.c_double(c["la"]), P(s), C.c_double(1.0), P(g), P(H), P(e))
        He = np.array(c["hessian"])
        Hl = np.tril(He) + np.tril(He, -1).T
        assert abs(e[0] - c["energy"]) <= 1e-12 * max(1.0, abs(c["energy"]))
        assert np.linalg.norm(g - c["gradient"]) <= 1e-12 * max(1e-300, np.linalg.norm(c["gradient"])) or np.linalg.norm(g) < 1e-9
        assert np.linalg.norm(H.reshape(n, n).T - Hl) <= 1e-12 * np.linalg.norm(Hl)


def test_device_psd_fix_matches_oracle(harness):
    rng = np.random.default_rng(1)
    for n, fn in ((6, harness.hh_psd6), (3, harness.hh_psd3)):
        for _ in range(300):
            A = rng.standard_normal((n, n)); A = A + A.T + rng.uniform(0, n) * np.eye(n)
            Af, Ai, ev = orc.psd_fix_invert(A, 1e-8)
            H = np.ascontiguousarray(A.T).reshape(-1).copy(); Hi = np.zeros(n * n); lm = np.zeros(1)
            sw = fn(P(H), P(Hi), C.c_double(1e-8), P(lm))
            assert 0 < sw < 12
            assert np.linalg.norm(H.reshape(n, n).T - Af) <= 1e-11 * np.linalg.norm(Af)
            assert np.linalg.norm(Hi.reshape(n, n).T - Ai) <= 1e-10 * np.linalg.norm(Ai)
            assert abs(lm[0] - ev[-1]) <= 1e-12 * ev[-1]


def _snh_hessian(s, mu, la, h2, harness):
    g = np.zeros(6); H = np.zeros(36); e = np.zeros(1)
    harness.hh_material6(orc.MODELS["Stable-Neohookean"], C.c_double(mu), C.c_double(la), P(np.ascontiguousarray(s)), C.c_double(h2), P(g), P(H), P(e))
    return H.reshape(6, 6).T.copy()


def test_device_psd_fast_path_matches_oracle_and_rejects_what_it_must(harness):
    """psd_fast6 (Cholesky + tridiagonal Newton for lambda_max) either reproduces psd_fix_invert (psd_fix.h:10-34) to
    rounding or refuses: indefinite, near-singular, clustered-top and non-finite inputs must go to the Jacobi path."""
    rng = np.random.default_rng(11)
    ip = C.POINTER(C.c_int)

    def run(A, relpiv=1e-5):
        Hc = np.ascontiguousarray(A.T).reshape(-1).copy(); L = np.zeros(36); Hi = np.zeros(36); lm = np.zeros(1); it = np.zeros(1, dtype=np.int32)
        ok = harness.hh_psd_fast6(P(Hc), P(L), P(Hi), C.c_double(1e-8), C.c_double(relpiv), P(lm), it.ctypes.data_as(ip))
        return ok, L.reshape(6, 6).T, Hi.reshape(6, 6).T, lm[0], int(it[0])

    mu, la = 344827.5862068966, 3103448.2758620693
    n_ok = 0
    for k in range(600):
        if k % 2 == 0:   # material Hessians around realistic stretches
            S = np.diag([1.2, .9, 1.05]) + 0.08 * rng.standard_normal((3, 3)); S = 0.5 * (S + S.T)
            s = np.array([S[0, 0], S[1, 1], S[2, 2], S[0, 1], S[0, 2], S[1, 2]])
            A = _snh_hessian(s, mu, la, 0.0333 ** 2, harness)
        else:            # random SPD with a spread spectrum
            Q, _ = np.linalg.qr(rng.standard_normal((6, 6)))
            A = (Q * np.exp(rng.uniform(-3, 3, size=6))) @ Q.T; A = 0.5 * (A + A.T)
        ev = np.linalg.eigvalsh(A)
        ok, L, Hi, lm, it = run(A)
        if ev[0] < 1e-8:
            assert not ok
            continue
        if not ok:
            continue
        n_ok += 1
        Af, Ai, evo = orc.psd_fix_invert(A, 1e-8)
        assert np.linalg.norm(L @ L.T - Af) <= 1e-13 * np.linalg.norm(Af)
        assert np.linalg.norm(Hi - Ai) <= 1e-11 * np.linalg.norm(Ai) * max(1.0, ev[-1] / ev[0] * 1e-4)
        assert abs(lm - evo[-1]) <= 1e-12 * evo[-1], (lm, evo[-1], it)
        assert it <= 24
    assert n_ok > 400
    # must refuse: indefinite, singular, negative definite, NaN
    Q, _ = np.linalg.qr(rng.standard_normal((6, 6)))
    for lam in ([-1, 1, 2, 3, 4, 5], [0, 1, 2, 3, 4, 5], [1e-9, 1, 2, 3, 4, 5], [-5, -4, -3, -2, -1, -.5], [1, 2, 3, 4, 5, np.nan]):
        A = (Q * np.array(lam, dtype=float)) @ Q.T; A = 0.5 * (A + A.T)
        assert not run(A)[0], lam
    # clustered top eigenvalue: the polynomial root is ill-conditioned unless the tridiagonal form deflates.  Either the
    # path refuses, or what it returns is right.
    for lam in ([1, 2, 3, 5, 5, 5], [1, 2, 3, 4, 5, 5 * (1 + 1e-9)], [1, 2, 5, 5, 5, 5], [3, 3, 3, 3, 3, 3], [1, 1, 2, 2, 4, 4]):
        A = (Q * np.array(lam, dtype=float)) @ Q.T; A = 0.5 * (A + A.T)
        ok, L, Hi, lm, it = run(A)
        if ok:
            assert abs(lm - max(lam)) <= 1e-11 * max(lam), (lam, lm)
    # constant Hessians of the reference's materials (h^2 mu diag(1,1,1,2,2,2) for ARAP, + la 1 1^T on the normal block
    # for Corot): multiple top eigenvalue made of 1x1 blocks, must stay on the fast path
    D = np.diag([1., 1, 1, 2, 2, 2]); N1 = np.zeros((6, 6)); N1[:3, :3] = 1.0
    for A in (0.0333 ** 2 * mu * D, 0.0333 ** 2 * (2 * mu * D + la * N1), 0.0333 ** 2 * (2 * mu * D + 0.1 * mu * N1), np.eye(6)):
        ok, L, Hi, lm, it = run(A)
        ev = np.linalg.eigvalsh(A)
        assert ok and abs(lm - ev[-1]) <= 1e-13 * ev[-1]
        assert np.linalg.norm(Hi - np.linalg.inv(A)) <= 1e-12 * np.linalg.norm(np.linalg.inv(A))
    # accepted: diagonal / already tridiagonal inputs (Householder steps with nothing to annihilate), separated top
    for A in (np.diag([1., 2, 3, 4, 5, 7]), np.diag([3., 1, 4, 1.5, 9, 2.6]) + np.diag([.1] * 5, 1) + np.diag([.1] * 5, -1)):
        ok, L, Hi, lm, it = run(A)
        assert ok and abs(lm - np.linalg.eigvalsh(A)[-1]) <= 1e-13 * lm
        assert np.linalg.norm(Hi - np.linalg.inv(A)) <= 1e-13 * np.linalg.norm(Hi)


def test_device_polar3_contracted_dsdf_matches_explicit_9x9(harness):
    rng = np.random.default_rng(2)
    ip = C.POINTER(C.c_int)
    for _ in range(300):
        Q = sl.expm(np.cross(np.eye(3), rng.standard_normal(3) * 0.5))
        F = Q @ (np.diag([1.2, .9, 1.05]) + 0.2 * rng.standard_normal((3, 3)))
        R0 = sl.expm(np.cross(np.eye(3), rng.standard_normal(3) * 0.1)) @ sl.polar(F)[0]
        Ro, so, Jo, ito, sto = orc.polar_svd3(R0, F)
        R = np.ascontiguousarray(R0.T).reshape(-1).copy(); Fc = np.ascontiguousarray(F.T).reshape(-1)
        s6 = np.zeros(6); Js = np.zeros(54); st = np.zeros(1, dtype=np.int32)
        it = harness.hh_polar3(P(R), P(Fc), P(s6), P(Js), st.ctypes.data_as(ip))
        assert it == ito and st[0] == sto
        assert np.abs(R.reshape(3, 3).T - Ro).max() <= 1e-13
        assert np.linalg.norm(s6 - so) <= 1e-13 * np.linalg.norm(so)
        assert np.linalg.norm(Js.reshape(9, 6).T - Jo) <= 1e-12 * np.linalg.norm(Jo)


def test_rolled_dsdf_column_is_bit_identical(harness):
    """dsdf3_column_rt (run-time column index, kernel option MFB_V4_ROLL_DSDF) against the unrolled dsdf3_column"""
    rng = np.random.default_rng(7)
    for _ in range(200):
        R = np.ascontiguousarray(sl.expm(np.cross(np.eye(3), rng.standard_normal(3))).T).reshape(-1)
        F = rng.standard_normal(9); Hi = rng.standard_normal(6)
        assert harness.hh_dsdf_rt_mismatches(P(R), P(F), P(Hi)) == 0


def test_device_polar2_matches_dsvd_route(harness):
    rng = np.random.default_rng(3)
    for _ in range(300):
        F = np.eye(2) + 0.4 * rng.standard_normal((2, 2))
        Ro, so, Jo = orc.polar_svd2(F)
        Fc = np.ascontiguousarray(F.T).reshape(-1); R = np.zeros(4); s3 = np.zeros(3); Js = np.zeros(12)
        harness.hh_polar2(P(Fc), P(R), P(s3), P(Js))
        assert np.abs(R.reshape(2, 2).T - Ro).max() <= 1e-13
        assert np.linalg.norm(s3 - so) <= 1e-13 * np.linalg.norm(so)
        assert np.linalg.norm(Js.reshape(4, 3).T - Jo) <= 1e-11 * max(np.linalg.norm(Jo), 1e-3)
    # inside the |s1^2 - s0^2| < 1e-4 band the reference freezes dR (dsvd.cpp:163-167)
    F = np.array([[1.00001, 0.0], [0.0, 1.0]])
    Ro, so, Jo = orc.polar_svd2(F)
    Fc = np.ascontiguousarray(F.T).reshape(-1); R = np.zeros(4); s3 = np.zeros(3); Js = np.zeros(12)
    harness.hh_polar2(P(Fc), P(R), P(s3), P(Js))
    assert np.linalg.norm(Js.reshape(4, 3).T - Jo) <= 1e-12


def test_host_cpp_shim_compiles_and_links(built, tmp_path):
    """mixed-fem_b200/host/mixed_stretch.hpp (the C++ mirror of mfem::MixedStretch<DIM>) instantiates for DIM = 2, 3
    and links against the C-ABI library (not run here: no GPU)."""
    src = tmp_path / "shim.cpp"
    src.write_text('#include "mixed-fem_b200/host/mixed_stretch.hpp"\n'
                   'template class mfb::MixedStretch<3>;\ntemplate class mfb::MixedStretch<2>;\n'
                   'int main(int argc, char**) { if (argc > 99) { mfb::MeshView m; mfb::MixedStretch<3> v(m); v.reset(); '
                   'std::vector<double> x; v.update(x, 0.1); auto r = v.rhs(); v.solve(r); '
                   'mfb::MeshView sh; sh.shell = true; mfb::MixedStretch<3> w(sh); '
                   'mfb::NewtonStep<3> nt(v, m, 1000.0, 0.0333); auto lg = nt.step(mfb::PcgSolver::by_name("eigen-pcg-block_jacobi")); '
                   'mfb::SystemMatrixIndefinite<3> kkt(v); auto kb = kkt.b(); auto kx = kkt.solve(1e-8, 100); auto kp = kkt.product(kx); '
                   'mfb::DualCondensedSystem<3> dc(v, 6); auto db = dc.b(); auto dl = dc.solve(1e-8, 100); auto dp = dc.product(dl); '
                   'return (int)v.lhs().nnz + lg.n_iters + (int)nt.positions().size() + (int)(kb.size() + kp.size() + db.size() + dp.size()); } return 0; }\n')
    exe = tmp_path / "shim"
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    libdir = os.path.join(ROOT, "mixed-fem_b200")
    r = subprocess.run([cxx, "-std=c++17", "-I", ROOT, str(src), "-o", str(exe), "-L", libdir, "-lmfem_b200",
                        "-Wl,-rpath," + libdir, "-L/usr/local/cuda/lib64", "-Wl,-rpath-link,/usr/local/cuda/lib64"],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    # the program that actually RUNS the shim on the GPU box (tests/test_gpu_cpp_shim.py) must build here too
    r = subprocess.run([cxx, "-std=c++17", "-I", ROOT, os.path.join(ROOT, "tests", "harness", "shim_run.cpp"), "-o", str(tmp_path / "shim_run"),
                        "-L", libdir, "-lmfem_b200", "-Wl,-rpath," + libdir, "-L/usr/local/cuda/lib64", "-Wl,-rpath-link,/usr/local/cuda/lib64"],
                       capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
