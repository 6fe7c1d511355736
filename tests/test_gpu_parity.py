"""GPU parity tests proper: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.
Bars (BASELINE.json north_star): per-element fields and assembled CSR values <= 1e-10 relative, CSR pattern and
element->nonzero map bit-exact, procrustes iteration counts identical."""
import numpy as np
import pytest

from mixed_fem_b200 import meshgen
from tests.common import DT, LA, MU, TOL, block_rel, mesh, pair, vec_rel

pytestmark = pytest.mark.gpu

MODELS3 = ["Stable-Neohookean", "Neohookean", "Corot", "FixedCorot", "ARAP", "Fung"]
MODELS2 = ["Stable-Neohookean", "Corot", "FixedCorot", "ARAP"]


def check_update(gpu, cpu, dim):
    N, M = (6, 9) if dim == 3 else (3, 4)
    errs = {
        "R": block_rel(gpu.get("R"), cpu.get("R"), M),
        "S": block_rel(gpu.get("S"), cpu.get("S"), N),
        "dSdF": block_rel(gpu.get("dSdF"), cpu.get("dSdF"), M * N),
        "g": block_rel(gpu.get("g"), cpu.get("g"), N),
        "H": block_rel(gpu.get("H"), cpu.get("H"), N * N),
        "Hinv": block_rel(gpu.get("Hinv"), cpu.get("Hinv"), N * N),
        "Hloc": block_rel(gpu.get("Hloc"), cpu.get("Hloc"), N * N),
        "evals": vec_rel(gpu.get("evals"), cpu.get("evals")),
        "grad": vec_rel(gpu.get("grad"), cpu.get("grad")),
        "grad_la": vec_rel(gpu.get("grad_la"), cpu.get("grad_la")),
        "grad_x": vec_rel(gpu.get("grad_x"), cpu.get("grad_x")),
        "A": vec_rel(gpu.lhs()[2], cpu.csr()[2]),
        "rhs": vec_rel(gpu.rhs(), cpu.rhs()),
        "gl": vec_rel(gpu.get("gl"), cpu.get("gl")),
    }
    bad = {k: v for k, v in errs.items() if not v <= TOL}
    assert not bad, (bad, errs)
    if dim == 3:
        assert int((gpu.get("iters").astype(int) != cpu.iters()).sum()) == 0
    st = gpu.status()
    assert st["polar_maxiter"] == 0 and st["polar_hfail"] == 0 and st["nonfinite"] == 0
    return errs


@pytest.mark.parametrize("dim,n,shuffle,pin", [(3, (6, 4, 3), False, 0.01), (3, (5, 3, 3), True, 0.2), (3, (3, 3, 3), True, None),
                                               (2, (9, 7, 0), False, 0.01), (2, (8, 8, 0), True, None)])
def test_pattern_and_scatter_map_bit_exact(built, dim, n, shuffle, pin):
    V, T, fm, h = mesh(dim, n, shuffle=shuffle, jitter=0.1 if shuffle else 0.0, pin=pin)
    gpu, cpu, _ = pair(V, T, fm, [("ARAP", MU, LA)], h=h, state=False)
    go, gi = gpu.csr_pattern()
    co, ci, _ = cpu.csr(values=False)
    assert go.dtype == np.int32 and gi.dtype == np.int32
    assert np.array_equal(go, co) and np.array_equal(gi, ci)
    gm, gs = gpu.scatter_map()
    cm, cs = cpu.scatter_map()
    assert np.array_equal(gm, cm)
    assert np.array_equal(gs, cs)
    gpu.close()


@pytest.mark.parametrize("model", MODELS3)
def test_update_parity_tets(built, model):
    V, T, fm, h = mesh(3, (7, 4, 4))
    gpu, cpu, st = pair(V, T, fm, [(model, MU, LA)], h=h)
    for o in (gpu, cpu):
        o.update(st["x"], DT)
    check_update(gpu, cpu, 3)
    gpu.close()


@pytest.mark.parametrize("model", MODELS2)
def test_update_parity_tri2d(built, model):
    V, T, fm, h = mesh(2, (12, 9, 0))
    gpu, cpu, st = pair(V, T, fm, [(model, MU, LA)], h=h)
    for o in (gpu, cpu):
        o.update(st["x"], DT)
    check_update(gpu, cpu, 2)
    gpu.close()


def test_update_parity_irregular_mesh_heterogeneous_materials(built):
    V, T, fm, h = mesh(3, (6, 5, 4), shuffle=True, jitter=0.15, pin=0.15)
    mats = [("Stable-Neohookean", MU, LA), ("ARAP", 2 * MU, LA), ("FixedCorot", 0.5 * MU, 0.3 * LA)]
    mat_id = (np.arange(T.shape[0]) % 3).astype(np.int32)
    gpu, cpu, st = pair(V, T, fm, mats, mat_id=mat_id, h=h)
    for o in (gpu, cpu):
        o.update(st["x"], DT)
    check_update(gpu, cpu, 3)
    gpu.close()


@pytest.mark.parametrize("variant,asm", [(3, 1), (3, 2), (3, 0)])
def test_update_kernel_variants_agree_with_oracle(built, variant, asm):
    """every assembly mode of the tet update kernel (option asm_mode; smem_b 3 = k_update_tet_v4 is the only kernel since the
    round-1 variants were removed): default = bulk reductions into the block-CSR upper triangle; the mesh is irregular with
    pinned vertices; about a third of the elements get an indefinite Hessian so both the Cholesky fast path and the Jacobi
    tail are exercised (status psd_slow counts the latter)."""
    V, T, fm, h = mesh(3, (9, 6, 5), shuffle=True, jitter=0.12, pin=0.12)
    gpu, cpu, st = pair(V, T, fm, [("Stable-Neohookean", MU, LA)], h=h)
    gpu.set_option("smem_b", variant)
    gpu.set_option("asm_mode", asm)
    rng = np.random.default_rng(7)
    s = st["s"].reshape(-1, 6).copy()
    sel = rng.uniform(size=s.shape[0]) < 0.33
    s[sel, :3] *= rng.uniform(0.45, 0.8, size=(int(sel.sum()), 1))
    for it in range(2):      # second pass: warm R, values accumulate into freshly zeroed targets again
        for o in (gpu, cpu):
            o.set("s", s.reshape(-1))
            o.update(st["x"] if it == 0 else st["xprev"], DT)
        check_update(gpu, cpu, 3)
    stt = gpu.status()
    if variant == 3:
        nslow = stt["psd_slow"]
        assert 0 < nslow < gpu.nE, stt
        ev = np.linalg.eigvalsh(cpu.get("H").reshape(-1, 6, 6))
        assert nslow >= int(sel.sum()) // 4
    gpu.close()


def test_removed_kernel_variants_are_rejected(built):
    import mixed_fem_b200 as mf
    V, T, fm, h = mesh(3, (3, 2, 2))
    gpu = mf.MixedStretch(V, T, fm, [("ARAP", MU, LA)])
    for v in (0, 1, 2):
        with pytest.raises(mf.MfbError):
            gpu.set_option("smem_b", v)
    gpu.set_option("smem_b", 3)
    gpu.close()


def test_indefinite_hessians_are_clamped_identically(built):
    """compressed state: Stable-Neohookean Hessians with negative eigenvalues exercise the 1e-8 clamp (psd_fix.h:20-25)"""
    V, T, fm, h = mesh(3, (5, 4, 3))
    gpu, cpu, st = pair(V, T, fm, [("Stable-Neohookean", MU, LA)], h=h)
    rng = np.random.default_rng(5)
    s = st["s"].reshape(-1, 6).copy()
    s[:, :3] *= rng.uniform(0.45, 0.8, size=(s.shape[0], 1))
    s[:, 3:] += rng.uniform(-0.3, 0.3, size=(s.shape[0], 3))
    for o in (gpu, cpu):
        o.set("s", s.reshape(-1))
        o.update(st["x"], DT)
    Hinv = cpu.get("Hinv").reshape(-1, 36)
    assert (np.abs(Hinv).max(axis=1) > 1e6).any(), "test state does not reach the clamp"
    check_update(gpu, cpu, 3)
    gpu.close()


def test_newton_like_sequence_rhs_solve_linesearch(built):
    """update -> rhs -> solve -> (s += a ds) -> update ... : the call order of NewtonOptimizer::step
    (newton_optimizer.cpp:36-101) with a random dx standing in for the linear solve."""
    V, T, fm, h = mesh(3, (6, 4, 3))
    gpu, cpu, st = pair(V, T, fm, [("Stable-Neohookean", MU, LA)], h=h)
    rng = np.random.default_rng(11)
    x = st["x"].copy()
    for it in range(3):
        for o in (gpu, cpu):
            o.update(x, DT)
        assert vec_rel(gpu.rhs(), cpu.rhs()) <= TOL
        dx = rng.standard_normal(3 * gpu.nfree) * 1e-3 * h
        for o in (gpu, cpu):
            o.solve(dx)
        assert block_rel(gpu.get("la"), cpu.get("la"), 6) <= TOL
        # ds = -Hinv (g - Sym la): where the PSD fix clamps an eigenvalue to 1e-8, |Hinv| ~ 1e8 amplifies the rounding of
        # (g - Sym la) in BOTH implementations, so the bar per element is TOL relative plus eps-level input noise
        # pushed through |Hinv|: |dds| <= TOL |ds| + 64 eps |Hinv|_F (|g| + |Sym la|)
        dg, dc = gpu.get("ds").reshape(-1, 6), cpu.get("ds").reshape(-1, 6)
        Hi = np.linalg.norm(cpu.get("Hinv").reshape(-1, 36), axis=1)
        gn = np.linalg.norm(cpu.get("g").reshape(-1, 6), axis=1) + 2 * np.linalg.norm(cpu.get("la").reshape(-1, 6), axis=1)
        bound = TOL * np.linalg.norm(dc, axis=1) + 64 * 2.2e-16 * Hi * gn
        assert (np.linalg.norm(dg - dc, axis=1) <= bound).all()
        s_trial = cpu.get("s") + 0.5 * cpu.get("ds")     # same trial point for both: this checks energy(), not ds
        eg = gpu.energy(x, s_trial)
        ec = cpu.energy(x, s_trial)
        assert abs(eg - ec) <= TOL * abs(ec)
        cg = gpu.constraint_value(x, gpu.get("s"))
        cc = cpu.constraint_value(x, cpu.get("s"))
        assert abs(cg - cc) <= 1e-9 * max(abs(cc), 1e-12)
        s_new = cpu.get("s") + 0.5 * cpu.get("ds")
        for o in (gpu, cpu):
            o.set("s", s_new)
        # move the free vertices a little
        xf = x.reshape(-1, 3)
        xf[fm >= 0] += dx.reshape(-1, 3)
    check_update(gpu, cpu, 3)
    gpu.close()


@pytest.mark.parametrize("dim", [3, 2])
def test_unfused_rhs_equals_fused(built, dim):
    V, T, fm, h = mesh(dim, (6, 4, 3))
    gpu, cpu, st = pair(V, T, fm, [("Stable-Neohookean", MU, LA)], h=h)
    gpu.update(st["x"], DT)
    r1 = gpu.rhs()
    gpu.set_option("fuse_rhs", 0)
    gpu.update(st["x"], DT)
    r2 = gpu.rhs()
    cpu.update(st["x"], DT)
    assert vec_rel(r2, r1) <= 1e-13 and vec_rel(r2, cpu.rhs()) <= TOL
    gpu.close()


@pytest.mark.parametrize("dim", [3, 2])
def test_reset_evaluate_constraint_and_products(built, dim):
    V, T, fm, h = mesh(dim, (5, 4, 3), shuffle=True, jitter=0.1, pin=0.2)
    gpu, cpu, st = pair(V, T, fm, [("FixedCorot", MU, LA)], h=h, state=False)
    N, M = (6, 9) if dim == 3 else (3, 4)
    for f, w in (("R", M), ("S", N), ("s", N), ("dSdF", M * N)):
        assert block_rel(gpu.get(f), cpu.get(f), w) <= TOL, f
    for o in (gpu, cpu):
        o.set("la", st["la"])
        o.update(st["x"], DT)
    assert vec_rel(gpu.evaluate_constraint(st["x"]), cpu.evaluate_constraint(st["x"])) <= TOL
    rng = np.random.default_rng(2)
    xe = rng.standard_normal(N * gpu.nE)
    xv = rng.standard_normal(dim * gpu.nfree)
    for which, xin, out0, tr in (("hessian", xe, xe * 0.5, False), ("hessian_inv", xe, xe * 0.1, False),
                                 ("jacobian_mixed", xe, xe * 2, False), ("jacobian_x", xe, xv, True), ("jacobian_x", xv, xe, False)):
        a = gpu.product(which, xin, out0, tr)
        b = cpu.product(which, xin, out0, tr)
        assert vec_rel(a, b) <= TOL, which
    gpu.close()


def test_post_solve_and_setters(built):
    V, T, fm, h = mesh(3, (4, 3, 3))
    gpu, cpu, st = pair(V, T, fm, [("ARAP", MU, LA)], h=h)
    assert np.array_equal(gpu.get("la"), st["la"]) and np.array_equal(gpu.get("s"), st["s"]) and np.array_equal(gpu.get("R"), st["R"])
    gpu.post_solve()
    assert not gpu.get("la").any()
    with pytest.raises(Exception):
        gpu.set("H", np.zeros(36 * gpu.nE))
    gpu.close()


def test_unsupported_2d_material_is_an_error(built):
    import mixed_fem_b200 as mf
    V, T, fm, h = mesh(2, (4, 4, 0))
    with pytest.raises(mf.MfbError):
        mf.MixedStretch(V, T, fm, [("Neohookean", MU, LA)])


@pytest.mark.parametrize("model", ["Stable-Neohookean", "ARAP", "Neohookean"])
def test_update_parity_shell(built, model):
    """config C5: TriMesh + MixedStretch<3> (triangles in 3D with rest normals, src/mesh/tri_mesh.cpp): F = J0 x + n (x) N,
    x-dependent local Jacobian, 9 x 9 element blocks.  update / rhs / solve / energy / constraint against the oracle."""
    import mixed_fem_b200 as mf
    from mixed_fem_b200 import meshgen
    from oracle import orc
    V, T, N = meshgen.shell_grid(14, 9)
    fm = meshgen.pinned_free_map(V, ratio=0.05)
    mu, la = meshgen.enu_to_lame(0.57e6, 0.3)          # apps/cloth_sim.cpp:64-71
    mats = [(model, mu, la)]
    gpu = mf.MixedStretch(V, T, fm, mats, normals=N, thickness=0.5e-3)
    cpu = orc.OracleMixedStretch(V, T, fm, mats, normals=N, thickness=0.5e-3)
    rng = np.random.default_rng(3)
    x = (V @ meshgen._affine().T + 0.01 * rng.standard_normal(V.shape)).reshape(-1)
    xprev = (V @ meshgen._affine().T + 0.005 * rng.standard_normal(V.shape)).reshape(-1)
    lam = rng.uniform(-1, 1, size=6 * T.shape[0]) * mu * 1e-3
    for o in (gpu, cpu):
        o.reset()
    go, gi = gpu.csr_pattern()
    co, ci, _ = cpu.csr(values=False)
    assert np.array_equal(go, co) and np.array_equal(gi, ci)
    gm, gs = gpu.scatter_map()
    cm, cs = cpu.scatter_map()
    assert np.array_equal(gm, cm) and np.array_equal(gs, cs)
    for f, w in (("R", 9), ("S", 6), ("s", 6), ("dSdF", 54)):
        assert block_rel(gpu.get(f), cpu.get(f), w) <= TOL, f
    for o in (gpu, cpu):
        o.update(xprev, DT)
        o.set("s", cpu.get("S") if o is cpu else gpu.get("S"))
        o.set("la", lam)
        o.update(x, DT)
    check_update(gpu, cpu, 3)
    dx = rng.standard_normal(3 * gpu.nfree) * 1e-3
    for o in (gpu, cpu):
        o.solve(dx)
    assert block_rel(gpu.get("la"), cpu.get("la"), 6) <= TOL
    assert block_rel(gpu.get("ds"), cpu.get("ds"), 6) <= 1e-8
    s_trial = cpu.get("s") + 0.5 * cpu.get("ds")
    x_trial = x + 1e-3 * rng.standard_normal(x.size)
    assert abs(gpu.energy(x_trial, s_trial) - cpu.energy(x_trial, s_trial)) <= TOL * abs(cpu.energy(x_trial, s_trial))
    cg, cc = gpu.constraint_value(x_trial, s_trial), cpu.constraint_value(x_trial, s_trial)
    assert abs(cg - cc) <= 1e-9 * max(abs(cc), 1e-12)
    # rhs() recomputed after the trial evaluation must still use the Jacobian of the update (TriMesh::update_jacobian)
    gpu.set_option("fuse_rhs", 0)
    gpu.update(x, DT)
    assert vec_rel(gpu.rhs(), cpu.rhs()) <= TOL
    xe = rng.standard_normal(6 * gpu.nE); xv = rng.standard_normal(3 * gpu.nfree)
    for which, xin, out0, tr in (("jacobian_x", xe, xv, True), ("jacobian_x", xv, xe, False)):
        assert vec_rel(gpu.product(which, xin, out0, tr), cpu.product(which, xin, out0, tr)) <= TOL, which
    gpu.close()


def test_host_update_with_chunked_upload_equals_device_update(built):
    """mfb_update(x_host) on >= 2^18 tets uploads x in pieces on a copy stream and launches the element chunks as their
    vertices arrive, and mfb_rhs(out_host) copies rhs_ out beside the CSR finalisation: same results as the plain
    device-pointer path (mfb_update_dev on the whole mesh), up to the summation order of the reductions."""
    import torch
    import mixed_fem_b200 as mf
    from mixed_fem_b200 import meshgen
    V, T = meshgen.kuhn_grid(36, 36, 36, h=1.0 / 36)
    assert T.shape[0] >= (1 << 18)
    fm = meshgen.pinned_free_map(V)
    xa = meshgen.deformed_positions(V, 1.0 / 36, scale=1.0)
    xb = meshgen.deformed_positions(V, 1.0 / 36, scale=0.5)
    res = []
    for host in (True, False):
        ms = mf.MixedStretch(V, T, fm, [("Stable-Neohookean", MU, LA)])
        ms.reset(xb)
        for it in range(3):                                   # repeated host updates reuse the staging buffers
            x = xa if it % 2 == 0 else xb
            if host:
                ms.update(x, DT)
                r = ms.rhs()
            else:
                xd = torch.as_tensor(x, device="cuda")
                ms.update_dev(xd.data_ptr(), DT)
                torch.cuda.synchronize()
                r = ms.rhs()
        res.append((r.copy(), ms.lhs()[2].copy(), ms.get("grad_x").copy(), ms.get("R").copy(), ms.get("Hinv").copy()))
        assert ms.status()["nonfinite"] == 0
        ms.close()
    for a, b in zip(res[0], res[1]):
        assert vec_rel(a, b) <= 1e-12


@pytest.mark.parametrize("kind", ["tet", "tri2d", "shell"])
def test_explicit_jacobian_x_matches_the_products(built, kind):
    """jacobian_x() = -PJW_ blockdiag(dSdF_) (mixed_stretch.cpp:385-390) exported as triplets: applied to random vectors it
    must reproduce product_jacobian_x in both directions, on the GPU and on the oracle."""
    import scipy.sparse as sp
    import mixed_fem_b200 as mf
    from mixed_fem_b200 import meshgen
    from oracle import orc
    rng = np.random.default_rng(9)
    kw = {}
    if kind == "tet":
        V, T = meshgen.kuhn_grid(5, 3, 2); d, N = 3, 6
    elif kind == "tri2d":
        V, T = meshgen.tri_grid(7, 5); d, N = 2, 3
    else:
        V, T, Nrm = meshgen.shell_grid(7, 5); d, N = 3, 6
        kw = dict(normals=Nrm, thickness=1e-3)
    fm = meshgen.pinned_free_map(V, ratio=0.05)
    mats = [("Stable-Neohookean", MU, LA)]
    gpu = mf.MixedStretch(V, T, fm, mats, **kw)
    cpu = orc.OracleMixedStretch(V, T, fm, mats, **kw)
    x = (V * 1.05 + 0.01 * rng.standard_normal(V.shape)).reshape(-1)
    for o in (gpu, cpu):
        o.reset()
        o.update(x, DT)
    r, c, v = gpu.jacobian_x()
    nE, nF = T.shape[0], gpu.nfree
    J = sp.coo_matrix((v, (r, c)), shape=(d * nF, N * nE)).tocsr()
    xe = rng.standard_normal(N * nE)
    xv = rng.standard_normal(d * nF)
    for o in (gpu, cpu):
        a = o.product("jacobian_x", xe, np.zeros(d * nF), transposed=True)
        assert vec_rel(J @ xe, a) <= 1e-11
        b = o.product("jacobian_x", xv, np.zeros(N * nE), transposed=False)
        assert vec_rel(J.T @ xv, b) <= 1e-11
    gpu.close()


def test_deterministic_assembly_is_bit_stable_and_matches_the_oracle(built):
    """asm_mode 2: owner-computes gather per CSR block and per vertex in ascending element order - the reference's own
    structure (sparse_utils.cpp:124-170, no atomics).  Two contexts, repeated updates: np.array_equal on the CSR values
    and the vertex vectors (the RED modes agree only to rounding), and the usual 1e-10 against the oracle.  Irregular mesh,
    pinned vertices, a third of the Hessians on the Jacobi tail."""
    import mixed_fem_b200 as mf
    V, T, fm, h = mesh(3, (9, 6, 5), shuffle=True, jitter=0.12, pin=0.12)
    mats = [("Stable-Neohookean", MU, LA)]
    gpu, cpu, st = pair(V, T, fm, mats, h=h)
    gpu.set_option("asm_mode", 2)
    rng = np.random.default_rng(7)
    s = st["s"].reshape(-1, 6).copy()
    sel = rng.uniform(size=s.shape[0]) < 0.33
    s[sel, :3] *= rng.uniform(0.45, 0.8, size=(int(sel.sum()), 1))
    runs = []
    for rep in range(3):
        g = gpu if rep == 0 else mf.MixedStretch(V, T, fm, mats)
        if rep:
            g.reset(); g.set("R", st["R"]); g.set("la", st["la"])
            g.set_option("asm_mode", 2)
        g.set("s", s.reshape(-1))
        g.update(st["x"], DT)
        runs.append((g.lhs()[2].copy(), g.rhs().copy(), g.get("grad_x").copy()))
        if rep == 0:
            cpu.set("s", s.reshape(-1)); cpu.update(st["x"], DT)
            check_update(g, cpu, 3)
            assert 0 < g.status()["psd_slow"] < g.nE
            g.set("R", st["R"])              # same warm start again (update() moved R_ to the converged rotations)
            g.update(st["x"], DT)            # again on the same context: nothing accumulates, same bits
            assert np.array_equal(g.lhs()[2], runs[0][0]) and np.array_equal(g.rhs(), runs[0][1])
        else:
            g.close()
    for a, b in zip(runs[0], runs[1]):
        assert np.array_equal(a, b)
    for a, b in zip(runs[0], runs[2]):
        assert np.array_equal(a, b)
    gpu.close()


@pytest.mark.parametrize("shuffle", [False, True])
def test_staged_finalize_matches_the_single_pass(built, shuffle, monkeypatch):
    """Above 2^18 tets update() can run in element chunks (8 here) and finalize, behind each one on a side stream, the block-rows
    whose last element lies in it (option stage_finalize, default 1; per-chunk Jacobi tails; double-buffered accumulator).
    Same CSR values, vectors and status as the single pass (stage_finalize 0) - to rounding, the reductions are atomic in
    both - on a coherent element order and on a shuffled one (every row final in the last chunks), through the host-x entry
    (chunked upload) and the device-x one, three updates in a row (both accumulators in use), a third of the Hessians slow."""
    import torch
    import mixed_fem_b200 as mf
    monkeypatch.setenv("MFB_CHUNKS", "8")        # a mesh of this size would get one chunk (one per 1.25M tets); read at mfb_create
    V, T, fm, h = mesh(3, (40, 40, 30), shuffle=shuffle, jitter=0.1, pin=0.02)
    assert T.shape[0] >= (1 << 18)
    mats = [("Stable-Neohookean", MU, LA)]
    st = meshgen.synthetic_state(V, T, h, MU)
    rng = np.random.default_rng(3)
    s = st["s"].reshape(-1, 6).copy()
    sel = rng.uniform(size=s.shape[0]) < 0.3
    s[sel, :3] *= rng.uniform(0.45, 0.8, size=(int(sel.sum()), 1))
    xs = [st["x"], 0.5 * (st["x"] + st["xprev"]), st["x"]]
    out = {}
    for mode in (0, 1):
        g = mf.MixedStretch(V, T, fm, mats)
        g.reset(); g.set("R", st["R"]); g.set("la", st["la"]); g.set("s", s.reshape(-1))
        g.set_option("stage_finalize", mode)
        res = []
        for i, x in enumerate(xs):
            if i == 1:
                xd = torch.as_tensor(x, device="cuda")
                g.update_dev(xd.data_ptr(), DT)
            else:
                g.update(x, DT)
            stt = g.status()
            res.append((g.lhs()[2].copy(), g.rhs().copy(), g.get("grad_x").copy(), stt["psd_slow"], g.get("Hinv").copy()))
        out[mode] = res
        g.close()
    for a, b in zip(out[0], out[1]):
        assert 0 < a[3] < T.shape[0] and a[3] == b[3]
        assert np.array_equal(a[4], b[4])                       # per-element results do not depend on the launch shape
        for i in range(3):
            assert np.abs(a[i] - b[i]).max() <= 1e-12 * np.abs(a[i]).max()
    # the three updates differ (nothing cached), and the third repeats the first state with a warm-started polar
    assert np.abs(out[1][0][0] - out[1][1][0]).max() > 1e-6 * np.abs(out[1][0][0]).max()
