"""GPU parity on the REFERENCE'S OWN inputs (copies under tests/golden/ref_inputs):
the configs of BASELINE.json as named - C1 scenes/2D/trampoline.json (two objects, two materials, hang BC), C2
models/mesh/beam_20k.mesh (100 793 unstructured tets), C5 models/obj/square_xz_16k.obj as a shell - the deformed
states the reference ships (boat, coarse bunny, gummy bear with per-tet material ids, the two-object drop scene,
the 2D diving board), and the one numeric fixture the reference holds (scripts/mats/*.mkt), against which the CUDA
path is compared DIRECTLY (no oracle in between).  Everything goes through the C ABI."""
import os

import numpy as np
import pytest

import mixed_fem_b200 as mf
from mixed_fem_b200 import meshgen, sceneio
from oracle import orc
from tests import refinputs as ri
from tests.common import DT, TOL, block_rel, vec_rel
from tests.test_gpu_parity import check_update

pytestmark = pytest.mark.gpu


def make_pair(V, T, fm, mats, mat_id=None, **kw):
    gpu = mf.MixedStretch(V, T, fm, mats, mat_id=mat_id, **kw)
    cpu = orc.OracleMixedStretch(V, T, fm, mats, mat_id=mat_id, **kw)
    for o in (gpu, cpu):
        o.reset()
    return gpu, cpu


def check_pattern(gpu, cpu):
    go, gi = gpu.csr_pattern()
    co, ci, _ = cpu.csr(values=False)
    assert np.array_equal(go, co) and np.array_equal(gi, ci)
    gm, gs = gpu.scatter_map()
    cm, cs = cpu.scatter_map()
    assert np.array_equal(gm, cm) and np.array_equal(gs, cs)


def run_state(gpu, cpu, st, dim, dt=DT, shell=False):
    for o in (gpu, cpu):
        if dim == 3:
            o.set("R", st["R"])
        o.set("s", st["s"]); o.set("la", st["la"])
        o.update(st["x"], dt)
    return check_update(gpu, cpu, dim)


def test_cuda_path_reproduces_the_reference_dump(built):
    """scripts/mats/*.mkt, written by the reference itself: Hloc_, jacobian_x(), jacobian_mixed(), gradient_mixed(),
    gradient_dual() of the CUDA path on the recovered state, and the device mass matrix P M P^T - no oracle involved."""
    st = ri.mkt_state()
    gpu = mf.MixedStretch(st["V"], st["T"], st["fm"], st["materials"])
    gpu.reset()
    gpu.set("s", st["s"])
    gpu.update(st["x"], st["h"])
    errs = ri.check_against_mkt_dump(gpu, st, jac_triplets=gpu.jacobian_x())
    print("CUDA path vs reference dump:", errs)
    # lhs_M: the Newton driver's P M P^T lives on the Assembler pattern (a superset of the reference's mass pattern)
    import scipy.sparse as sp
    gpu.newton_init(density=st["density"], h=st["h"])
    _, mptr = gpu.newton_positions_dev()
    outer, inner = gpu.csr_pattern()
    mvals = _copy_dev(mptr, gpu.nnz)
    Mg = sp.csr_matrix((mvals, inner, outer), shape=(2 * gpu.nfree, 2 * gpu.nfree))
    Mref = ri.mkt("lhs_M")
    assert abs(Mg - Mref).max() <= 1e-13 * abs(Mref).max()
    gpu.close()


def _copy_dev(ptr, n):
    """n doubles at a raw device pointer -> numpy (cudaMemcpy through torch's runtime)"""
    import ctypes
    import torch
    out = np.empty(n)
    rt = torch.cuda.cudart()
    torch.cuda.synchronize()
    err = rt.cudaMemcpy(out.ctypes.data, ptr, 8 * n, 2) if hasattr(rt, "cudaMemcpy") else None
    if err is None or int(err) != 0:
        lib = ctypes.CDLL("libcudart.so.12")
        lib.cudaMemcpy.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int]
        assert lib.cudaMemcpy(out.ctypes.data, ptr, 8 * n, 2) == 0
    return out


def test_C2_beam_20k_mesh_as_named(built):
    """models/mesh/beam_20k.mesh: 20 000 V / 100 793 unstructured tets (real valence distribution, graded elements),
    Stable-Neohookean, pinned end; pattern and scatter map bit-exact, every update field 1e-10."""
    V, T, _ = sceneio.read_mesh(os.path.join(ri.RI, "models/mesh/beam_20k.mesh"))
    V = V / V.max()                                        # simulation_state.cpp:69-73
    fm = sceneio.free_map_from_fixed(sceneio.fixed_vertices(V, "hangends", 0.1, 0))
    mu, la = meshgen.enu_to_lame(1e6, 0.45)
    gpu, cpu = make_pair(V, T, fm, [("Stable-Neohookean", mu, la)])
    check_pattern(gpu, cpu)
    edge = np.linalg.norm(V[T[:, 1]] - V[T[:, 0]], axis=1).mean()
    st = meshgen.synthetic_state(V, T, edge, mu)
    errs = run_state(gpu, cpu, st, 3)
    print("beam_20k:", errs, gpu.status())
    assert vec_rel(gpu.get("vols"), cpu.get("vols")) <= 1e-14
    # rhs / solve / line-search reductions on the same mesh
    dx = np.random.default_rng(3).standard_normal(3 * gpu.nfree) * 1e-3 * edge
    for o in (gpu, cpu):
        o.solve(dx)
    assert block_rel(gpu.get("la"), cpu.get("la"), 6) <= TOL
    s_trial = cpu.get("s") + 0.5 * cpu.get("ds")
    assert abs(gpu.energy(st["x"], s_trial) - cpu.energy(st["x"], s_trial)) <= TOL * abs(cpu.energy(st["x"], s_trial))
    cg, cc = gpu.constraint_value(st["x"], s_trial), cpu.constraint_value(st["x"], s_trial)
    assert abs(cg - cc) <= 1e-9 * max(abs(cc), 1e-12)
    gpu.close()


def test_C1_trampoline_scene_two_objects_two_materials_hang_bc(built):
    sc = ri.scene("scenes/2D/trampoline.json", 2)
    gpu, cpu = make_pair(sc.V, sc.T, sc.free_map, sc.materials, mat_id=sc.mat_id)
    check_pattern(gpu, cpu)
    edge = np.linalg.norm(sc.V[sc.T[:, 1]] - sc.V[sc.T[:, 0]], axis=1).mean()
    st = meshgen.synthetic_state(sc.V, sc.T, edge, sc.materials[0][1])
    errs = run_state(gpu, cpu, st, 2, dt=sc.h)
    print("trampoline:", errs)
    gpu.close()


def test_2d_diving_board_deformed_state(built):
    """scenes/2D/diving_board_deformed.json (FixedCorot x 2, hangends) at the saved states sim_x0_0001 / 0054"""
    sc = ri.scene("scenes/2D/diving_board_deformed.json", 2)
    gpu, cpu = make_pair(sc.V, sc.T, sc.free_map, sc.materials, mat_id=sc.mat_id)
    for name in ("sim_x0_0001", "sim_x0_0054"):
        x = sceneio.read_dmat(os.path.join(ri.RI, "scenes/2D/data/diving_board/%s.dmat" % name)).reshape(-1)
        st = ri.seeded_state(sc.V, sc.T, x, sc.materials[0][1])
        run_state(gpu, cpu, st, 2, dt=sc.h)
    gpu.close()


def test_C5_square_xz_16k_obj_as_shell(built):
    """models/obj/square_xz_16k.obj (8 321 V / 16 384 F) as TriMesh + MixedStretch<3>, thickness 1e-3, rest normals of
    the mesh; cloth constants of apps/cloth_sim.cpp:64-71"""
    V, T = sceneio.read_obj(os.path.join(ri.RI, "models/obj/square_xz_16k.obj"))
    V, T = sceneio.remove_unreferenced(V, T)
    N = meshgen.face_normals(V, T)
    fm = sceneio.free_map_from_fixed(sceneio.fixed_vertices(V, "hang", 0.1, 0))
    mu, la = meshgen.enu_to_lame(0.57e6, 0.3)
    mats = [("Stable-Neohookean", mu, la)]
    gpu, cpu = make_pair(V, T, fm, mats, normals=N, thickness=1e-3)
    check_pattern(gpu, cpu)
    rng = np.random.default_rng(4)
    L = V[:, 0].max() - V[:, 0].min()
    bump = 0.05 * L * np.sin(np.pi * (V[:, 0] - V[:, 0].min()) / L)
    x = (V * np.array([1.03, 1.0, 0.98]) + np.stack([0 * bump, bump, 0 * bump], axis=1)).reshape(-1)
    edge = np.linalg.norm(V[T[:, 1]] - V[T[:, 0]], axis=1).mean()
    x = x + 0.02 * edge * rng.standard_normal(x.size)
    st = ri.seeded_state(V, T, x, mu, normals=N)
    for o in (gpu, cpu):
        o.set("R", st["R"]); o.set("s", st["s"]); o.set("la", st["la"])
        o.update(st["x"], DT)
    errs = check_update(gpu, cpu, 3)
    print("square_xz_16k shell:", errs)
    gpu.close()


@pytest.mark.parametrize("mesh,states,E", [
    ("models/mesh/boat.mesh", ["scenes/data/boat_x_0009.dmat"], 1e5),
    ("models/mesh/coarse_bunny.mesh", ["scenes/data/bunny_x_0002.dmat", "scenes/data/bunny_x_0027.dmat", "scenes/data/bunny_x_0031.dmat"], 3e4),
])
def test_3d_deformed_states_on_their_meshes(built, mesh, states, E):
    """boat.mesh (11 295 T) + boat_x_0009, coarse_bunny.mesh (2 274 T) + bunny_x_{0002,0027,0031}: det F in (0.37, 2)"""
    V, T, _ = sceneio.read_mesh(os.path.join(ri.RI, mesh))
    V = V / V.max()
    fm = sceneio.free_map_from_fixed(sceneio.fixed_vertices(V, "hangends", 0.1, 0))
    mu, la = meshgen.enu_to_lame(E, 0.45)
    for model in ("Stable-Neohookean", "FixedCorot"):
        gpu, cpu = make_pair(V, T, fm, [(model, mu, la)])
        check_pattern(gpu, cpu)
        for sfile in states:
            x = sceneio.read_dmat(os.path.join(ri.RI, sfile)).reshape(-1)
            st = ri.seeded_state(V, T, x, mu)
            errs = run_state(gpu, cpu, st, 3)
        print(mesh, model, errs, gpu.status())
        gpu.close()


def test_gummy_bear_heterogeneous_materials_deformed(built):
    """scenes/3D/gummy_bear_deformed.json: per-tet material ids (E = 6e4 / 1e14), saved state sim_x0_0013, hangends"""
    sc = ri.scene("scenes/3D/gummy_bear_deformed.json", 3)
    gpu, cpu = make_pair(sc.V, sc.T, sc.free_map, sc.materials, mat_id=sc.mat_id)
    check_pattern(gpu, cpu)
    st = ri.seeded_state(sc.V, sc.T, sc.x0, sc.materials[0][1])
    errs = run_state(gpu, cpu, st, 3, dt=sc.h)
    print("gummy bear:", errs, gpu.status())
    gpu.close()


def test_two_object_scene_concatenation(built):
    """scenes/3D/collision_drop_bunny.json: cube_14.mesh + coarse_bunny.mesh concatenated (meshes.cpp:31-50), two
    materials, saved state sim_x0_0016 (contact itself is out of scope: only the elastic variable runs)"""
    sc = ri.scene("scenes/3D/collision_drop_bunny.json", 3)
    gpu, cpu = make_pair(sc.V, sc.T, sc.free_map, sc.materials, mat_id=sc.mat_id)
    check_pattern(gpu, cpu)
    x = sceneio.read_dmat(os.path.join(ri.RI, "scenes/3D/data/sim_x0_0016.dmat")).reshape(-1)   # the scene's (disabled) initial_state0
    st = ri.seeded_state(sc.V, sc.T, x, sc.materials[0][1])
    run_state(gpu, cpu, st, 3, dt=sc.h)
    gpu.close()


class _AsOracleVar:
    """lets oracle/newton.py - the restatement of the REFERENCE'S Newton loop (NewtonOptimizer::step, line search, BDF1,
    SystemMatrixPD with a direct solve) - drive the CUDA variable through the C ABI: the drop-in scenario itself."""

    def __init__(self, gpu):
        self.g = gpu

    def csr(self):
        return self.g.lhs()

    def __getattr__(self, name):
        return getattr(self.g, name)


def _trampoline():
    sc = ri.scene("scenes/2D/trampoline.json", 2)
    assert sc.h == 0.0333 and sorted(set(sc.density_e.tolist())) == [1e3, 1e5]
    return sc


def test_C1_trampoline_newton_history_reference_loop_drives_the_cuda_variable(built):
    """scenes/2D/trampoline.json as written (mixed-stretch only, enable_ccd off: contact is out of scope): densities
    1e5 / 1e3 per object, h = 0.0333, float gravity, hang BC; two time steps of the reference's Newton loop with an exact
    linear solve, once around the oracle variable and once around the CUDA variable: histories within 1e-8."""
    from oracle import newton as onewton
    from tests.common import compare_histories
    sc = _trampoline()
    gpu, cpu = make_pair(sc.V, sc.T, sc.free_map, sc.materials, mat_id=sc.mat_id)
    no = onewton.OracleNewton(cpu, sc.V, sc.T, sc.free_map, density=sc.density_e, h=sc.h, force=sc.body_force)
    ng = onewton.OracleNewton(_AsOracleVar(gpu), sc.V, sc.T, sc.free_map, density=sc.density_e, h=sc.h, force=sc.body_force)
    for outer in (2, 4):        # only the last step runs into the converged tail: a coin-flip alpha on a 7e-9 step moves x by 3e-9, and BDF1 turns that into a 1e-7 velocity difference for the next step
        lo = no.step(outer_steps=outer, ls_iters=sc.ls_iters)
        lg = ng.step(outer_steps=outer, ls_iters=sc.ls_iters)
        compare_histories(lg, lo, min_len=min(outer - 1, 2))
        assert vec_rel(ng.x, no.x) <= 1e-9
        assert vec_rel(gpu.get("s"), cpu.get("s")) <= 1e-9
    gpu.close()


def test_C1_trampoline_device_resident_newton_step_per_object_density(built):
    """the same scene through mfb_newton_init_density / mfb_newton_step (everything on the device, block-Jacobi PCG to
    1e-13 from a zero guess) against the oracle's loop with its direct solve"""
    from oracle import newton as onewton
    from tests.common import compare_histories
    sc = _trampoline()
    gpu, cpu = make_pair(sc.V, sc.T, sc.free_map, sc.materials, mat_id=sc.mat_id)
    no = onewton.OracleNewton(cpu, sc.V, sc.T, sc.free_map, density=sc.density_e, h=sc.h, force=sc.body_force)
    gpu.newton_init(density=sc.density_e, h=sc.h, force=sc.body_force)
    gpu.set_option("cg_zero_guess", 1)
    for step, outer in enumerate((2, 4)):
        lo = no.step(outer_steps=outer, ls_iters=sc.ls_iters)
        lg = gpu.newton_step(outer_steps=outer, ls_iters=sc.ls_iters, precond=2, cg_tol=1e-14, cg_max_iters=20000)
        print("trampoline device step", step, "cg iters", lg["cg_iters"], "energy", lg["energy"], lo["energy"])
        compare_histories(lg, lo, min_len=min(outer - 1, 2))
        assert vec_rel(gpu.newton_positions(), no.x) <= 1e-9
    # per-material densities through mfb_material::density (density=None) give the same mass matrix here, because
    # every object of this scene has ONE material
    mats4 = [m + (c["density"],) for m, c in zip(sc.materials, sc.material_configs)]
    g2 = mf.MixedStretch(sc.V, sc.T, sc.free_map, mats4, mat_id=sc.mat_id); g2.reset()
    g2.newton_init(density=None, h=sc.h, force=sc.body_force)
    g3 = mf.MixedStretch(sc.V, sc.T, sc.free_map, sc.materials, mat_id=sc.mat_id); g3.reset()
    g3.newton_init(density=sc.density_e, h=sc.h, force=sc.body_force)
    m2 = _copy_dev(g2.newton_positions_dev()[1], g2.nnz)
    m3 = _copy_dev(g3.newton_positions_dev()[1], g3.nnz)
    assert np.array_equal(m2, m3) and m2.max() > 50 * m2[m2 > 0].min()
    for g in (gpu, g2, g3):
        g.close()


def test_convergence_driver_gradient_norms_on_the_trampoline_scene(built, tmp_path):
    """apps/convergence.cpp (the reference's integration harness, DIM = 2): per Newton iteration the gradient norm of the
    NON-mixed system (Stretch variable, each element's own F-form material) at the mixed iterate.  tools/convergence.py on
    scenes/2D/trampoline.json against the oracle's Newton loop + OracleStretch; the DMAT it writes round-trips."""
    import importlib.util
    from oracle import newton as onewton
    from oracle import stretch as ostretch
    spec = importlib.util.spec_from_file_location("convergence", os.path.join(ri.HERE, "..", "tools", "convergence.py"))
    conv = importlib.util.module_from_spec(spec); spec.loader.exec_module(conv)
    path = os.path.join(ri.RI, "scenes/2D/trampoline.json")
    G = conv.run(path, 2, root=ri.SCENES_ROOT, dim=2, cg_tol=1e-14)
    sc = _trampoline()
    assert G.shape == (2, sc.outer_steps)
    cpu = orc.OracleMixedStretch(sc.V, sc.T, sc.free_map, sc.materials, mat_id=sc.mat_id); cpu.reset()
    no = onewton.OracleNewton(cpu, sc.V, sc.T, sc.free_map, density=sc.density_e, h=sc.h, force=sc.body_force)
    st = ostretch.OracleStretch(sc.V, sc.T, sc.free_map, materials=sc.materials, mat_id=sc.mat_id)
    for i in range(2):
        lo = no.step(outer_steps=sc.outer_steps, ls_iters=sc.ls_iters, stretch=st)
        ref = lo["fem_gradient_norm"]
        # the first iterations are far from the noise floor: 1e-8; once the step is converged (norms 1e-6 of the first) both
        # sides sit on rounding noise of the linear solves and only the order of magnitude is comparable
        big = ref > 1e-5 * ref[0]
        k = int(big.sum())
        assert k >= 2
        assert np.all(np.abs(G[i, :k] - ref[:k]) <= 1e-8 * ref[:k]), (G[i], ref)
    p = str(tmp_path / "convergence.dmat")
    sceneio.write_dmat(p, G)
    assert np.array_equal(sceneio.read_dmat(p), G)
