"""CPU tests on the reference's OWN inputs: the readers (.mesh / .obj / .dmat / .mkt / scene JSON), the scene loader's
BC / material / density bookkeeping, and the oracle against the one numeric artefact the reference holds
(scripts/mats/*.mkt, a 2D KKT dump) - which makes the 2D path of the oracle REFERENCE-PINNED:
  lhs_M  = P M P^T ............ consistent mass matrix on doublearea 'volumes' x density (tri2d_mesh.cpp:32-41)  exact
  lhs_K  = hessian() .......... Hloc_ = vol h^2 d2psi/ds2 (FixedCorot 3-vector form, psd_fix_invert)              1e-10
  lhs_Gs = jacobian_mixed() ... vol Sym                                                                          exact
  lhs_Gx = jacobian_x() ....... -P J^T W blockdiag(dSdF_): dphi/dX, local Jacobian, 2D polar + its derivative     1e-10
  rhs_gs, rhs_gl .............. -gradient_mixed(), -gradient_dual()                                               1e-10
The state (x, s) behind the dump is recovered by tests/golden/make_mkt_state.py."""
import gzip
import hashlib
import json
import os

import numpy as np
import pytest
import scipy.sparse as sp

from mixed_fem_b200 import meshgen, sceneio
from oracle import orc
from oracle.newton import mass_matrix
from tests import refinputs as ri


def test_manifest_matches_the_copies():
    man = json.load(open(os.path.join(ri.RI, "MANIFEST.json")))["files"]
    assert len(man) >= 35
    for rel, meta in man.items():
        raw = gzip.open(os.path.join(ri.RI, rel + ".gz"), "rb").read()
        assert hashlib.sha256(raw).hexdigest() == meta["sha256"], rel
        ref = os.path.join("/root/reference", rel)
        if os.path.exists(ref):                      # CPU container only: the copy IS the reference's file
            assert open(ref, "rb").read() == raw, rel


def test_readers_mesh_obj_dmat(tmp_path):
    V, T, F = sceneio.read_mesh(os.path.join(ri.RI, "models/mesh/beam_20k.mesh"))
    assert V.shape == (20000, 3) and T.shape == (100793, 4) and T.min() == 0 and T.max() == 19999
    V2, T2, _ = sceneio.read_mesh(os.path.join(ri.RI, "models/mesh/two_tets.mesh"))
    assert V2.shape == (5, 3) and T2.shape == (2, 4)
    Vo, Fo = sceneio.read_obj(os.path.join(ri.RI, "models/obj/square_xz_16k.obj"))     # 'f a/b/c' records
    assert Vo.shape == (8321, 3) and Fo.shape == (16384, 3) and Fo.max() == 8320
    x = sceneio.read_dmat(os.path.join(ri.RI, "scenes/data/boat_x_0009.dmat"))
    assert x.shape == (3674, 3)
    ids = sceneio.read_dmat(os.path.join(ri.RI, "models/mesh/gummy_bear/gummy_bear_material_ids.dmat"))
    assert ids.shape == (24393, 1) and set(np.unique(ids)) == {0.0, 1.0}
    # round trips through the writers (the driver's state in/out format, SURVEY.md section 5 checkpoint row)
    p = str(tmp_path / "m.dmat")
    sceneio.write_dmat(p, x[:17])
    assert np.array_equal(sceneio.read_dmat(p), x[:17])
    pm = str(tmp_path / "m.mesh")
    sceneio.write_mesh(pm, V2, T2)
    V3, T3, _ = sceneio.read_mesh(pm)
    assert np.array_equal(V3, V2) and np.array_equal(T3, T2)
    # binary DMAT body (readDMAT.cpp:111-128)
    pb = str(tmp_path / "b.dmat")
    A = np.arange(6, dtype=np.float64).reshape(2, 3)
    with open(pb, "wb") as fh:
        fh.write(b"0 0\n3 2\n" + np.ascontiguousarray(A.T).tobytes())
    assert np.array_equal(sceneio.read_dmat(pb), A)


def test_scene_trampoline_C1():
    """scenes/2D/trampoline.json: two objects, two Stable-Neohookean materials, densities 1e5 / 1e3, hang BC on the
    second object only (the LAST vertex of each end group, fixed_boundary_conditions.h:70-81)"""
    sc = ri.scene("scenes/2D/trampoline.json", 2)
    assert sc.V.shape == (181, 2) and sc.T.shape == (304, 3)
    assert sc.object_vertex_offsets.tolist() == [0, 145, 181]
    assert [m[0] for m in sc.materials] == ["Stable-Neohookean"] * 2
    mu0, la0 = meshgen.enu_to_lame(79e5, 0.42)
    assert sc.materials[0][1:] == (mu0, la0)
    assert np.array_equal(np.unique(sc.mat_id[:256]), [0]) and np.array_equal(np.unique(sc.mat_id[256:]), [1])
    assert np.all(sc.density_e[:256] == 1e5) and np.all(sc.density_e[256:] == 1e3)
    assert sc.is_fixed.sum() == 2 and np.all(np.nonzero(sc.is_fixed)[0] >= 145)
    box = sc.V[145:]
    g = sceneio.boundary_groups(box, 0.1, 0)
    assert np.nonzero(sc.is_fixed)[0].tolist() == sorted([145 + g[0][-1], 145 + g[1][-1]])
    assert sc.free_map.max() == 178 and (sc.free_map == -1).sum() == 2
    assert sc.h == 0.0333 and sc.outer_steps == 10 and sc.ls_iters == 20
    assert sc.body_force[1] == float(np.float32(-9.8))            # body_force passes through std::vector<float>
    assert np.all(sc.V[:145, 1] >= 4.0)                             # offset (0,5,0) on the first object


def test_scene_3d_normalize_offset_initial_state_material_ids():
    sc = ri.scene("scenes/3D/gummy_bear_deformed.json", 3)
    assert sc.V.shape == (5701, 3) and sc.T.shape == (24393, 4)
    assert abs((sc.V[:, 1] - 5.0).max() - 1.0) < 1.0 and sc.V[:, 0].max() <= 1.0 + 1e-15   # V /= V.maxCoeff(), then offset
    assert np.bincount(sc.mat_id).tolist() == [22400, 1993]
    assert np.all(sc.density_e == 1e3)
    assert sc.x0.shape == (3 * 5701,) and not np.allclose(sc.x0, sc.V.reshape(-1))
    J = np.linalg.det(meshgen.deformation_gradients(sc.V, sc.T, sc.x0))
    assert 0.5 < J.min() and J.max() < 1.8 and abs(np.median(J) - 1.0) < 1e-3
    # hangends: the whole high-x group
    g = sceneio.boundary_groups(sc.V, 0.1, 0)
    assert np.array_equal(np.nonzero(sc.is_fixed)[0], g[1])
    sc2 = ri.scene("scenes/3D/collision_drop_bunny.json", 3)        # two objects concatenated (meshes.cpp:31-50)
    assert sc2.V.shape[0] == 713 and sc2.T.min() == 0 and sc2.T.max() == 712
    assert sc2.object_vertex_offsets[1] == 14 and np.all(sc2.T[sc2.object_vertex_offsets[1] <= sc2.T.min(axis=1)].min() >= 14)


def test_mkt_dump_mass_matrix_bit_exact():
    st = ri.mkt_state()
    V, T, fm = st["V"], st["T"], st["fm"]
    cpu = orc.OracleMixedStretch(V, T, fm, st["materials"])
    cpu.reset()
    vols = cpu.get("vols")
    assert np.all(vols == 0.03125)                                  # igl::doublearea, NOT the area (tri2d_mesh.cpp:32-34)
    M = mass_matrix(V, T, vols, st["density"])
    free = np.nonzero(fm >= 0)[0]
    fd = (2 * free[:, None] + np.arange(2)[None, :]).reshape(-1)
    PMP = M[fd][:, fd]
    Mref = ri.mkt("lhs_M")
    assert abs(PMP - Mref).max() == 0.0
    # the reference's pattern carries the explicit x-y zeros of the dense element blocks (assemble.cpp:68-74)
    assert Mref.nnz == 2 * PMP.nnz


def test_oracle_reproduces_the_reference_dump():
    st = ri.mkt_state()
    assert st["fit_residuals"].max() < 1e-12
    cpu = orc.OracleMixedStretch(st["V"], st["T"], st["fm"], st["materials"])
    cpu.reset()
    cpu.set("s", st["s"])
    cpu.update(st["x"], st["h"])
    errs = ri.check_against_mkt_dump(cpu, st)
    print("oracle vs reference dump:", errs)


def test_dumped_solution_is_only_an_approximate_solve():
    """SURVEY.md section 8c(5) hoped (x,s,la)_sub solves the dumped KKT system; it does to 5 % only (an iterative /
    subspace solve) - recorded here so nobody leans on it: the matrices and right-hand sides above are the fixture."""
    M, K, Gx, Gs = ri.mkt("lhs_M"), ri.mkt("lhs_K"), ri.mkt("lhs_Gx"), ri.mkt("lhs_Gs")
    b = np.concatenate([ri.mkt(n).toarray().ravel() for n in ("rhs_x", "rhs_gs", "rhs_gl")])
    sol = np.concatenate([ri.mkt(n).toarray().ravel() for n in ("x_sub", "s_sub", "la_sub")])
    A = sp.bmat([[M, None, Gx], [None, K, Gs], [Gx.T, Gs, None]]).tocsr()
    rel = np.linalg.norm(A @ sol - b) / np.linalg.norm(b)
    assert 1e-3 < rel < 0.1
    # rhs_x = -P M (x - x~ - h^2 f) needs x~ (not recorded, and the stretch script moves the pinned ends): not a fixture
