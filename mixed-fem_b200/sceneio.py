"""Readers for the reference's own input files and the part of its scene loader that feeds the mixed-stretch path.

What each function replaces in evouga/mixed-fem (host-side, numpy only; nothing here touches the device):

  read_mesh ............ igl::readMESH as called by SimState<3>::load_mesh (src/simulation_state.cpp:62-73)
  read_obj ............. igl::read_triangle_mesh on a Wavefront .obj (src/simulation_state.cpp:49-60)
  read_dmat/write_dmat . igl::readDMAT / writeDMAT (deps/libigl/include/igl/readDMAT.cpp: "cols rows", column-major)
  read_mkt ............. Eigen::loadMarket counterpart for the dumps of scripts/mats/*.mkt
  remove_unreferenced .. igl::remove_unreferenced (2D objects only, src/simulation_state.cpp:52-54)
  boundary_groups ...... BoundaryCondition::init_boundary_groups (src/boundary_conditions/boundary_condition.cpp:5-21)
  fixed_vertices ....... is_fixed_ of NullBC / OnePointBC / HangBC / HangEndsBC and of the scripted BCs at t = 0
                         (src/boundary_conditions/fixed_boundary_conditions.h:9-97, bend/stretch/twist/moving *.h)
  load_scene ........... SimState<DIM>::load (src/simulation_state.cpp:95-405) + Meshes::Meshes / init_bcs
                         (src/mesh/meshes.cpp:11-55,151-186): objects concatenated, per-object material / density / BC

Files ending in .gz are opened through gzip (the copies under tests/golden/ref_inputs are stored compressed).
"""
import gzip
import json
import os

import numpy as np

from .meshgen import enu_to_lame, face_normals


def _open_text(path):
    if str(path).endswith(".gz"):
        return gzip.open(path, "rt")
    return open(path, "r")


def _resolve(path):
    """path, or path + '.gz' when only the compressed copy exists"""
    if os.path.exists(path):
        return path
    if os.path.exists(path + ".gz"):
        return path + ".gz"
    raise FileNotFoundError(path)


# ---------------------------------------------------------------------------------------------------------------
# MEDIT .mesh (ASCII): Vertices / Triangles / Tetrahedra sections, 1-based ids, one trailing reference per record
# ---------------------------------------------------------------------------------------------------------------
def read_mesh(path):
    """(V (nV,3) float64, T (nT,4) int32, F (nF,3) int32), 0-based, in file order (igl::readMESH keeps both)."""
    with _open_text(_resolve(path)) as fh:
        tok = fh.read().split()
    V = np.zeros((0, 3)); T = np.zeros((0, 4), dtype=np.int32); F = np.zeros((0, 3), dtype=np.int32)
    i, n = 0, len(tok)
    dim = 3
    while i < n:
        key = tok[i].lower()
        i += 1
        if key == "meshversionformatted":
            i += 1
        elif key == "dimension":
            dim = int(tok[i]); i += 1
        elif key == "vertices":
            cnt = int(tok[i]); i += 1
            a = np.array(tok[i:i + cnt * (dim + 1)], dtype=np.float64).reshape(cnt, dim + 1)
            V = np.ascontiguousarray(a[:, :dim]); i += cnt * (dim + 1)
        elif key == "triangles":
            cnt = int(tok[i]); i += 1
            a = np.array(tok[i:i + cnt * 4], dtype=np.int64).reshape(cnt, 4)
            F = np.ascontiguousarray(a[:, :3] - 1).astype(np.int32); i += cnt * 4
        elif key == "tetrahedra":
            cnt = int(tok[i]); i += 1
            a = np.array(tok[i:i + cnt * 5], dtype=np.int64).reshape(cnt, 5)
            T = np.ascontiguousarray(a[:, :4] - 1).astype(np.int32); i += cnt * 5
        elif key in ("edges", "quadrilaterals", "hexahedra", "corners", "requiredvertices", "ridges", "requirededges"):
            width = {"edges": 3, "quadrilaterals": 5, "hexahedra": 9, "corners": 1, "requiredvertices": 1, "ridges": 1,
                     "requirededges": 1}[key]
            cnt = int(tok[i]); i += 1 + cnt * width
        elif key == "end":
            break
        elif key.startswith("#"):
            continue
        else:
            raise ValueError("%s: unknown MEDIT section %r" % (path, tok[i - 1]))
    if V.shape[0] == 0:
        raise ValueError("%s: no Vertices section" % path)
    return V, T, F


def write_mesh(path, V, T):
    with open(path, "w") as fh:
        fh.write("MeshVersionFormatted 1\nDimension 3\nVertices\n%d\n" % V.shape[0])
        for v in V:
            fh.write("%.17g %.17g %.17g 0\n" % tuple(v))
        fh.write("Triangles\n0\nTetrahedra\n%d\n" % T.shape[0])
        for t in T:
            fh.write("%d %d %d %d 0\n" % tuple(int(q) + 1 for q in t))
        fh.write("End\n")


# ---------------------------------------------------------------------------------------------------------------
# Wavefront .obj: positions and faces (v, f with v/vt/vn index forms, negative indices, polygons fan-triangulated)
# ---------------------------------------------------------------------------------------------------------------
def read_obj(path):
    """(V (nV,3) float64, F (nF,3) int32); texture / normal records are ignored like read_triangle_mesh does."""
    vs, fs = [], []
    with _open_text(_resolve(path)) as fh:
        for line in fh:
            if line.startswith("v "):
                p = line.split()
                vs.append((float(p[1]), float(p[2]), float(p[3]) if len(p) > 3 else 0.0))
            elif line.startswith("f "):
                idx = []
                for w in line.split()[1:]:
                    k = int(w.split("/")[0])
                    idx.append(k - 1 if k > 0 else len(vs) + k)
                for q in range(1, len(idx) - 1):
                    fs.append((idx[0], idx[q], idx[q + 1]))
    return np.array(vs, dtype=np.float64).reshape(-1, 3), np.array(fs, dtype=np.int32).reshape(-1, 3)


def remove_unreferenced(V, T):
    """igl::remove_unreferenced: keep referenced vertices in their original order, renumber T."""
    used = np.zeros(V.shape[0], dtype=bool)
    used[T.ravel()] = True
    new_id = np.cumsum(used) - 1
    return np.ascontiguousarray(V[used]), new_id[T].astype(np.int32)


# ---------------------------------------------------------------------------------------------------------------
# DMAT: "cols rows" then the values column-major; an all-zero ASCII header followed by a second header = binary body
# ---------------------------------------------------------------------------------------------------------------
def read_dmat(path):
    p = _resolve(path)
    opener = gzip.open if p.endswith(".gz") else open
    with opener(p, "rb") as fh:
        raw = fh.read()
    nl = raw.index(b"\n")
    cols, rows = (int(t) for t in raw[:nl].split())
    if rows * cols > 0:
        vals = np.array(raw[nl + 1:].split()[: rows * cols], dtype=np.float64)
        if vals.size != rows * cols:
            raise ValueError("%s: expected %d values, found %d" % (path, rows * cols, vals.size))
        return np.ascontiguousarray(vals.reshape(cols, rows).T)
    rest = raw[nl + 1:]
    if not rest.strip():
        return np.zeros((rows, cols))
    nl2 = rest.index(b"\n")
    cols, rows = (int(t) for t in rest[:nl2].split())
    body = np.frombuffer(rest[nl2 + 1: nl2 + 1 + 8 * rows * cols], dtype=np.float64)
    return np.ascontiguousarray(body.reshape(cols, rows).T)


def write_dmat(path, M):
    M = np.atleast_2d(np.asarray(M, dtype=np.float64))
    with open(path, "w") as fh:
        fh.write("%d %d\n" % (M.shape[1], M.shape[0]))
        for v in M.T.ravel():
            fh.write("%.17g\n" % v)


def read_mkt(path):
    """MatrixMarket file -> dense ndarray ('array' format: rows x cols, column-major) or (rows, cols, data, shape)
    triplets ('coordinate' format, 0-based; symmetric files are NOT expanded - Eigen::saveMarket writes general ones)."""
    with _open_text(_resolve(path)) as fh:
        head = fh.readline().lower().split()
        if head[0] != "%%matrixmarket":
            raise ValueError("%s: not a MatrixMarket file" % path)
        fmt = head[2]
        line = fh.readline()
        while line.startswith("%"):
            line = fh.readline()
        dims = [int(t) for t in line.split()]
        body = np.array(fh.read().split(), dtype=np.float64)
    if fmt == "array":
        r, c = dims[0], (dims[1] if len(dims) > 1 else 1)
        return np.ascontiguousarray(body.reshape(c, r).T)
    r, c, nnz = dims
    body = body.reshape(nnz, -1)
    return body[:, 0].astype(np.int64) - 1, body[:, 1].astype(np.int64) - 1, body[:, 2].copy(), (r, c)


# ---------------------------------------------------------------------------------------------------------------
# boundary conditions: which vertices are pinned
# ---------------------------------------------------------------------------------------------------------------
def boundary_groups(V, ratio=0.1, axis=0):
    """the two vertex groups at the low / high end of `axis` (boundary_condition.cpp:5-21), ascending vertex order"""
    lo, hi = V.min(axis=0), V.max(axis=0)
    rng = hi - lo
    c = V[:, axis]
    g0 = np.nonzero(c < lo[axis] + rng[axis] * ratio)[0]
    g1 = np.nonzero((c > hi[axis] - rng[axis] * ratio) & ~(c < lo[axis] + rng[axis] * ratio))[0]
    return [g0, g1]


def fixed_vertices(V, bc_type="null", ratio=0.1, axis=0):
    """is_fixed_ (nV,) int32 of the named BC on REFERENCE vertices V (the BC is built on Vref_, mesh.cpp:90-92)."""
    fixed = np.zeros(V.shape[0], dtype=np.int32)
    t = bc_type.lower()
    if t in ("null", "scale", "randomize"):
        return fixed
    if t == "onepoint":
        fixed[0] = 1
        return fixed
    g = boundary_groups(V, ratio, axis)
    if t == "hang":                       # the LAST vertex of each end group (fixed_boundary_conditions.h:70-81)
        for grp in g:
            fixed[grp[-1]] = 1
    elif t == "hangends":                 # group_id_ = 1: the high end (fixed_boundary_conditions.h:84-97)
        fixed[g[1]] = 1
    elif t in ("bend", "stretch", "twist", "twist-and-stretch", "twistandstretch"):
        fixed[g[0]] = 1; fixed[g[1]] = 1
    elif t in ("translate", "moving"):    # MovingBC group_id_ (moving_boundary_condition.h:20-29)
        fixed[g[1]] = 1
    else:
        raise ValueError("unknown boundary condition %r" % bc_type)
    return fixed


def free_map_from_fixed(is_fixed):
    """Meshes::init_bcs / Mesh::init_bcs (meshes.cpp:175-185): ascending free ids, -1 where pinned"""
    fm = np.full(is_fixed.shape[0], -1, dtype=np.int32)
    free = is_fixed == 0
    fm[free] = np.arange(int(free.sum()), dtype=np.int32)
    return fm


# ---------------------------------------------------------------------------------------------------------------
# scene JSON
# ---------------------------------------------------------------------------------------------------------------
class Scene:
    """What SimState<DIM>::load leaves behind for the mixed-stretch path (concatenated over objects):
    V (= Vref_), T, free_map, mat_id + materials [(name, mu, la)], density_e (per ELEMENT: the density of the material
    its OBJECT was created with - a per-tet material_ids file does not change it, tet_mesh.cpp:36-40), x0 / v0
    (initial_state or vec(V), zeros), object_vertex_offsets, and the SimConfig scalars."""
    pass


def _material_list(args):
    mats = []
    for m in args.get("material_models", [{}]):
        ym = float(m.get("youngs_modulus", 1e6)); pr = float(m.get("poissons_ratio", 0.45))
        mu, la = enu_to_lame(ym, pr)
        mats.append({"name": m.get("energy", "Stable-Neohookean"), "mu": mu, "la": la, "density": float(m.get("density", 1000.0)),
                     "ym": ym, "pr": pr})
    return mats


def load_scene(json_path, dim, root=None, shell=False, thickness=1e-3):
    """dim = 2 (sim2d: objects are .obj, z dropped) or 3 (sim3d: .mesh).  `root`: directory the scene's relative paths
    ('../models/...') are resolved against (default: the nearest ancestor of the json file that has a sibling
    'models' directory, i.e. what running from <repo>/build resolves to).
    shell=True loads 3D triangle objects (.obj) as TriMesh shells (apps/cloth_sim.cpp pattern; no scene file of the
    reference does this, so only the mesh part of the loader applies)."""
    with _open_text(_resolve(json_path)) as fh:
        args = json.load(fh)
    if root is None:      # the scenes are written for a run from <repo>/build: find the ancestor where '../models' exists
        root = os.path.dirname(os.path.abspath(json_path))
        probe = root
        for _ in range(4):
            if os.path.isdir(os.path.join(probe, "..", "models")):
                root = probe
                break
            probe = os.path.dirname(probe)
    sc = Scene()
    sc.args = args
    sc.dim = dim
    # load_params (simulation_state.cpp:380-405); body_force goes through std::vector<float>
    sc.h = float(args.get("dt", 0.034))
    sc.body_force = np.array([float(np.float32(v)) for v in args.get("body_force", [0.0, -9.8, 0.0])])[:dim if not shell else 3]
    sc.outer_steps = int(args.get("max_newton_iterations", 5))
    sc.ls_iters = int(args.get("max_linesearch_iterations", 20))
    sc.cg_max_iters = int(args.get("max_iterative_solver_iters", 500))
    sc.cg_tol = float(args.get("iterative_solver_tolerance", 1e-4))
    sc.enable_ccd = bool(args.get("enable_ccd", False))
    sc.linear_solver = args.get("linear_solver", "eigen-llt")
    sc.mixed_variables = list(args.get("mixed_variables", ["mixed-stretch", "mixed-collision"]))
    sc.variables = list(args.get("variables", []))
    mats = _material_list(args)
    sc.materials = [(m["name"], m["mu"], m["la"]) for m in mats]
    sc.material_configs = mats
    Vs, Ts, mids, dens, fixed, offs, v0s = [], [], [], [], [], [0], []
    for obj in args.get("objects", []):
        path = os.path.normpath(os.path.join(root, obj["path"]))
        idx = int(obj.get("material_index", 0))
        if dim == 2 or shell:
            V, T = read_obj(path)
            V, T = remove_unreferenced(V, T)
            if not shell:
                V = np.ascontiguousarray(V[:, :2])
        else:
            V, T, _ = read_mesh(path)
            if obj.get("normalize", True):
                V = V / V.max()
        off = np.asarray(obj.get("offset", [0.0, 0.0, 0.0]), dtype=np.float64)
        V = V + off[None, :V.shape[1]]
        if "transformation" in obj:
            d = V.shape[1]
            A = np.asarray(obj["transformation"], dtype=np.float64).reshape(d, d).T   # Map<Matrix> is column-major
            c = V.sum(axis=0) / V.shape[0]
            V = (V - c) @ A.T + c
        if "material_ids" in obj and dim == 3 and not shell:
            mid = read_dmat(os.path.normpath(os.path.join(root, obj["material_ids"]))).reshape(-1).astype(np.int32)
            rho = mats[0]["density"]       # Mesh(V,T,ids,materials): material_ = materials[0] (mesh.cpp:17-19)
        else:
            mid = np.full(T.shape[0], idx, dtype=np.int32)
            rho = mats[idx]["density"]
        bc = obj.get("boundary_condition", {})
        fx = fixed_vertices(V, bc.get("type", "null"), float(bc.get("ratio", 0.1)), int(bc.get("axis", 0)))
        v0 = np.zeros_like(V)
        if "initial_velocity" in obj:
            v0[:] = np.asarray(obj["initial_velocity"], dtype=np.float64)[None, :V.shape[1]]
        Vs.append(V); Ts.append(T + offs[-1]); mids.append(mid); dens.append(np.full(T.shape[0], rho)); fixed.append(fx); v0s.append(v0)
        offs.append(offs[-1] + V.shape[0])
    sc.V = np.ascontiguousarray(np.concatenate(Vs, axis=0))
    sc.T = np.ascontiguousarray(np.concatenate(Ts, axis=0)).astype(np.int32)
    sc.mat_id = np.concatenate(mids).astype(np.int32)
    sc.density_e = np.concatenate(dens)
    sc.is_fixed = np.concatenate(fixed).astype(np.int32)
    sc.free_map = free_map_from_fixed(sc.is_fixed)
    sc.object_vertex_offsets = np.array(offs)
    sc.normals = face_normals(sc.V, sc.T) if shell else None
    sc.thickness = thickness
    sc.x0 = sc.V.reshape(-1).copy()
    sc.v0 = np.concatenate(v0s, axis=0).reshape(-1)
    init = args.get("initial_state", {})
    if "x_path" in init:
        x = read_dmat(os.path.normpath(os.path.join(root, init["x_path"])))
        if x.shape != sc.V.shape:
            raise ValueError("initial_state x_path has shape %s, mesh has %s" % (x.shape, sc.V.shape))
        sc.x0 = np.ascontiguousarray(x).reshape(-1)
    if "v_path" in init:
        vpath = os.path.normpath(os.path.join(root, init["v_path"]))
        try:
            v = read_dmat(vpath)
            sc.v0 = np.ascontiguousarray(v).reshape(-1)
        except FileNotFoundError:
            pass
    return sc
