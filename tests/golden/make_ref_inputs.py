"""Copies the reference's own INPUT DATA files (meshes, saved states, scene descriptions, the one numeric matrix dump)
that the parity tests run on into tests/golden/ref_inputs/, gzip-compressed, with a manifest of SHA-256 sums of the
uncompressed bytes.  These are data, not source: no reference code is copied.  /root/reference does not exist on the
GPU box, so the tests read these copies (mixed_fem_b200.sceneio opens *.gz transparently).

Run here (CPU container):  python tests/golden/make_ref_inputs.py
"""
import gzip
import hashlib
import json
import os

REF = "/root/reference"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_inputs")
FILES = [
    # config C1 (scenes/2D/trampoline.json) and the deformed 2D states that share its two meshes
    "scenes/2D/trampoline.json", "scenes/2D/diving_board_deformed.json",
    "models/obj/square_xy_145.obj", "models/obj/box_wide_145_thin_xy.obj",
    "scenes/2D/data/diving_board/sim_x0_0001.dmat", "scenes/2D/data/diving_board/sim_v_0001.dmat",
    "scenes/2D/data/diving_board/sim_x0_0054.dmat", "scenes/2D/data/diving_board/sim_v_0054.dmat",
    # config C2 as named
    "models/mesh/beam_20k.mesh",
    # config C5
    "models/obj/square_xz_16k.obj",
    # deformed 3D states paired with their meshes
    "scenes/3D/boat.json", "models/mesh/boat.mesh", "scenes/data/boat_x_0009.dmat", "scenes/data/boat_v_0009.dmat",
    "models/mesh/coarse_bunny.mesh", "scenes/data/bunny_x_0002.dmat", "scenes/data/bunny_x_0027.dmat", "scenes/data/bunny_x_0031.dmat",
    "scenes/3D/collision_drop_bunny.json", "models/mesh/cube_14.mesh", "scenes/3D/data/sim_x0_0016.dmat", "scenes/3D/data/sim_v_0016.dmat",
    "scenes/3D/gummy_bear_deformed.json", "models/mesh/gummy_bear/gummy_bear.mesh", "models/mesh/gummy_bear/gummy_bear_material_ids.dmat",
    "models/mesh/gummy_bear/sim_x0_0013.dmat", "models/mesh/gummy_bear/sim_v_0013.dmat",
    # tiny meshes of the reference's (disabled) unit tests
    "models/mesh/tet.mesh", "models/mesh/two_tets.mesh",
    # the one numeric artefact the reference holds: a 2D KKT dump (scripts/subspace_solver.m reads it)
    "scripts/mats/lhs_M.mkt", "scripts/mats/lhs_K.mkt", "scripts/mats/lhs_Gx.mkt", "scripts/mats/lhs_Gs.mkt",
    "scripts/mats/rhs_x.mkt", "scripts/mats/rhs_gs.mkt", "scripts/mats/rhs_gl.mkt",
    "scripts/mats/x_sub.mkt", "scripts/mats/s_sub.mkt", "scripts/mats/la_sub.mkt",
]


def main():
    manifest = {}
    for rel in FILES:
        src = os.path.join(REF, rel)
        if not os.path.exists(src):
            print("missing in the reference tree, skipped:", rel)
            continue
        raw = open(src, "rb").read()
        dst = os.path.join(OUT, rel + ".gz")
        os.makedirs(os.path.dirname(dst), exist_ok=True)
        with open(dst, "wb") as fh:
            with gzip.GzipFile(filename="", mode="wb", fileobj=fh, mtime=0, compresslevel=9) as gz:
                gz.write(raw)
        manifest[rel] = {"sha256": hashlib.sha256(raw).hexdigest(), "bytes": len(raw)}
    with open(os.path.join(OUT, "MANIFEST.json"), "w") as fh:
        json.dump({"reference_commit": "489dab3", "files": manifest}, fh, indent=1, sort_keys=True)
    print("%d files, %.1f MB compressed" % (len(manifest), sum(os.path.getsize(os.path.join(OUT, r + ".gz")) for r in manifest) / 1e6))


if __name__ == "__main__":
    main()
