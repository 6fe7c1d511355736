#!/usr/bin/env python
"""Headless twin of the reference's apps/convergence.cpp (:53-127) for the mixed-stretch variable on the GPU:
    python tools/convergence.py scene.json -n N [--root DIR] [--out output/convergence.dmat] [--dim 2|3]
loads the scene (mixed_fem_b200.sceneio: objects, materials, per-object densities, BCs, initial_state), runs N time steps of
the device-resident Newton optimizer (mfb_newton_step) and records, at the top of every Newton iteration, the gradient norm
|x.rhs() + Stretch.rhs()|_2 of the NON-mixed system at the mixed iterate (option convergence_log); a step that stops early
repeats its last value up to max_newton_iterations (convergence.cpp:111-117).  Output: the N x M matrix as an ASCII DMAT.
Contact (mixed-collision, CCD) is out of scope: scenes are run with the elastic variable only."""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mixed_fem_b200 as mf                    # noqa: E402
from mixed_fem_b200 import sceneio             # noqa: E402


def run(scene_path, nsteps, root=None, dim=2, precond=2, cg_tol=1e-12, cg_max_iters=20000):
    sc = sceneio.load_scene(scene_path, dim, root=root)
    M = sc.outer_steps
    ms = mf.MixedStretch(sc.V, sc.T, sc.free_map, sc.materials, mat_id=sc.mat_id)
    ms.reset(sc.x0)
    ms.newton_init(density=sc.density_e, h=sc.h, force=sc.body_force, x0=sc.x0, v0=sc.v0)
    ms.set_option("convergence_log", 1)
    G = np.zeros((nsteps, M))
    for i in range(nsteps):
        lg = ms.newton_step(outer_steps=M, ls_iters=sc.ls_iters, precond=precond, cg_tol=cg_tol, cg_max_iters=cg_max_iters)
        g = list(lg["fem_gradient_norm"][:M])
        g += [g[-1]] * (M - len(g))
        G[i] = g
    ms.close()
    return G


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("json")
    ap.add_argument("-n", type=int, default=1, help="number of time steps")
    ap.add_argument("--root", default=None, help="directory the scene's relative paths resolve against")
    ap.add_argument("--dim", type=int, default=2, help="apps/convergence.cpp is compiled for DIM = 2")
    ap.add_argument("--out", default=os.path.join("output", "convergence.dmat"))
    a = ap.parse_args()
    G = run(a.json, a.n, root=a.root, dim=a.dim)
    print("gradients:\n", G)
    os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
    sceneio.write_dmat(a.out, G)


if __name__ == "__main__":
    main()
