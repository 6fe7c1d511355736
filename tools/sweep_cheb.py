"""C3 (998 250 tets, ARAP): CG iterations and ms per Newton iteration of one time step for eigen-pcg-diag, block-Jacobi
and the Chebyshev preconditioner over a sweep of (degree, ratio)."""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mixed_fem_b200 as mf
from mixed_fem_b200 import meshgen

n = int(sys.argv[1]) if len(sys.argv) > 1 else 55
V, T = meshgen.kuhn_grid(n, n, n)
fm = meshgen.pinned_free_map(V)
mu, la = meshgen.enu_to_lame(1e6, 0.45)
ms = mf.MixedStretch(V, T, fm, [("ARAP", mu, la)])
cases = [(1, 0, 0), (2, 0, 0)] + [(3, d, r) for d in (3, 4, 6, 8) for r in (30, 60, 120)]
for precond, deg, ratio in cases:
    ms.reset()
    ms.newton_init(density=1000.0, h=0.0333)
    if precond == 3:
        ms.set_option("cheb_degree", deg); ms.set_option("cheb_ratio", ratio)
    ms.newton_step(outer_steps=1, precond=precond)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    lg = ms.newton_step(outer_steps=3, precond=precond, cg_tol=1e-4, cg_max_iters=500)
    torch.cuda.synchronize()
    el = (time.perf_counter() - t0) * 1e3 / max(lg["n_iters"], 1)
    print("precond %d degree %d ratio %3d : cg %s  spmv-equivalents %d  %.2f ms / Newton iteration  energy %.10g" %
          (precond, deg, ratio, lg["cg_iters"].tolist(), int(lg["cg_iters"].sum()) * max(deg, 1), el, lg["energy"][-1]), flush=True)
ms.close()
