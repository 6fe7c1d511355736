"""Single-GPU timing of update() on ONE part of a 2-way partition of the C4 grid: library partitioner (interface elements
first) against the numpy slab partition (ascending elements) against the standalone half grid."""
import sys, os, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mixed_fem_b200 as mf
from mixed_fem_b200 import meshgen, partition

def run(V, T, fm, x, tag):
    mu, la = meshgen.enu_to_lame(1e6, 0.45)
    ms = mf.MixedStretch(V, T, fm, [("Stable-Neohookean", mu, la)])
    ms.reset(x[1])
    xs = [torch.as_tensor(a, device="cuda") for a in x]
    for i in range(3):
        ms.update_dev(xs[i % 2].data_ptr(), 0.0333)
    torch.cuda.synchronize()
    ms.set_option("profile", 1); ms.kernel_time_ms()
    for i in range(10):
        ms.update_dev(xs[(i + 1) % 2].data_ptr(), 0.0333)
    torch.cuda.synchronize()
    kt, kn = ms.kernel_time_ms()
    print(tag, "nE", T.shape[0], "kernel ms", kt / kn, flush=True)
    ms.close()

n = 119
V, T = meshgen.kuhn_grid(n, n, n, h=1.0 / 119)
fm = meshgen.pinned_free_map(V)
xa = meshgen.deformed_positions(V, 1.0 / 119, scale=1.0); xb = meshgen.deformed_positions(V, 1.0 / 119, scale=0.5)
NP = int(sys.argv[1]) if len(sys.argv) > 1 else 2
for rank in range(NP):
    c, _, _ = partition.make_local(V, T, fm, NP, rank)
    run(c.V, c.T, c.free_map, [c.restrict_vertex_vector(xa, 3), c.restrict_vertex_vector(xb, 3)], "library partition rank %d (n_if_elems %d)" % (rank, c.n_if_elems))
    a, _, _ = partition.make_local_numpy(V, T, fm, NP, rank)
    run(a.V, a.T, a.free_map, [a.restrict_vertex_vector(xa, 3), a.restrict_vertex_vector(xb, 3)], "numpy partition rank %d (ascending elements)" % rank)
