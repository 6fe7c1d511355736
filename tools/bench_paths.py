#!/usr/bin/env python
"""Device timings of the §8(f) paths at config C3 size (55^3 x 6 = 998 250 tets): the fused KKT product, the matrix-free
dual-condensed product and the non-mixed Stretch update, CUDA events on torch's current stream (= the context stream),
inputs resident.  Prints one JSON line.  Not part of bench.py's contract; numbers go to DESIGN.md / profiles/."""
import ctypes as C
import json
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import mixed_fem_b200 as mf  # noqa: E402
from mixed_fem_b200 import meshgen  # noqa: E402


def timed(fn, reps=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 55
    V, T = meshgen.kuhn_grid(n, n, n)
    fm = meshgen.pinned_free_map(V)
    mu, la = meshgen.enu_to_lame(1e6, 0.45)
    ms = mf.MixedStretch(V, T, fm, [("Stable-Neohookean", mu, la)])
    ms.set_stream(torch.cuda.current_stream().cuda_stream)
    ms.reset()
    h = 0.0333
    ms.newton_init(density=1000.0, h=h)
    x = meshgen.deformed_positions(V, 1.0 / n)
    xd = torch.from_numpy(x).cuda()
    ms.update_dev(xd.data_ptr(), h)
    L, hnd = ms.L, ms.h
    ox, os_, ola, tot = C.c_int64(), C.c_int64(), C.c_int64(), C.c_int64()
    ms._check(L.mfb_kkt_layout_dev(hnd, C.byref(ox), C.byref(os_), C.byref(ola), C.byref(tot)))
    # random KKT vector with zero padding: fill through the host entry point's layout is not needed for timing; zeros + noise
    # on the x part keep the padding rule (padding of ld must be zero)
    v = torch.zeros(tot.value, dtype=torch.float64, device="cuda")
    v[: 3 * ms.nfree] = torch.randn(3 * ms.nfree, dtype=torch.float64, device="cuda")
    out = torch.zeros_like(v)
    t_kkt = timed(lambda: ms._check(L.mfb_kkt_product_dev(hnd, C.c_void_p(v.data_ptr()), C.c_void_p(out.data_ptr()))))
    ne = (tot.value - os_.value) // 2
    lav = torch.zeros(ne, dtype=torch.float64, device="cuda")
    lo = torch.zeros_like(lav)
    t_dual = timed(lambda: ms._check(L.mfb_dual_product_dev(hnd, C.c_void_p(lav.data_ptr()), C.c_void_p(lo.data_ptr()))))
    t_str = timed(lambda: ms._check(L.mfb_stretch_update_dev(hnd, C.c_void_p(xd.data_ptr()), h)), reps=5, warm=1)
    t_upd = timed(lambda: ms.update_dev(xd.data_ptr(), h))
    nE = ms.nE
    print(json.dumps({"workload": "%d^3 x 6 = %d tets, Stable-Neohookean" % (n, nE), "ms_kkt_product": t_kkt, "ms_dual_product": t_dual,
                      "ms_stretch_update": t_str, "ms_mixed_update": t_upd, "kkt_products_tets_per_s": nE / t_kkt * 1e3,
                      "stretch_updates_tets_per_s": nE / t_str * 1e3}))
    ms.close()


if __name__ == "__main__":
    main()
