"""GPU: mfb_energy_noalloc_dev - the allocation-free variant of mfb_energy_dev (no counterpart in the reference; added
for callers that cannot call cudaMalloc) - through the C ABI against the existing
entry point, on two calls in a row and across two contexts that share the library's static scratch array (a larger mesh
first, so the second call finds stale partial sums beyond its own range).  Collected last on purpose
(file name)."""
import ctypes as C

import numpy as np
import pytest

import mixed_fem_b200 as mf
from mixed_fem_b200 import meshgen
from tests.common import DT, LA, MU

pytestmark = pytest.mark.gpu
TOL = 1e-12        # energies are compared at 1e-12 relative elsewhere in the suite


def _energy_both(ms, s_ptr, e_dev):
    """(existing entry point, new entry point) for the stretches at s_ptr (None: the context's s_)"""
    ref = ms.energy_dev(s_ptr)
    ms.energy_noalloc_dev(e_dev.data_ptr(), s_ptr)
    return ref, float(e_dev.cpu()[0])


def test_energy_noalloc_matches_energy_dev_on_consecutive_calls(built):
    import torch
    L = mf.lib()
    assert L.mfb_noalloc_max_elements() == 131072 * 128
    e_dev = torch.full((1,), float("nan"), dtype=torch.float64, device="cuda")
    results = []
    # a larger tet mesh first, then a small 2D mesh: the second context's sum must ignore the first one's leftovers
    for dim, n, model in ((3, (24, 12, 10), "Stable-Neohookean"), (2, (9, 7, 0), "FixedCorot")):
        V, T = meshgen.kuhn_grid(*n) if dim == 3 else meshgen.tri_grid(n[0], n[1])
        fm = meshgen.pinned_free_map(V)
        st = meshgen.synthetic_state(V, T, 1.0 / n[0], MU)
        ms = mf.MixedStretch(V, T, fm, [(model, MU, LA)])
        ms.reset()
        ms.set("s", st["s"])
        ms.update(st["x"], DT)
        # call 1: the context's own s_ ; call 2, directly after: a trial state s + 0.5 ds in a caller-owned device buffer
        a_ref, a_new = _energy_both(ms, None, e_dev)
        ms.solve(np.random.default_rng(1).standard_normal(dim * ms.nfree) * 1e-3 / n[0])
        sp, ld, nc = ms.field_dev("s")
        dp, _, _ = ms.field_dev("ds")
        s_all = torch.as_tensor(_Dev(sp, ld * nc), device="cuda").clone()
        ds_all = torch.as_tensor(_Dev(dp, ld * nc), device="cuda")
        trial = s_all + 0.5 * ds_all
        b_ref, b_new = _energy_both(ms, trial.data_ptr(), e_dev)
        c_ref, c_new = _energy_both(ms, None, e_dev)             # and back: same value as call 1
        for ref, new in ((a_ref, a_new), (b_ref, b_new), (c_ref, c_new)):
            assert np.isfinite(new) and abs(new - ref) <= TOL * abs(ref), (ref, new)
        assert a_new == c_new and a_ref != b_ref
        results.append((a_new, b_new))
        ms.close()
    assert results[0] != results[1]


def test_energy_noalloc_rejects_bad_arguments(built):
    L = mf.lib()
    V, T = meshgen.kuhn_grid(3, 2, 2)
    ms = mf.MixedStretch(V, T, meshgen.pinned_free_map(V), [("ARAP", MU, LA)])
    ms.reset()
    assert L.mfb_energy_noalloc_dev(ms.h, None, None) == -1                       # MFB_ERR_INVALID: no output pointer
    assert b"e_dev" in L.mfb_last_error_string()
    assert L.mfb_energy_noalloc_dev(None, None, C.c_void_p(8)) == -1              # null context
    ms.close()


class _Dev:
    """view of a raw device pointer for torch (__cuda_array_interface__, as bench.py's DevArray)"""

    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (int(n),), "typestr": "<f8", "data": (int(ptr), False), "version": 2}
