"""Launched by torchrun on N GPUs (tests/test_gpu_dist.py): MixedStretch::update() on N element parts with the interface
sum of grad_x_ / rhs_ done INSIDE update (option update_exchange: interface elements first, point-to-point exchange
on a side stream under the interior elements) against the single-GPU update of the same global problem on rank 0 -
every per-vertex vector, the sub-assembled matrix applied to a vector, and the per-element fields.  Also the three
exchange paths against each other: overlapped p2p, plain p2p, all-reduce over all shared vertices.  Slab and RCB parts,
a state with a third of the Hessians indefinite (the Jacobi tail list is processed in two pieces on the overlapped path).
Prints 'DIST_OK' on every rank when the checks pass."""
import os
import sys
import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import mixed_fem_b200 as mf                      # noqa: E402
from mixed_fem_b200 import meshgen, partition    # noqa: E402


def gather_free(local, vec, nFg, dev):
    """per-vertex (nFl x 3) local vector -> global mean over the copies + max deviation between copies"""
    xg = torch.zeros(3 * nFg, dtype=torch.float64, device=dev)
    cnt = torch.zeros(nFg, dtype=torch.float64, device=dev)
    idx = torch.as_tensor(local.free_l2g, device=dev)
    v = torch.as_tensor(vec, device=dev).view(-1, 3)
    xg.view(-1, 3)[idx] = v
    cnt[idx] = 1.0
    dist.all_reduce(xg); dist.all_reduce(cnt)
    mean = xg.view(-1, 3) / cnt[:, None]
    dev_max = (mean[idx] - v).abs().max().item() if v.numel() else 0.0
    t = torch.tensor([dev_max], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return mean.reshape(-1).cpu().numpy(), t.item()


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    V, T = meshgen.kuhn_grid(20, 8, 6)
    fm = meshgen.pinned_free_map(V)
    mu, la = meshgen.enu_to_lame(1e6, 0.45)
    mats = [("Stable-Neohookean", mu, la)]
    st = meshgen.synthetic_state(V, T, 1.0 / 20, mu)
    rng = np.random.default_rng(7)
    s = st["s"].reshape(-1, 6).copy()
    sel = rng.uniform(size=s.shape[0]) < 0.33            # indefinite Hessians -> Jacobi tail
    s[sel, :3] *= rng.uniform(0.45, 0.8, size=(int(sel.sum()), 1))
    h = 0.0333
    nFg = int(fm.max()) + 1
    vg = np.sin(0.37 * np.arange(3 * nFg) + 0.11)
    ok = True
    ref = None
    if rank == 0:
        g = mf.MixedStretch(V, T, fm, mats, device=dev.index)
        g.reset(); g.set("R", st["R"]); g.set("s", s.reshape(-1)); g.set("la", st["la"])
        g.update(st["x"], h)
        vd = torch.as_tensor(vg, device=dev); yd = torch.zeros_like(vd)
        g.spmv_dev(vd.data_ptr(), yd.data_ptr()); torch.cuda.synchronize()
        ref = {"grad_x": g.get("grad_x"), "rhs": g.rhs(), "Av": yd.cpu().numpy(), "Hinv": g.get("Hinv"), "dSdF": g.get("dSdF"),
               "slow": g.status()["psd_slow"]}
        g.close()
    for method in (0, 1):
        local, parts, shared = partition.make_local(V, T, fm, world, rank, method=method)
        assert local.n_if_elems > 0 and local.nbr_rank.size > 0
        res = {}
        for mode in ("overlap", "p2p", "allreduce"):
            ms = mf.MixedStretch(local.V, local.T, local.free_map, mats, device=dev.index)
            ms.reset()
            ms.comm_init(local)
            ms.set("R", st["R"].reshape(-1, 9)[local.elems].reshape(-1))
            ms.set("s", local.restrict_element_vector(s.reshape(-1), 6)); ms.set("la", local.restrict_element_vector(st["la"], 6))
            xl = torch.as_tensor(local.restrict_vertex_vector(st["x"], 3), device=dev)
            if mode == "overlap":
                ms.set_option("update_exchange", 1)
            if mode == "allreduce":
                ms.set_option("p2p", 0)
            for _ in range(2):             # twice: buffers, events and the tail marker are reused
                ms.update_dev(xl.data_ptr(), h)
            if mode != "overlap":
                ms.interface_sum_dev(ms.field_dev("grad_x")[0], 3, ms.rhs_dev(), 3)
            torch.cuda.synchronize()
            vl = torch.as_tensor(vg.reshape(-1, 3)[local.free_l2g].reshape(-1).copy(), device=dev)
            yl = torch.zeros_like(vl)
            ms.spmv_dev(vl.data_ptr(), yl.data_ptr())
            ms.interface_sum_dev(yl.data_ptr(), 3)
            torch.cuda.synchronize()
            out = {}
            for name, vec in (("grad_x", ms.get("grad_x")), ("rhs", ms.rhs()), ("Av", yl.cpu().numpy())):
                out[name], dmax = gather_free(local, vec, nFg, dev)
                scale = np.abs(out[name]).max()
                assert dmax <= 1e-13 * scale, (mode, name, dmax, scale)        # copies of a shared vertex agree
            slow = torch.tensor([float(ms.status()["psd_slow"])], device=dev); dist.all_reduce(slow)
            out["slow"] = slow.item()
            out["Hinv"] = ms.get("Hinv").reshape(-1, 36); out["dSdF"] = ms.get("dSdF").reshape(-1, 54)
            res[mode] = out
            ms.close()
        if rank == 0:
            for mode, out in res.items():
                for name in ("grad_x", "rhs", "Av"):
                    err = np.linalg.norm(out[name] - ref[name]) / np.linalg.norm(ref[name])
                    print("method", method, mode, name, "rel diff vs 1 GPU", err, flush=True)
                    ok = ok and err <= 1e-12
                ok = ok and out["slow"] == ref["slow"] and ref["slow"] > 0
            for name, w in (("Hinv", 36), ("dSdF", 54)):
                a = res["overlap"][name]; b = ref[name].reshape(-1, w)[local.elems]
                ok = ok and np.abs(a - b).max() <= 1e-12 * np.abs(b).max()
        for name in ("grad_x", "rhs", "Av"):       # the three exchange paths agree to the last bits of a + b vs an all-reduce
            for mode in ("p2p", "allreduce"):
                d = np.abs(res["overlap"][name] - res[mode][name]).max() / np.abs(res["overlap"][name]).max()
                ok = ok and d <= 1e-14
        local.close()
    flag = torch.tensor([1.0 if ok else 0.0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.barrier()
    if flag.item() == 1.0:
        print("DIST_OK", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1.0 else 1)


if __name__ == "__main__":
    main()
