"""Launched by torchrun on N GPUs (tests/test_gpu_dist.py): the distributed PCG of the library (sub-assembled operator,
interface sums and scalar all-reduces over NCCL inside the library) against the single-GPU solve of the same global
system on rank 0.  Prints 'DIST_OK' on every rank when the checks pass."""
import os
import sys
import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import mixed_fem_b200 as mf                      # noqa: E402
from mixed_fem_b200 import meshgen, partition    # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    V, T = meshgen.kuhn_grid(24, 6, 6)
    fm = meshgen.pinned_free_map(V)
    E, nu = 1e6, 0.45
    mu, la = meshgen.enu_to_lame(E, nu)
    mats = [("Stable-Neohookean", mu, la)]
    rng = np.random.default_rng(5)
    x = (V @ meshgen._affine().T + 0.002 * rng.standard_normal(V.shape)).reshape(-1)
    nFg = int(fm.max()) + 1
    bg = rng.standard_normal(3 * nFg)
    h = 0.0333

    local, parts, shared = partition.make_local(V, T, fm, world, rank)
    ms = mf.MixedStretch(local.V, local.T, local.free_map, mats, device=dev.index)
    ms.reset()
    ms.comm_init(local)
    ms.newton_init(density=1000.0, h=h)            # builds the local (sub-assembled) mass values
    ms.update(local.restrict_vertex_vector(x, 3), h)
    _, Mptr = ms.newton_positions_dev()
    nFl = ms.nfree
    bl = torch.as_tensor(bg.reshape(-1, 3)[local.free_l2g].reshape(-1).copy(), device=dev)
    out = {}
    for precond in (1, 2):
        xl = torch.zeros(3 * nFl, dtype=torch.float64, device=dev)
        it, rel = ms.pcg_dev(bl.data_ptr(), xl.data_ptr(), M_ptr=Mptr, precond=precond, tol=1e-10, max_iters=3000)
        out[precond] = (it, rel, xl.cpu().numpy())
    # the default path above is the peer-mailbox one when CUDA IPC works on this node (peer_state 1; -1: NCCL fallback, same
    # on every rank).  The NCCL exchange on the same system: same iteration count, and the same solution to rounding
    # (the partial sums are added in rank order in one path and by NCCL's tree in the other)
    state = ms.peer_state()
    states = [None] * world
    dist.all_gather_object(states, state)
    assert state in (1, -1) and len(set(states)) == 1, states
    print("rank", rank, "peer_state", state, flush=True)
    if os.environ.get("MFB_REQUIRE_PEER") == "1":
        assert state == 1
    # (this system is small: the default above is also the persistent cooperative CG kernel; cg_persist 0 = the three
    # launches per iteration with the same mailboxes, peer 0 = three launches and NCCL)
    for name, val, label in (("cg_persist", 0, "peer, 3 launches"), ("peer", 0, "nccl")):
        ms.set_option(name, val)
        for precond in (1, 2):
            xl = torch.zeros(3 * nFl, dtype=torch.float64, device=dev)
            it, rel = ms.pcg_dev(bl.data_ptr(), xl.data_ptr(), M_ptr=Mptr, precond=precond, tol=1e-10, max_iters=3000)
            a = out[precond]
            d = np.abs(xl.cpu().numpy() - a[2]).max() / np.abs(a[2]).max()
            print("rank", rank, "precond", precond, "iters default /", label, a[0], it, "max rel diff", d, flush=True)
            assert abs(it - a[0]) <= 2 and d <= 1e-9, (it, a[0], d)
    ms.set_option("peer", 1)
    ms.set_option("cg_persist", 2)
    # and again through the mailboxes: sequence numbers carry on across solves
    for precond in (1,):
        xl = torch.zeros(3 * nFl, dtype=torch.float64, device=dev)
        it, rel = ms.pcg_dev(bl.data_ptr(), xl.data_ptr(), M_ptr=Mptr, precond=precond, tol=1e-10, max_iters=3000)
        assert it == out[precond][0] and np.array_equal(xl.cpu().numpy(), out[precond][2]), (it, out[precond][0])
    # gather the distributed solution on rank 0 (interface copies must agree between ranks)
    sol = {}
    for precond in (1, 2):
        xg = torch.zeros(3 * nFg, dtype=torch.float64, device=dev)
        cnt = torch.zeros(nFg, dtype=torch.float64, device=dev)
        idx = torch.as_tensor(local.free_l2g, device=dev)
        xg.view(-1, 3)[idx] = torch.as_tensor(out[precond][2], device=dev).view(-1, 3)
        cnt[idx] = 1.0
        xs = xg.clone(); dist.all_reduce(xs); dist.all_reduce(cnt)
        mean = xs.view(-1, 3) / cnt[:, None]
        # every copy equals the mean  <=>  interface copies are identical
        dev_max = (mean[idx] - xg.view(-1, 3)[idx]).abs().max().item()
        assert dev_max <= 1e-12 * (mean.abs().max().item() + 1e-300), dev_max
        sol[precond] = mean.reshape(-1).cpu().numpy()
    ok = True
    if rank == 0:
        ref = mf.MixedStretch(V, T, fm, mats, device=dev.index)
        ref.reset()
        ref.newton_init(density=1000.0, h=h)
        ref.update(x, h)
        _, Mref = ref.newton_positions_dev()
        bt = torch.as_tensor(bg, device=dev)
        for precond in (1, 2):
            xr = torch.zeros(3 * nFg, dtype=torch.float64, device=dev)
            it1, rel1 = ref.pcg_dev(bt.data_ptr(), xr.data_ptr(), M_ptr=Mref, precond=precond, tol=1e-10, max_iters=3000)
            xr = xr.cpu().numpy()
            err = np.linalg.norm(sol[precond] - xr) / np.linalg.norm(xr)
            print("precond", precond, "iters dist/single", out[precond][0], it1, "relres", out[precond][1], rel1, "rel diff", err, flush=True)
            ok = ok and err <= 1e-8 and abs(out[precond][0] - it1) <= max(3, it1 // 20)
        ref.close()
    flag = torch.tensor([1.0 if ok else 0.0], device=dev)
    dist.broadcast(flag, 0)
    ms.close()
    dist.barrier()
    if flag.item() == 1.0:
        print("DIST_OK", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1.0 else 1)


if __name__ == "__main__":
    main()
