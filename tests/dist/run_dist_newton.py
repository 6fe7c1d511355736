"""Launched by torchrun on N GPUs (tests/test_gpu_dist.py): two device-resident Newton time steps with the mesh split
into N element slabs (mfb_newton_step as a collective call) against the single-GPU run of the same problem on rank 0:
energy / decrement / alpha histories at the 1e-8 bar, final positions.  Prints 'DIST_OK' on every rank."""
import os
import sys
import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import mixed_fem_b200 as mf                      # noqa: E402
from mixed_fem_b200 import meshgen, partition    # noqa: E402
from tests.common import compare_histories       # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dev = torch.device("cuda", torch.cuda.current_device())
    dist.init_process_group("nccl", device_id=dev)
    V, T = meshgen.kuhn_grid(12, 4, 4)
    fm = meshgen.pinned_free_map(V)
    mu, la = meshgen.enu_to_lame(1e6, 0.45)
    ok = True
    for model in ("Stable-Neohookean", "ARAP"):
        mats = [(model, mu, la)]
        local, parts, shared = partition.make_local(V, T, fm, world, rank)
        ms = mf.MixedStretch(local.V, local.T, local.free_map, mats, device=dev.index)
        ms.reset()
        ms.comm_init(local)
        ms.newton_init(density=1000.0, h=0.02)
        ms.set_option("cg_zero_guess", 1)
        logs = [ms.newton_step(outer_steps=o, newton_tol=1e-10, precond=2, cg_tol=1e-13, cg_max_iters=5000) for o in (3, 6)]
        xl = ms.newton_positions()
        if rank == 0:
            ref = mf.MixedStretch(V, T, fm, mats, device=dev.index)
            ref.reset()
            ref.newton_init(density=1000.0, h=0.02)
            ref.set_option("cg_zero_guess", 1)
            for k, o in enumerate((3, 6)):     # only the last step runs into the termination test (tests/test_gpu_newton.py::_run)
                lr = ref.newton_step(outer_steps=o, newton_tol=1e-10, precond=2, cg_tol=1e-13, cg_max_iters=5000)
                compare_histories(logs[k], lr)
            xr = ref.newton_positions().reshape(-1, 3)[local.vmap].reshape(-1)
            err = np.linalg.norm(xl - xr) / np.linalg.norm(xr)
            print(model, "newton iterations", [l["n_iters"] for l in logs], "positions rel diff", err, flush=True)
            ok = ok and err <= 1e-9
            ref.close()
        ms.close()
    flag = torch.tensor([1.0 if ok else 0.0], device=dev)
    dist.broadcast(flag, 0)
    dist.barrier()
    if flag.item() == 1.0:
        print("DIST_OK", flush=True)
    dist.destroy_process_group()
    sys.exit(0 if flag.item() == 1.0 else 1)


if __name__ == "__main__":
    main()
