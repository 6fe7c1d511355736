#!/usr/bin/env python
"""Benchmark of the mixed-stretch hot path: tet updates+assemblies per second (FP64).

    python bench.py --gpus N --steps K --warmup W            # CUDA path (this repo)
    python bench.py --impl reference --steps K --warmup W     # CPU arm: the oracle port on the box's host cores

A step = one MixedStretch::update() (F -> polar -> material g,H -> PSD fix -> element blocks -> CSR scatter,
grad_x_, grad_la_, fused Schur rhs) over every tet of the workload, inputs resident in HBM.  The workload is
BASELINE.json configs[3]: the synthetic 10M-tet Kuhn box grid (119^3 x 6), Stable-Neohookean; with N > 1 the SAME grid
is element-partitioned into N slabs by the library's partitioner (strong scaling) and the interface-vertex sums of
grad_x_/rhs_ go point-to-point to the neighbouring slabs, under the interior elements.  Consecutive steps alternate
between two vertex states so the warm-started polar solve does real Newton iterations every step (a repeated x would
converge in zero iterations).  The same line carries the ms-per-Newton-iteration legs (C3, C4: device-resident Newton
driver with the distributed CG - peer mailboxes over CUDA IPC when N > 1), a cross-rank parity object, the roofline of
the update kernels, the tail roofline, clocks, and the CPU baseline (rank 0, N = 1).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "tet_updates_assemblies_per_s"
UNIT = "tets/s"
DT = 0.0333
E_YOUNG, NU = 1e6, 0.45
_NATIVE_LIB = None


def algorithmic_bytes(nE, nV, nF, nb):
    """1372 nE + 24 nV + 24 nF + 72 nb (each datum once, FP64 / int32, symmetric 6x6 packed)."""
    return 1372.0 * nE + 24.0 * nV + 24.0 * nF + 72.0 * nb


class ClockSampler:
    """nvidia-smi clocks + throttle reasons sampled DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        self.rows = []
        self.proc = None
        self.index = index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q, "--format=csv,noheader,nounits",
                                          "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            p = [t.strip() for t in r.split(",")]
            if len(p) < 7:
                continue
            try:
                sm.append(float(p[0])); mx.append(float(p[1]))
            except ValueError:
                continue
            for n, v in zip(names, p[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(n)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_arm(nx, ny, nz, seconds, threads=None, native=True, min_reps=1):
    """Time the oracle (CPU restatement, OpenMP) on the workload grid (or a bounded sample of it): same generator, same
    material, same alternating states.  Returns (tets/s, dict)."""
    from mixed_fem_b200 import meshgen
    from oracle import orc
    global _NATIVE_LIB
    libpath = None
    if native:   # rebuild for this host's ISA into a scratch file (the shipped liborc.so targets x86-64-v3)
        if _NATIVE_LIB is None:
            try:
                _NATIVE_LIB = orc.build(march="native", out="/tmp/liborc_native_%d.so" % os.getpid(), force=True)
            except Exception:
                _NATIVE_LIB = ""
        libpath = _NATIVE_LIB or None
    L = orc.lib(libpath)
    if not threads:
        # all host cores this process may use: torchrun exports OMP_NUM_THREADS=1 to every rank, which would make the
        # CPU arm 10x slower than the machine allows
        try:
            threads = len(os.sched_getaffinity(0))
        except Exception:
            threads = os.cpu_count() or 1
    L.orc_set_num_threads(int(threads))
    cores = L.orc_num_threads()
    V, T = meshgen.kuhn_grid(nx, ny, nz, h=1.0 / 119)
    fm = meshgen.pinned_free_map(V)
    mu, la = meshgen.enu_to_lame(E_YOUNG, NU)
    xa = meshgen.deformed_positions(V, 1.0 / 119, scale=1.0)
    xb = meshgen.deformed_positions(V, 1.0 / 119, scale=0.5)
    o = orc.OracleMixedStretch(V, T, fm, [("Stable-Neohookean", mu, la)], libpath=libpath)
    o.reset()
    o.update_rotations(xb)
    o.set("s", o.get("S"))
    rng = np.random.default_rng(meshgen.SEED)
    o.set("la", rng.uniform(-1, 1, size=6 * T.shape[0]) * mu * 1e-3)
    o.update(xa, DT); o.rhs()          # warm-up
    o.timers(reset=True)
    t0 = time.perf_counter()
    reps = 0
    while True:
        o.update(xb if reps % 2 == 0 else xa, DT)
        o.rhs()
        reps += 1
        if time.perf_counter() - t0 >= seconds and reps >= min_reps:
            break
    el = time.perf_counter() - t0
    tm = o.timers()
    info = {"value": T.shape[0] * reps / el, "unit": UNIT, "cores": int(cores), "kind": "port",
            "sample": "%dx%dx%d x6 = %d tets of the same Kuhn grid, %d update()+rhs() calls in %.1f s, OpenMP, -O3 -march=%s; "
                      "oracle = Eigen-free restatement (generous: no per-element heap matrices)" %
                      (nx, ny, nz, T.shape[0], reps, el, "native" if libpath else "x86-64-v3"),
            "phase_ms_per_call": {k: v / reps for k, v in tm.items()}}
    return info["value"], info


class DevArray:
    """__cuda_array_interface__ view of a raw device pointer so torch can wrap context-owned vectors without copies."""

    def __init__(self, ptr, shape):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": "<f8", "data": (int(ptr), False), "version": 2}


def gpu_arm(args):
    import torch
    import torch.distributed as dist
    import mixed_fem_b200 as mf
    from mixed_fem_b200 import meshgen, partition

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the CUDA path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    nx, ny, nz = args.grid
    h = 1.0 / 119
    V, T = meshgen.kuhn_grid(nx, ny, nz, h=h)
    fm = meshgen.pinned_free_map(V)
    mu, la = meshgen.enu_to_lame(E_YOUNG, NU)
    nE_total, nV_total = T.shape[0], V.shape[0]
    xa = meshgen.deformed_positions(V, h, scale=args.jitter)
    xb = meshgen.deformed_positions(V, h, scale=0.5 * args.jitter)
    tail_states = None
    if world == 1 and args.tail_jitter > 0 and not args.no_tail:
        tail_states = (meshgen.deformed_positions(V, h, scale=args.tail_jitter), meshgen.deformed_positions(V, h, scale=0.5 * args.tail_jitter))
    rng = np.random.default_rng(meshgen.SEED)
    lam = rng.uniform(-1, 1, size=6 * nE_total) * mu * 1e-3
    if world > 1:
        loc, parts, shared = partition.make_local(V, T, fm, world, rank)
        Vl, Tl, fml = loc.V, loc.T, loc.free_map
        xa, xb = loc.restrict_vertex_vector(xa, 3), loc.restrict_vertex_vector(xb, 3)
        lam = loc.restrict_element_vector(lam, 6)
    else:
        loc, Vl, Tl, fml = None, V, T, fm
    del V, T
    ms = mf.MixedStretch(Vl, Tl, fml, [(args.material, mu, la)], device=local_rank)
    ms.set_stream(torch.cuda.current_stream().cuda_stream)
    if args.variant is not None:
        ms.set_option("smem_b", args.variant)
    if args.asm is not None:
        ms.set_option("asm_mode", args.asm)
    if args.exp:
        ms.set_option("exp", args.exp)
    for kv in args.opt:
        name, val = kv.split("=")
        ms.set_option(name, int(val))
    # state: R, s from the 'previous' positions (polar from identity on the device), lambda seeded
    ms.reset(xb)
    ms.set("la", lam)
    nb = ms.nnz // 9
    # device-resident inputs + pinned host copies for the e2e leg
    xs_host = [torch.from_numpy(xa).pin_memory(), torch.from_numpy(xb).pin_memory()]
    xs_dev = [t.to(dev) for t in xs_host]
    rhs_host = torch.empty(3 * ms.nfree, dtype=torch.float64).pin_memory()
    iface = None
    if world > 1:
        # the exchange step lives in the library: pack -> ncclAllReduce -> unpack of grad_x_ | rhs_ at the shared vertices
        # the exchange step lives in the library: interface elements first, point-to-point sum of their grad_x_ | rhs_ with
        # the neighbouring slabs on a side stream under the interior elements (option update_exchange); --opt
        # update_exchange=0 falls back to an explicit exchange after the update
        ms.comm_init(loc)
        if "update_exchange=0" in args.opt:
            gx_ptr, rh_ptr = ms.field_dev("grad_x")[0], ms.rhs_dev()
            iface = lambda: ms.interface_sum_dev(gx_ptr, 3, rh_ptr, 3)
        else:
            ms.set_option("update_exchange", 1)

    def step(i):
        ms.update_dev(xs_dev[i % 2].data_ptr(), DT)
        if iface is not None:
            iface()

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    for i in range(args.warmup):
        step(i)
    sync()
    st0 = ms.status()
    ms.set_option("profile", 1)
    ms.kernel_time_ms()
    ms.launches(reset=True)
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync()
    e0.record()
    for i in range(args.steps):
        step(args.warmup + i)
    e1.record()
    sync()
    ms_total = e0.elapsed_time(e1)
    clocks = sampler.stop() if rank == 0 else None
    launches = ms.launches()
    kt, kn = ms.kernel_time_ms()
    ms.set_option("profile", 0)
    # ---- e2e leg: host buffers through the C ABI, H2D of x and D2H of rhs inside the timed region
    import ctypes as C
    def e2e_step(i):
        xh = xs_host[i % 2]
        ms._check(ms.L.mfb_update(ms.h, C.cast(xh.data_ptr(), C.POINTER(C.c_double)), DT))
        if iface is not None:
            iface()
        ms._check(ms.L.mfb_rhs(ms.h, C.cast(rhs_host.data_ptr(), C.POINTER(C.c_double))))
    for i in range(2):
        e2e_step(i)
    sync()
    t0 = time.perf_counter()
    e0.record()
    for i in range(args.steps):
        e2e_step(i)
    e1.record()
    sync()
    e2e_ms = max(e0.elapsed_time(e1), 0.0)
    e2e_wall = (time.perf_counter() - t0) * 1e3
    parity = parity_pass(ms, loc, xb, xs_dev[0], lam, iface, world, dev, args)
    tail = None
    if tail_states is not None:
        tail = tail_leg(ms, tail_states, lam, dev, nE_total, Vl.shape[0], nb, jitter=args.tail_jitter)
    t = torch.tensor([ms_total, e2e_ms, kt / max(kn, 1)], dtype=torch.float64, device=dev)
    cnt = torch.tensor([launches, 3 * Vl.shape[0] * 8, 3 * ms.nfree * 8, nb], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(cnt, op=dist.ReduceOp.SUM)
    ms_total, e2e_ms, k_ms = t.tolist()
    launches_all, h2d, d2h, nb_all = cnt.tolist()
    if rank == 0:
        per_step = ms_total / args.steps
        value = nE_total / (per_step * 1e-3)
        nF_total = int((fm >= 0).sum())
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        peak = float(peaks.get("hbm_gbs", 6650.0))
        # roofline of the dominant kernel (k_update_tet) on rank 0's share, per launch
        nE_l, nV_l, nF_l = Tl.shape[0], Vl.shape[0], ms.nfree
        abytes = algorithmic_bytes(nE_l, nV_l, nF_l, nb)
        achieved = abytes / (k_ms * 1e-3) / 1e9 if k_ms > 0 else None
        traffic = None
        if world == 1 and nE_l == 10110954:      # the ncu capture in profiles/ is of the 1-GPU C4 workload only
            try:
                traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get("k_update_tet_bytes_per_launch")
            except Exception:
                pass
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "synthetic",
            "config": {"workload": ("C4: synthetic %dx%dx%d Kuhn box grid = %d tets / %d vertices, " % (nx, ny, nz, nE_total, nV_total))
                                   + args.material + " E=1e6 nu=0.45, dt=0.0333, MixedStretch::update() + fused Schur rhs",
                       "partition": "1 GPU" if world == 1 else "%d element slabs (mfb_partition_create), interface elements first, NCCL send/recv of their grad_x_/rhs_ sums with the neighbouring slabs under the interior elements" % world,
                       "l2": "inputs larger than L2 (%.1f GB of per-element state per step), no flush needed" % (1372.0 * nE_total / world / 1e9),
                       "states": "x alternates between two seeded states each step so the warm-started polar iterates"},
            "e2e": {"value": nE_total / (e2e_ms / args.steps * 1e-3), "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                    "api": "mfb_update(x_host) + mfb_rhs(out_host), pinned host buffers", "ms_per_step": e2e_ms / args.steps,
                    "wall_ms_per_step": e2e_wall / args.steps},
            "gpu_launches": int(launches_all),
            "clocks": clocks,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": (achieved / peak) if achieved else None,
                         "traffic": traffic, "kernel": "k_update_tet_v4<fuse_rhs,coop-RED> + k_update_tet_v4_tail + k_finalize_rows + k_deinterleave_vec (one update())", "kernel_ms": k_ms,
                         "algorithmic_bytes_per_launch": abytes, "bytes_per_tet": abytes / nE_l,
                         "peak_source": "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 GB/s (of fallback)",
                         "kernel_share_of_step": k_ms / per_step if per_step > 0 else None},
            "polar_iters_per_tet": st0["polar_iters"] / max(nE_l, 1), "eig_sweeps_per_tet": st0["eig_sweeps"] / max(nE_l, 1),
            "psd_slow_fraction": st0.get("psd_slow", 0) / max(nE_l, 1),
            "parity": parity,
        }
        if tail is not None:
            tail["frac"] = tail["achieved"] / peak
            out["roofline_tail"] = tail
    if not args.no_newton:
        ms.close()
        del ms
        torch.cuda.empty_cache()
        nl = safe_newton_leg(args, args.newton_grid, "ARAP", "C3", rank, world, local_rank)
        if rank == 0:
            if world == 1 and not args.no_cpu:
                nl["cpu_baseline"] = newton_cpu_baseline(args)
            out["newton_iteration"] = nl
        if not args.no_newton_c4:
            nl4 = safe_newton_leg(args, args.grid, args.material, "C4", rank, world, local_rank, steps=2)
            if rank == 0:
                out["newton_iteration_C4"] = nl4
    if rank == 0:
        if world == 1 and not args.no_cpu:
            cx, cy, cz = args.cpu_grid
            _, info = cpu_arm(cx, cy, cz, args.cpu_seconds)
            v1, _ = cpu_arm(cx, cy, cz, min(6.0, args.cpu_seconds), threads=1)
            info["value_1_thread"] = v1            # OMP_NUM_THREADS=1 beside all cores
            out["cpu_baseline"] = info
        print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


def parity_pass(ms, loc, xb, x_dev, lam, iface, world, dev, args):
    """Cross-rank correctness evidence printed with every run: a fixed pass - reset(xb), seeded
    lambda, ONE update at the first state, interface sums - and global scalars of its results that must not depend on the
    number of GPUs: owner-weighted |rhs_|^2 and |grad_x_|^2, v.Av and |Av|^2 for a fixed pseudo-random v (A applied
    sub-assembled + interface sum), the elastic energy and the constraint value.  Compared with the constants of a 1-GPU
    run stored in tests/golden/bench_parity.json (written by --write-parity)."""
    import torch
    import torch.distributed as dist
    ms.reset(xb)
    ms.set("la", lam)
    ms.update_dev(x_dev.data_ptr(), DT)
    if iface is not None:
        iface()
    torch.cuda.synchronize()
    nF = ms.nfree
    if loc is not None:
        w = np.repeat(loc.dot_weight, 3)
        gid = loc.free_l2g
    else:
        w = np.ones(3 * nF)
        gid = np.arange(nF, dtype=np.int64)
    rhs = ms.rhs()
    gx = ms.get("grad_x")
    v = np.sin(0.37 * (3.0 * gid[:, None] + np.arange(3)[None, :]) + 0.11).reshape(-1)
    vd = torch.as_tensor(v, device=dev)
    yd = torch.zeros_like(vd)
    ms.spmv_dev(vd.data_ptr(), yd.data_ptr())
    if world > 1:
        ms.interface_sum_dev(yd.data_ptr(), 3)
    torch.cuda.synchronize()
    y = yd.cpu().numpy()
    e_psi = ms.energy_dev()
    e_con = ms.constraint_value_dev(x_dev.data_ptr())
    vals = torch.tensor([np.dot(w * rhs, rhs), np.dot(w * gx, gx), np.dot(w * v, y), np.dot(w * y, y), e_psi, e_con],
                        dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(vals)
    names = ["rhs_norm2", "grad_x_norm2", "vAv", "Av_norm2", "energy", "constraint_value"]
    got = dict(zip(names, vals.tolist()))
    key = "%dx%dx%d:%s:j%g" % (tuple(args.grid) + (args.material, args.jitter))
    path = os.path.join(ROOT, "tests", "golden", "bench_parity.json")
    try:
        table = json.load(open(path))
    except Exception:
        table = {}
    if args.write_parity and world == 1:
        table[key] = got
        json.dump(table, open(path, "w"), indent=1, sort_keys=True)
    ref = table.get(key)
    out = {"values": got, "reference": "1-GPU run, tests/golden/bench_parity.json" if ref else None}
    if ref:
        err = {k: abs(got[k] - ref[k]) / max(abs(ref[k]), 1e-300) for k in names}
        out["rel_err_vs_1gpu"] = err
        out["ok"] = bool(max(err.values()) <= 1e-10)
    return out


def tail_leg(ms, states, lam, dev, nE_total, nV, nb, steps=8, jitter=None):
    """The same update on a state where a large part of the Hessians is indefinite (vertex jitter x tail_jitter): those
    elements leave the Cholesky fast path for the Jacobi tail kernel.  Reported beside the headline so that the cost of
    the eigen path is driver-recorded."""
    import torch
    xa, xb = states
    xs = [torch.as_tensor(xa, device=dev), torch.as_tensor(xb, device=dev)]
    ms.reset(xb)
    ms.set("la", lam)
    for i in range(3):
        ms.update_dev(xs[i % 2].data_ptr(), DT)
    torch.cuda.synchronize()
    st = ms.status()
    ms.set_option("profile", 1)
    ms.kernel_time_ms()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(steps):
        ms.update_dev(xs[(3 + i) % 2].data_ptr(), DT)
    e1.record()
    torch.cuda.synchronize()
    kt, kn = ms.kernel_time_ms()
    ms.set_option("profile", 0)
    per = e0.elapsed_time(e1) / steps
    k_ms = kt / max(kn, 1)
    ab = algorithmic_bytes(nE_total, nV, ms.nfree, nb)
    return {"jitter": jitter, "psd_slow_fraction": st.get("psd_slow", 0) / nE_total, "eig_sweeps_per_tet": st["eig_sweeps"] / nE_total,
            "value": nE_total / (per * 1e-3), "unit": UNIT, "ms_per_step": per, "kernel_ms": k_ms, "achieved": ab / (k_ms * 1e-3) / 1e9}


def cg_bytes_per_iteration(nb, nF):
    """one PCG iteration: the block-CSR operator once (72 B values + 4 B column index per block) + 7 passes over d*nF vectors"""
    return 76.0 * nb + 7.0 * 24.0 * nF


def safe_newton_leg(args, grid, material, label, rank=0, world=1, local_rank=0, steps=3):
    """newton_leg, but a failure (on any rank) does not cost the headline line: all ranks agree on it, retry once with the
    NCCL exchange if the peer mailboxes were in use, and report the error text otherwise."""
    import torch
    import torch.distributed as dist

    def attempt():
        err, res = None, None
        try:
            res = newton_leg(args, grid, material, label, rank, world, local_rank, steps)
        except Exception as e:          # noqa: BLE001 - reported in the JSON line
            err = "%s: %s" % (type(e).__name__, e)
        bad = err is not None
        if world > 1:
            f = torch.tensor([1.0 if bad else 0.0], device="cuda")
            dist.all_reduce(f, op=dist.ReduceOp.MAX)
            bad = f.item() > 0
        return bad, err, res

    bad, err, res = attempt()
    if bad and world > 1 and "peer=0" not in args.opt:
        first = err or "failure on another rank"
        args.opt.append("peer=0")
        bad, err, res = attempt()
        if not bad and res is not None:
            res["note"] = "first attempt failed (%s); this is the NCCL exchange (option peer = 0)" % first
    if bad:
        return {"workload": label, "error": err or "failure on another rank"}
    return res


def newton_leg(args, grid, material, label, rank=0, world=1, local_rank=0, steps=3):
    """Metric m2 of SURVEY.md section 8d: ms per Newton iteration - `steps` time steps of the device-resident Newton driver
    with the reference's default solver settings (eigen-pcg-diag, itr_tol 1e-4, <= 500 CG iterations, outer_steps 3):
    update + rhs + lhs = M + A + PCG + solve(dx) + line search per iteration.  Timed with CUDA events around the time
    steps (max over ranks) and, per phase, with the library's event spans under the reference's Timer keys.
    N > 1: the same mesh in N element slabs, mfb_newton_step as a collective call (distributed PCG inside the library)."""
    import torch
    import torch.distributed as dist
    import mixed_fem_b200 as mf
    from mixed_fem_b200 import meshgen, partition
    nx, ny, nz = grid
    h = 1.0 / nx
    V, T = meshgen.kuhn_grid(nx, ny, nz, h=h)
    fm = meshgen.pinned_free_map(V)
    mu, la = meshgen.enu_to_lame(E_YOUNG, NU)
    nE = T.shape[0]
    nF_total = int((fm >= 0).sum())
    if world > 1:
        loc, _, _ = partition.make_local(V, T, fm, world, rank)
        ms = mf.MixedStretch(loc.V, loc.T, loc.free_map, [(material, mu, la)], device=local_rank)
        del V, T
        ms.reset()
        ms.comm_init(loc)
    else:
        ms = mf.MixedStretch(V, T, fm, [(material, mu, la)], device=local_rank)
        del V, T
        ms.reset()
    for kv in args.opt:                               # library options for A/B runs (e.g. peer=0: NCCL exchange in the CG)
        name, val = kv.split("=")
        ms.set_option(name, int(val))
    ms.set_stream(torch.cuda.current_stream().cuda_stream)
    ms.newton_init(density=1000.0, h=DT)
    ms.newton_step(outer_steps=1)                     # warm-up time step
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
    ms.launches(reset=True)
    ms.set_option("profile", 1)
    ms.phase_times(reset=True)
    ms.kernel_time_ms()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    logs = []
    e0.record()
    for _ in range(steps):
        logs.append(ms.newton_step(outer_steps=3, precond=1, cg_tol=1e-4, cg_max_iters=500))
    e1.record()
    torch.cuda.synchronize()
    el = e0.elapsed_time(e1)
    ph = ms.phase_times()
    ms.set_option("profile", 0)
    # the same time step with the Chebyshev preconditioner (precond 3, not a reference solver): fewer CG iterations and
    # fewer scalar reductions per SpMV, more SpMVs - reported beside the reference-default line, never in place of it
    cheb = None
    if label == "C3" or world > 1:
        ec0, ec1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ec0.record()
        lc = ms.newton_step(outer_steps=3, precond=3, cg_tol=1e-4, cg_max_iters=500)
        ec1.record()
        torch.cuda.synchronize()
        tc = torch.tensor([ec0.elapsed_time(ec1)], dtype=torch.float64, device=torch.device("cuda", local_rank))
        if world > 1:
            dist.all_reduce(tc, op=dist.ReduceOp.MAX)
        cheb = {"precond": "chebyshev degree 6 over block-Jacobi (mfb_pcg_dev precond 3; not a reference solver)",
                "ms_per_newton_iteration": float(tc[0]) / max(int(lc["n_iters"]), 1), "cg_iterations": [int(v) for v in lc["cg_iters"]]}
    nb_l, nF_l = ms.nnz // 9, ms.nfree
    n = max(sum(int(l["n_iters"]) for l in logs), 1)
    cg_total = max(sum(int(l["cg_iters_total"]) for l in logs), 1)
    tt = torch.tensor([el, ph["linsolve"][0]], dtype=torch.float64, device=torch.device("cuda", local_rank))
    if world > 1:
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    el, lin_ms = tt.tolist()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    cg_ms = lin_ms / cg_total
    cgcyc = None
    try:                                              # block 0's SM cycles per phase of the persistent CG kernel (debug export)
        import ctypes as C
        buf = (C.c_longlong * 12)()
        ms.L.mfb_debug_pcg_cycles.argtypes = [C.c_void_p, C.POINTER(C.c_longlong), C.c_int]
        if ms.L.mfb_debug_pcg_cycles(ms.h, buf, 1) == 0 and sum(buf) > 0:
            names = ["spmv", "barrier1", "sum1", "post1", "collect1", "update", "barrier2", "sum2", "post2", "collect2", "dir", "barrier3"]
            tot_it = max(1, sum(int(v) for l in logs for v in l["cg_iters"]))
            cgcyc = {k: round(buf[i] / tot_it / 1e3, 2) for i, k in enumerate(names)}
    except Exception:
        cgcyc = None
    cgb = cg_bytes_per_iteration(nb_l, nF_l)          # this rank's share
    out = {"workload": "%s: %dx%dx%d Kuhn grid = %d tets, %s, dt=%.4f, eigen-pcg-diag tol 1e-4, %d time steps x 3 Newton iterations" %
                       (label, nx, ny, nz, nE, material, DT, steps),
           "n_gpus": world, "ms_per_newton_iteration": el / n, "newton_iterations": n,
           "cg_iterations": [int(v) for l in logs for v in l["cg_iters"]],
           "ms_per_cg_iteration": cg_ms,
           "cg_persistent_kcycles_per_iteration": cgcyc,
           "cg_exchange": (None if world == 1 else {1: "peer mailboxes (CUDA IPC over NVLink, inside the CG kernels; no NCCL call per iteration)",
                                                    -1: "NCCL (peer mailboxes unavailable on this node)", 0: "NCCL (option peer = 0)"}[ms.peer_state()]
                           if "peer=0" not in args.opt else "NCCL (option peer = 0)"),
           "phase_ms_per_newton_iteration": {k: v[0] / n for k, v in ph.items()},
           "roofline_cg": {"bound": "hbm", "unit": "GB/s", "peak": peak, "achieved": cgb / (cg_ms * 1e-3) / 1e9,
                           "frac": cgb / (cg_ms * 1e-3) / 1e9 / peak, "algorithmic_bytes_per_iteration": cgb,
                           "note": "per rank: 76 B per 3x3 block (values + column index) + 7 vector passes of 24 B per free vertex; "
                                   "time = CUDA-event span of the PCG (key linsolve) / CG iterations, max over ranks"},
           "energy": [float(v) for v in logs[-1]["energy"]], "decrement": [float(v) for v in logs[-1]["decrement"]],
           "alpha": [float(v) for v in logs[-1]["alpha"]], "gpu_launches": ms.launches(), "chebyshev": cheb,
           "timing": "CUDA events around %d time steps (the loop itself branches on device scalars read by the host), max over ranks" % steps}
    ms.close()
    return out


def newton_cpu_baseline(args):
    """CPU arm of the SAME leg on a bounded sample: the oracle's restatement of NewtonOptimizer::step (oracle/newton.py)
    around the C++/OpenMP oracle variable, linear solve = the Eigen ConjugateGradient restatement (diagonal preconditioner,
    tol 1e-4, <= 500 iterations) on scipy's CSR product (one thread: Eigen's RowMajor SpMV is OpenMP-parallel, so this
    part is pessimistic for the reference; update / rhs / solve / energy / constraint run on all cores)."""
    from mixed_fem_b200 import meshgen
    from oracle import newton as onewton
    from oracle import orc
    nx, ny, nz = args.newton_cpu_grid
    V, T = meshgen.kuhn_grid(nx, ny, nz, h=1.0 / 55)
    fm = meshgen.pinned_free_map(V)
    mu, la = meshgen.enu_to_lame(E_YOUNG, NU)
    L = orc.lib(_NATIVE_LIB or None)
    try:
        threads = len(os.sched_getaffinity(0))
    except Exception:
        threads = os.cpu_count() or 1
    L.orc_set_num_threads(int(threads))
    var = orc.OracleMixedStretch(V, T, fm, [("ARAP", mu, la)], libpath=_NATIVE_LIB or None)
    var.reset()
    nt = onewton.OracleNewton(var, V, T, fm, density=1000.0, h=DT)
    nt.step(outer_steps=1, cg=(1, 1e-4, 500))
    for k in nt.timers:
        nt.timers[k] = 0.0
    var.timers(reset=True)
    t0 = time.perf_counter()
    lg = nt.step(outer_steps=3, cg=(1, 1e-4, 500))
    el = time.perf_counter() - t0
    n = max(int(lg["n_iters"]), 1)
    inner = var.timers()
    ph = {k: v * 1e3 / n for k, v in nt.timers.items()}
    ph.update({"energy": inner["energy"] / n, "constraint": inner["constraint"] / n, "rotations": inner["rotations"] / n,
               "derivatives": inner["derivatives"] / n, "Aloc": inner["Aloc"] / n, "assemble": inner["assemble"] / n})
    nE = T.shape[0]
    return {"ms_per_newton_iteration": el * 1e3 / n, "ms_per_newton_iteration_per_Mtet": el * 1e3 / n / (nE / 1e6), "cores": int(L.orc_num_threads()),
            "kind": "port", "cg_iterations": [int(v) for v in lg["cg_iters"]],
            "sample": "%dx%dx%d x6 = %d tets of the C3 grid (ARAP), one time step of 3 Newton iterations; scale by tets for C3" % (nx, ny, nz, nE),
            "phase_ms_per_newton_iteration": ph}


def reference_arm(args):
    """bench.py --impl reference: the CPU implementation of the path (the oracle port: the reference itself cannot be
    compiled here, no Eigen) on the host cores, SAME workload as the CUDA arm (the full --grid, default C4 = 119^3 x 6
    tets; about 25 GB of host memory and 3-4 s per update()+rhs() call on 16 cores), every step one full call."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cx, cy, cz = args.cpu_grid if args.cpu_grid_given else args.grid
    nE = cx * cy * cz * 6
    # one cpu_arm call builds the oracle once and loops update()+rhs() for the requested number of calls
    reps = max(1, args.warmup + args.steps)
    budget = 1.0              # seconds: min_reps decides
    if nE * reps / 2.5e6 > 420.0:      # keep the run inside a few minutes on a small host: fewer calls, said in `sample`
        reps = max(2, int(420.0 * 2.5e6 / nE))
    value, info = cpu_arm(cx, cy, cz, budget, min_reps=reps)
    v1 = None
    if nE <= 2_000_000:
        v1, _ = cpu_arm(cx, cy, cz, 4.0, threads=1)
    info["value"] = value
    info["value_1_thread"] = v1
    same = [cx, cy, cz] == list(args.grid)
    out = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": nE / value * 1e3, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
           "dtype": "f64", "data": "synthetic",
           "config": {"workload": ("C4: synthetic %dx%dx%d Kuhn box grid = %d tets, " % (cx, cy, cz, nE)) + "Stable-Neohookean E=1e6 nu=0.45, dt=0.0333, "
                                  "MixedStretch::update() + rhs() on the host cores (reference cannot be compiled here: no Eigen; oracle port timed instead)",
                      "same_config_as_cuda_arm": same},
           "cpu_baseline": info,
           "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="cuda", choices=["cuda", "reference"])
    ap.add_argument("--grid", type=int, nargs=3, default=[119, 119, 119])
    ap.add_argument("--cpu-grid", type=int, nargs=3, default=None, help="grid of the CPU sample (default: 64 32 32 for the in-run cpu_baseline, the full --grid for --impl reference)")
    ap.add_argument("--newton-cpu-grid", type=int, nargs=3, default=[32, 16, 16])
    ap.add_argument("--tail-jitter", type=float, default=8.0, help="vertex jitter multiplier of the roofline_tail line (about 30 %% of the Hessians indefinite)")
    ap.add_argument("--no-tail", action="store_true")
    ap.add_argument("--no-newton-c4", action="store_true", help="skip the ms-per-Newton-iteration leg on the C4 grid")
    ap.add_argument("--write-parity", action="store_true", help="1 GPU: store this run's parity scalars as the reference constants")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--opt", action="append", default=[], help="library option name=value (mfb_set_option), e.g. no_overlap=1")
    ap.add_argument("--exp", type=int, default=0, help="what-if experiment bits of the update kernel (results are WRONG; profiling aid only)")
    ap.add_argument("--jitter", type=float, default=1.0, help="multiplies the vertex jitter of the synthetic state (SURVEY 8d: 1.0); large values make Hessians indefinite -> Jacobi tail")
    ap.add_argument("--material", default="Stable-Neohookean", help="material of the timed update (the headline metric is quoted on the default)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-newton", action="store_true", help="skip the ms-per-Newton-iteration leg (config C3)")
    ap.add_argument("--newton-grid", type=int, nargs=3, default=[55, 55, 55])
    ap.add_argument("--variant", type=int, default=None, help="kernel variant (library option smem_b): only 3 (k_update_tet_v4) exists; the round-1 kernels were removed")
    ap.add_argument("--asm", type=int, default=None, help="v4 assembly mode (library option asm_mode): 1 bulk reductions into block-CSR (default), 0 FP64 RED into scalar CSR")
    args = ap.parse_args()
    args.cpu_grid_given = args.cpu_grid is not None
    if args.cpu_grid is None:
        args.cpu_grid = [64, 32, 32]
    args.warmup = max(args.warmup, 3) if args.impl == "cuda" else args.warmup
    if args.impl == "reference":
        reference_arm(args)
    else:
        gpu_arm(args)


if __name__ == "__main__":
    main()
