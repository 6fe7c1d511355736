"""GPU: the host C++ mirror (mixed-fem_b200/host/mixed_stretch.hpp) EXECUTED - tests/harness/shim_run.cpp drives
mfb::MixedStretch<3> and mfb::NewtonStep<3> call for call like the reference drives a MixedVariable<3>
(src/variables/mixed_variable.h:11-69), writing through value() / lambda() - and every vector it produces is compared
with the oracle at 1e-10 (Newton history 1e-8)."""
import os
import struct
import subprocess

import numpy as np
import pytest

from mixed_fem_b200 import meshgen
from oracle import newton as onewton
from oracle import orc
from tests.common import DT, LA, MU, TOL, block_rel, vec_rel

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def build_shim_run(outdir):
    exe = os.path.join(str(outdir), "shim_run")
    libdir = os.path.join(ROOT, "mixed-fem_b200")
    cxx = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else "g++"
    r = subprocess.run([cxx, "-O1", "-std=c++17", "-I", ROOT, os.path.join(ROOT, "tests", "harness", "shim_run.cpp"), "-o", exe,
                        "-L", libdir, "-lmfem_b200", "-Wl,-rpath," + libdir, "-L/usr/local/cuda/lib64",
                        "-Wl,-rpath-link,/usr/local/cuda/lib64"], capture_output=True, text=True)
    assert r.returncode == 0, r.stderr[-2000:]
    return exe


def read_records(path):
    raw = open(path, "rb").read()
    out, pos = [], 0
    while pos < len(raw):
        (n,) = struct.unpack_from("<q", raw, pos)
        pos += 8
        out.append(np.frombuffer(raw, dtype=np.float64, count=n, offset=pos).copy())
        pos += 8 * n
    return out


def test_cpp_shim_runs_on_the_gpu_and_matches_the_oracle(built, tmp_path):
    exe = build_shim_run(tmp_path)
    V, T = meshgen.kuhn_grid(8, 4, 3)
    fm = meshgen.pinned_free_map(V)
    st = meshgen.synthetic_state(V, T, 1.0 / 8, MU)
    nF = int(fm.max()) + 1
    dx = np.random.default_rng(4).standard_normal(3 * nF) * 1e-3 / 8
    fin, fout = str(tmp_path / "in.bin"), str(tmp_path / "out.bin")
    with open(fin, "wb") as fh:
        fh.write(np.array([V.shape[0], T.shape[0]], dtype=np.int32).tobytes())
        fh.write(np.ascontiguousarray(T, dtype=np.int32).tobytes()); fh.write(np.ascontiguousarray(V).tobytes())
        fh.write(np.ascontiguousarray(fm, dtype=np.int32).tobytes()); fh.write(np.array([MU, LA, DT]).tobytes())
        fh.write(st["x"].tobytes()); fh.write(st["s"].tobytes()); fh.write(st["la"].tobytes())
        fh.write(np.array([nF], dtype=np.int32).tobytes()); fh.write(dx.tobytes())
    r = subprocess.run([exe, fin, fout], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, (r.stdout[-1000:], r.stderr[-1000:])
    rec = read_records(fout)
    assert len(rec) == 15
    cpu = orc.OracleMixedStretch(V, T, fm, [("Stable-Neohookean", MU, LA)])
    cpu.reset()
    assert block_rel(rec[0], cpu.get("s"), 6) <= TOL
    cpu.set("s", st["s"]); cpu.set("la", st["la"])
    cpu.update(st["x"], DT)
    assert vec_rel(rec[1], cpu.rhs()) <= TOL
    assert vec_rel(rec[2], cpu.csr()[2]) <= TOL
    assert vec_rel(rec[3], cpu.get("grad_x")) <= TOL
    assert vec_rel(rec[4], cpu.get("grad")) <= TOL
    assert vec_rel(rec[5], cpu.get("grad_la")) <= TOL
    cpu.solve(dx)
    assert block_rel(rec[7], cpu.get("la"), 6) <= TOL
    assert block_rel(rec[6], cpu.get("ds"), 6) <= 1e-8
    s_trial = cpu.get("s") + 0.5 * cpu.get("ds")
    assert abs(rec[8][0] - cpu.energy(st["x"], s_trial)) <= TOL * abs(cpu.energy(st["x"], s_trial))
    cc = cpu.constraint_value(st["x"], s_trial)
    assert abs(rec[8][1] - cc) <= 1e-9 * max(abs(cc), 1e-12)
    assert vec_rel(rec[9], cpu.evaluate_constraint(st["x"])) <= TOL
    assert vec_rel(rec[10], cpu.get("evals")) <= TOL
    assert not rec[11].any()
    # NewtonStep::step from the rest state against the oracle's restatement of NewtonOptimizer::step
    cpu.reset()
    nt = onewton.OracleNewton(cpu, V, T, fm, density=1000.0, h=0.02)
    lo = nt.step(outer_steps=3, newton_tol=1e-10)
    hist = rec[12]
    n = int(hist[0])
    assert n == lo["n_iters"]
    h = hist[1:].reshape(n, 3)
    assert np.all(np.abs(h[:, 0] - lo["energy"]) <= 1e-8 * np.abs(lo["energy"]))
    assert np.all(np.abs(h[:, 1] - lo["decrement"]) <= 1e-8 * np.abs(lo["decrement"]) + 1e-12)
    assert np.array_equal(h[:, 2], lo["alpha"])
    assert vec_rel(rec[13], nt.x) <= 1e-9
    assert not rec[14].any()
