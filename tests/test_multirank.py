"""N > 1 paths on CPU: (i) the library's contiguous sharding of one batch over several devices (host emulation with
two pretend devices), (ii) bench.py's rank plumbing over gloo with world_size 2 (barrier, max over ranks, rank 0
prints, the reference arm runs on rank 0 only)."""
import json
import os
import subprocess
import sys

import numpy as np

from qsearch_b200 import _lib, build, gatesets, lowering, unitaries

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_batch_is_sharded_over_two_devices(monkeypatch):
    rng = np.random.default_rng(2)
    g = gatesets.QubitCNOTLinear()
    c0 = gatesets.ProductGate(g.initial_layer(3))
    c1 = g.successors(c0)[0][0]
    gps = [lowering.lower(c0), lowering.lower(c1)]
    poi = np.array([0, 1, 1, 0, 1, 0, 0], dtype="int32")
    flat = np.concatenate([rng.random(gps[i].n_params) * 2 * np.pi for i in poi])
    T = unitaries.qft(8)
    one = _lib.Library(build.build_emu())
    assert one.num_devices == 1
    ref = one.eval_residuals_jac(gps, poi, flat, T) + one.eval_hs_grad(gps, poi, flat, T)
    ref_solve = one.solve("lm", gps, poi, flat, T)
    monkeypatch.setenv("QSB_EMU_DEVICES", "2")
    two = _lib.Library(build.build_emu())
    assert two.num_devices == 2
    got = two.eval_residuals_jac(gps, poi, flat, T) + two.eval_hs_grad(gps, poi, flat, T)
    for a, b in zip(ref, got):
        assert np.array_equal(a, b)
    for a, b in zip(ref_solve, two.solve("lm", gps, poi, flat, T)):
        assert np.array_equal(a, b)


_WORKER = r"""
import os, sys
sys.path.insert(0, {root!r})
import bench
world = int(os.environ["WORLD_SIZE"]); rank = int(os.environ["RANK"])
dist = bench.dist_setup(world)
bench.barrier(dist)
m = bench.all_max(dist, 10.0 + rank)
assert m == 10.0 + world - 1, m
bench.barrier(dist)
if rank == 0:
    print("MAX", m, flush=True)
"""


def _torchrun(args, port):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port)] + args
    env = dict(os.environ, OMP_NUM_THREADS="1")
    return subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)


def test_gloo_rank_plumbing(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_WORKER.format(root=ROOT))
    out = _torchrun([str(script)], 29731)
    assert out.returncode == 0, out.stderr[-2000:]
    assert "MAX 11.0" in out.stdout


def test_reference_arm_runs_on_rank0_only():
    out = _torchrun(["bench.py", "--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0", "--cpu-evals", "32"], 29732)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["value"] > 0 and d["cpu_baseline"]["kind"] == "port"
    assert d["e2e"]["h2d_bytes_per_step"] == 0
