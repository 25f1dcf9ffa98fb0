#!/usr/bin/env python
"""Benchmark of the qsearch hot path on B200: batched circuit objective + full parameter Jacobian.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--qubits 4|5]

Metric (BASELINE.json): objective+Jacobian evaluations / second, fp64, on the microbenchmark of configs[3]:
random ring circuits of depth 10 (4 qubits by default: G = 11 layers, P = 62 parameters), 1e5 parameter vectors.
One "eval" = LM form: residual vector r (2N^2) and residual Jacobian J_r (2N^2 x P) of one (circuit, vector).
One "step" = one pass over all instances.  `value` is timed with CUDA events with inputs resident in HBM and J_r
written to HBM; `e2e` goes through the host-buffer C-ABI call (H2D of parameters, D2H of r and J_r inside the
timed region).  Multi-GPU: one process per GPU, each rank owns its own 1e5 instances (weak scaling, no collective:
SURVEY.md section 8e); torch.distributed (gloo) is used only for the barrier and the max-over-ranks of the times.
"""
import argparse
import json
import multiprocessing as mp
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from qsearch_b200 import workloads, lowering  # noqa: E402  (host-only imports; no CUDA yet)


def flops_lm(n, G, P):
    """Algorithmic flops per LM-form eval, dense prefix/suffix formulation."""
    N3 = (1 << n) ** 3
    return 8 * N3 * (2 * (G - 1) + 2 * P) + 8 * N3 * (P + 1)


def flops_hs(n, G, P):
    N = 1 << n
    return 8 * N ** 3 * (2 * (G - 1) + 2 * P) + 8 * N * N * (P + 1)


# ----------------------------------------------------------------------------- CPU arm (oracle port)
_CPU = {}


def _cpu_init(n, n_circuits, target_name):
    from threadpoolctl import threadpool_limits
    _CPU["limit"] = threadpool_limits(1)
    circuits, params, targets = workloads.microbench_workload(n, n_circuits, 8)
    _CPU["circuits"], _CPU["params"], _CPU["T"] = circuits, params, targets[target_name]


def _cpu_work(idx):
    from oracle import qs_oracle as O
    t0 = time.perf_counter()
    for ci, vi in idx:
        O.eval_lm(_CPU["circuits"][ci], _CPU["T"], _CPU["params"][ci, vi])
    return time.perf_counter() - t0


def cpu_arm(n, n_evals, procs, target_name="qft"):
    """Time the CPU restatement of the reference's NumPy path (oracle/qs_oracle.py) on `procs` forked workers."""
    n_circuits = 16
    idx = [(i % n_circuits, (i // n_circuits) % 8) for i in range(n_evals)]
    chunks = [idx[p::procs] for p in range(procs)]
    ctx = mp.get_context("fork")
    with ctx.Pool(procs, initializer=_cpu_init, initargs=(n, n_circuits, target_name)) as pool:
        pool.map(_cpu_work, [c[:2] for c in chunks])  # warm-up
        t0 = time.perf_counter()
        pool.map(_cpu_work, chunks)
        dt = time.perf_counter() - t0
    return n_evals / dt, dt


# ----------------------------------------------------------------------------- helpers
class ClockSampler:
    def __init__(self, gpu_index):
        self.gpu = gpu_index
        self.rows = []
        self._stop = threading.Event()
        self._th = None

    def start(self):
        def loop():
            q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
                "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
            while not self._stop.is_set():
                try:
                    out = subprocess.run(["nvidia-smi", "-i", str(self.gpu), "--query-gpu=" + q, "--format=csv,noheader,nounits"],
                                         capture_output=True, text=True, timeout=5).stdout.strip()
                    if out:
                        self.rows.append([x.strip() for x in out.split(",")])
                except Exception:
                    pass
                self._stop.wait(0.2)
        self._th = threading.Thread(target=loop, daemon=True)
        self._th.start()

    def stop(self):
        self._stop.set()
        if self._th:
            self._th.join(timeout=6)
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 7 for i in range(4) if r[3 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def dist_setup(world):
    if world <= 1:
        return None
    import torch.distributed as dist
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    dist.init_process_group("gloo", rank=int(os.environ["RANK"]), world_size=world)
    return dist


def all_max(dist, x):
    if dist is None:
        return x
    import torch
    t = torch.tensor([x], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t[0])


def barrier(dist):
    if dist is not None:
        dist.barrier()


def load_peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        return {}


# ----------------------------------------------------------------------------- arms
def workload_config(args, n, G, P):
    return {"workload": "configs[3]: {}q QubitCNOTRing depth-10 circuits (G={}, P={}), {} circuits x {} parameter vectors "
                        "= {} instances per GPU; LM form (r, J_r) per instance; target qft({})".format(
                            n, G, P, args.circuits, args.vectors, args.circuits * args.vectors, 1 << n),
            "n_qubits": n, "layers": G, "params": P, "instances_per_gpu": args.circuits * args.vectors,
            "algorithmic_flop_per_eval": flops_lm(n, G, P),
            "l2": "each step writes {:.1f} GB of J_r through L2 (>> 126 MB) between re-reads of the parameters".format(
                args.circuits * args.vectors * (P + 1) * 2 * (1 << (2 * n)) * 8 / 1e9)}


def run_reference(args, rank, world):
    """--impl reference: the reference's CPU path for the same workload on the host cores.  /root/reference does not
    exist on the GPU box and qsrs (Rust) is not buildable, so this times oracle/qs_oracle.py, the NumPy restatement
    pinned to reference-generated fixtures (cpu_baseline.kind = "port")."""
    if rank != 0:
        return
    n = args.qubits
    circuits, _ = workloads.microbench_circuits(n, 1, 10)
    G, P = len(circuits[0]._subgates), circuits[0].num_inputs
    procs = os.cpu_count() or 1
    per_step = max(procs * 4, int(args.cpu_evals))
    for _ in range(max(0, args.warmup - 1)):
        cpu_arm(n, procs * 2, procs)
    vals, t_total = [], 0.0
    for _ in range(args.steps):
        v, dt = cpu_arm(n, per_step, procs)
        vals.append(v)
        t_total += dt
    value = per_step * args.steps / t_total
    line = {"impl": "reference", "metric": "objective+Jacobian evals/sec", "value": value, "unit": "evals/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_total / args.steps,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(args, n, G, P),
            "cpu_baseline": {"value": value, "unit": "evals/s", "cores": procs, "kind": "port",
                             "sample": "{} LM-form evals per step over {} forked workers (BLAS threads = 1)".format(per_step, procs)},
            "e2e": {"value": value, "unit": "evals/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def run_ours(args, rank, world, local_rank):
    n = args.qubits
    # CPU baseline first (forks workers): must happen before this process creates a CUDA context.
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu:
        procs = os.cpu_count() or 1
        n_evals = max(procs * 4, int(args.cpu_evals))
        v, dt = cpu_arm(n, n_evals, procs)
        cpu = {"value": v, "unit": "evals/s", "cores": procs, "kind": "port",
               "sample": "{} LM-form evals of the same workload in {:.1f} s on {} forked workers, oracle/qs_oracle.py "
                         "(NumPy restatement of the reference path; qsrs is not buildable offline)".format(n_evals, dt, procs)}
    dist = dist_setup(world)

    from qsearch_b200 import _lib
    lib = _lib.Library(devices=[local_rank])
    circuits, params, targets = workloads.microbench_workload(n, args.circuits, args.vectors)
    if world > 1:  # every rank owns different parameter vectors of the same circuit set
        params = np.random.default_rng(workloads.SEED_BASE + 1000 + rank).random(params.shape) * 2 * np.pi
    G, P = len(circuits[0]._subgates), circuits[0].num_inputs
    N = 1 << n
    gps = [lowering.lower(c) for c in circuits]
    poi = np.repeat(np.arange(args.circuits, dtype="int32"), args.vectors)
    n_inst = len(poi)
    T = targets["qft"]
    flat = np.ascontiguousarray(params.reshape(n_inst, P))

    # -- parity gate before any timing (small slice, NumPy oracle) ------------------------------------------
    from oracle import qs_oracle as O
    chk = [0, n_inst // 2, n_inst - 1]
    r, Jr = lib.eval_residuals_jac(gps, poi[chk], flat[chk], T)
    Jr = Jr.reshape(len(chk), P, 2 * N * N)
    worst = 0.0
    for s, i in enumerate(chk):
        ro, Jo = O.eval_lm(circuits[poi[i]], T, flat[i])
        worst = max(worst, np.abs(r[s] - ro).max(), np.abs(Jr[s].T - Jo).max())
    if not worst < 1e-12:
        raise SystemExit("parity gate failed: max abs error {:.3e} vs oracle".format(worst))

    # -- device-resident timing -------------------------------------------------------------------------
    batch = lib.batch(gps, poi, slot=0)
    batch.set_params(flat)
    batch.set_target(T)
    fp64_peak = lib.measure_fp64_peak(0, False)
    dmma_peak = lib.measure_fp64_peak(0, True)
    for _ in range(args.warmup):
        batch.run(_lib.RUN_RESID, timed=True)
    launches0 = lib.launch_count()
    sampler = ClockSampler(local_rank)
    sampler.start()
    barrier(dist)
    batch.sync()
    t_dev, t_wall0 = 0.0, time.perf_counter()
    for _ in range(args.steps):
        t_dev += batch.run(_lib.RUN_RESID, timed=True)  # CUDA events on the launching stream
    batch.sync()
    t_wall = time.perf_counter() - t_wall0
    barrier(dist)
    clocks = sampler.stop()
    launches = lib.launch_count() - launches0
    t_dev = all_max(dist, t_dev * 1e-3)
    t_wall = all_max(dist, t_wall)
    t = max(t_dev, 1e-12)
    value = world * n_inst * args.steps / t
    kern_ms = 1e3 * t_dev / args.steps
    out_bytes = batch.output_bytes(_lib.RUN_RESID, n_inst)

    # HS form (f, grad f) on the same instances, device-resident
    for _ in range(2):
        batch.run(_lib.RUN_HS, timed=True)
    hs_ms = min(batch.run(_lib.RUN_HS, timed=True) for _ in range(3))
    # LM-system form (r'r, J_r'r, J_r'J_r per instance; J_r stays in L2) on a quarter of the instances
    n_sys = max(1, n_inst // 4)
    batch.run(_lib.RUN_LMSYS, count=n_sys, timed=True)
    sys_ms = min(batch.run(_lib.RUN_LMSYS, count=n_sys, timed=True) for _ in range(2))
    batch.close()

    # -- end to end through the host-buffer C-ABI call ------------------------------------------------------
    e_inst = min(n_inst, args.e2e_instances)
    sel = slice(0, e_inst)
    out_r = lib.pinned_empty((e_inst, 2 * N * N))      # page-locked host buffers (qsb_host_alloc)
    out_J = lib.pinned_empty((e_inst * P, 2 * N * N))
    # warm-up: page faults of the output buffers, device buffers, host threads.  On a fresh box the first calls of a process
    # run at 0.1-0.8x of the steady rate for a while (profiles/r1/as_e2e_ab.log), so warm up until two consecutive calls
    # agree within 5 % (at least max(3, W) calls, at most 16)
    prev = None
    for w in range(16):
        tw = time.perf_counter()
        lib.eval_residuals_jac(gps, poi[sel], flat[sel], T, out_r, out_J)
        tw = time.perf_counter() - tw
        if w + 1 >= max(3, args.warmup) and prev is not None and abs(tw - prev) <= 0.05 * max(tw, prev):
            break
        prev = tw
    barrier(dist)
    t0 = time.perf_counter()
    e_steps = max(1, min(args.steps, 3))
    for _ in range(e_steps):
        lib.eval_residuals_jac(gps, poi[sel], flat[sel], T, out_r, out_J)
    t_e2e = all_max(dist, time.perf_counter() - t0)
    e2e_value = world * e_inst * e_steps / t_e2e
    for _ in range(2):
        lib.eval_hs_grad(gps, poi, flat, T)
    t0 = time.perf_counter()
    f, g = lib.eval_hs_grad(gps, poi, flat, T)
    t_hs_e2e = all_max(dist, time.perf_counter() - t0)

    if rank != 0:
        return
    F = flops_lm(n, G, P)
    peaks = load_peaks()
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    achieved_tf = (n_inst / (kern_ms * 1e-3)) * F / 1e12
    # what the warp kernel really issues on the FP64 tensor pipe: P half-size complex products (N x N/2 x N, rank-one form
    # of the derivative) + one dense product for U T^dagger = (4 P + 8) N^3 flop per eval
    executed_dmma_tf = (n_inst / (kern_ms * 1e-3)) * (4.0 * P + 8.0) * (N ** 3) / 1e12
    ncu_traffic = {4: 255610.0, 5: 1068462.0}.get(n)  # dram read+write bytes per eval: profiles/r1/an_ncu_full_resid_n{4,5}_warp.txt
    line = {
        "metric": "objective+Jacobian evals/sec", "value": value, "unit": "evals/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": kern_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": workload_config(args, n, G, P),
        "clocks": clocks, "gpu_launches": int(launches), "wall_s_timed_region": t_wall,
        "e2e": {"value": e2e_value, "unit": "evals/s", "h2d_bytes_per_step": int(e_inst * P * 8 + 2 * N * N * 8),
                "d2h_bytes_per_step": int(e_inst * (P + 1) * 2 * N * N * 8),
                "note": "qsb_eval_residuals_jac with host buffers, {} instances per step (r and J_r copied back: "
                        "PCIe-bound by the {:.2f} MB of Jacobian per eval)".format(e_inst, (P + 1) * 2 * N * N * 8 / 1e6)},
        "roofline": {"bound": "tensor", "achieved": achieved_tf, "peak": dmma_peak, "unit": "TFLOP/s",
                     "frac": achieved_tf / dmma_peak if dmma_peak else None,
                     "traffic": int(ncu_traffic * n_inst) if ncu_traffic else None,
                     "peak_source": "FP64 tensor pipe (DMMA m8n8k4) micro-benchmark run by this command (qsb_measure_fp64_peak); "
                                    "FP64 is not in MEASURED_PEAKS.json (bf16 and HBM only)",
                     "dfma_peak": fp64_peak,
                     "flops": "achieved = ALGORITHMIC flops (dense prefix/suffix formulation, SURVEY 8d: 8N^3(2(G-1)+2P) + 8N^3(P+1)) per "
                              "second; gate structure lets the kernel execute fewer, so frac can exceed 1. Executed on the tensor pipe: "
                              "(4 P + 8) N^3 flop per eval (see executed); DFMA work shares the same FP64 pipe",
                     "executed": {"dmma_tflops": executed_dmma_tf, "frac_of_dmma_peak": executed_dmma_tf / dmma_peak if dmma_peak else None,
                                  "ncu_dmma_pipe_active_pct": {4: 44.9, 5: 54.2}.get(n),
                                  "ncu_fp64_dfma_pipe_active_pct": {4: 21.9, 5: 19.3}.get(n)},
                     "traffic_note": "ncu dram bytes per eval (one --set full capture, profiles/r1) x instances of this launch",
                     "hbm": {"achieved": out_bytes / (kern_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                             "frac": out_bytes / (kern_ms * 1e-3) / 1e9 / hbm_peak,
                             "bytes_per_eval": out_bytes / n_inst}},
        "hs_form": {"value": world * n_inst / (hs_ms * 1e-3), "unit": "evals/s", "ms_per_step": hs_ms,
                    "e2e_value": world * n_inst / t_hs_e2e, "algorithmic_flop_per_eval": flops_hs(n, G, P)},
        "lm_system_form": {"value": world * n_sys / (sys_ms * 1e-3), "unit": "evals/s", "instances": n_sys,
                           "what": "r'r, J_r'r, J_r'J_r per instance (QSB_RUN_LMSYS: DMMA Jacobian + DMMA SYRK, J_r never written to HBM)",
                           "algorithmic_flop_per_eval": F + 2 * 2 * N * N * (P + 1) * (P + 1)},
        "parity_max_abs_err": worst,
    }
    if cpu is not None:
        line["cpu_baseline"] = cpu
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--qubits", type=int, default=4)
    ap.add_argument("--circuits", type=int, default=100)
    ap.add_argument("--vectors", type=int, default=1000)
    ap.add_argument("--e2e-instances", type=int, default=16384)
    ap.add_argument("--cpu-evals", type=float, default=0)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.cpu_evals <= 0:  # about 10-20 s of CPU work: ~100 evals/s/core for n=4 in NumPy
        args.cpu_evals = (1200 if args.qubits <= 4 else 700) * max(1, (os.cpu_count() or 1)) / 8 * 1.0
        args.cpu_evals = max(args.cpu_evals, 8 * (os.cpu_count() or 1))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
