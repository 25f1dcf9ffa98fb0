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


# ----------------------------------------------------------------------------- solve form: CPU arm (scipy LM on the oracle)
QSRS_CAVEAT = ("CPU arm = NumPy restatement of the reference's NumPy backend (kind 'port'); the reference's own Rust backend "
               "qsrs, which cannot be built offline, claims 'up to a 20x speedup' over it (qsrs/README.md:5): divide CPU "
               "ratios by up to 20 for a qsrs-equivalent figure")


SOLVE_FORM_SEEDS = (0, 1, 2)  # three seeded frontiers: the wall time of a batch is its slowest instance, one draw of which is noisy


def _solve_target(n, which, seed=0):
    children, x0s, targets = workloads.frontier_workload(n, seed=seed)
    if which == "unreachable":
        return children, x0s, targets["unreachable"]
    from oracle import qs_oracle as O
    c0, x_star = targets["reachable"]
    return children, x0s, O.matrix(c0, x_star)


def _cpu_solve_work(job):
    """One reference-style solve: scipy.optimize.least_squares(method='lm') on the oracle's residuals / residual Jacobian
    (qsearch/solvers.py:155-157: the call LeastSquares_Jac_Solver makes)."""
    from threadpoolctl import threadpool_limits
    import scipy.optimize
    from oracle import qs_oracle as O
    n, which, ci, si = job
    with threadpool_limits(1):
        children, x0s, T = _solve_target(n, which)
        c = children[ci]
        t0 = time.perf_counter()
        res = scipy.optimize.least_squares(lambda x: O.matrix_residuals(T, O.matrix(c, x)), x0s[ci][si],
                                           jac=lambda x: O.matrix_residuals_jac(T, *O.mat_jac(c, x)), method="lm")
        dt = time.perf_counter() - t0
    return dt, int(res.njev), float(O.matrix_distance_squared(T, O.matrix(c, res.x)))


def cpu_solve_arm(n, procs):
    """One solve per host core (bounded sample of the solve-form workload), all cores busy at once."""
    children, _, _ = workloads.frontier_workload(n)
    jobs = [(n, "unreachable" if k % 2 == 0 else "reachable", (k // 2 * 5) % len(children) if k % 2 == 0 else 0, k % 4) for k in range(procs)]
    ctx = mp.get_context("fork")
    t0 = time.perf_counter()
    with ctx.Pool(procs) as pool:
        out = pool.map(_cpu_solve_work, jobs)
    wall = time.perf_counter() - t0
    njev = sum(o[1] for o in out)
    return {"solves_per_s": len(jobs) / wall, "solves": len(jobs), "wall_s": wall, "cores": procs, "jac_evals": njev,
            "us_per_lm_iteration_per_core": 1e6 * sum(o[0] for o in out) / max(1, njev),
            "solver": "scipy.optimize.least_squares(method='lm') on oracle/qs_oracle.py, one solve per core", "caveat": QSRS_CAVEAT}


def solve_form(lib, n, dist, world, cpu_solve):
    """Frontier-shaped batch (48 children x 4 starts; once against an unreachable target, once against a reachable one)
    through Solver.solve_batch end to end: host start vectors in, (xopt, distance) out."""
    import qsearch_b200 as q
    from qsearch_b200 import solvers, _lib as L
    prev = L.use_library(lib)
    try:
        out = {}
        for frontier in (False, True):
            tot_t, tot_inst, tot_it, max_it, best = 0.0, 0, 0, 0, []
            for which, seed in [(w, s) for s in SOLVE_FORM_SEEDS for w in ("unreachable", "reachable")]:
                children, x0s, T = _solve_target(n, which, seed)
                opts = q.Options(target=T, threshold=1e-10)
                solver = solvers.LeastSquares_Jac_Solver(starts=4)
                solver.solve_batch(children[:2], opts, x0s[:2], want_matrices=False, frontier=frontier)  # warm-up: module load, buffers
                barrier(dist)
                t0 = time.perf_counter()
                res = solver.solve_batch(children, opts, x0s, want_matrices=False, frontier=frontier)
                dt = all_max(dist, time.perf_counter() - t0)
                tot_t += dt
                tot_inst += len(res.all_iterations)
                tot_it += int(res.all_iterations.sum())
                max_it += int(res.all_iterations.max())
                best.append(float(min(res.distance)))
            key = "frontier_rules" if frontier else "reference_semantics"
            out[key] = {"solves_per_s": world * tot_inst / tot_t, "instances": tot_inst, "wall_s": tot_t, "lm_iterations": tot_it,
                        "us_per_lm_iteration_latency": 1e6 * tot_t / max(1, max_it),
                        "instance_iterations_per_s": world * tot_it / tot_t, "best_distance": best}
        out["what"] = ("n={}: three seeded frontiers of 48 children x 4 starts, each once against qft (unreachable) and once against a reachable target, through "
                       "LeastSquares_Jac_Solver.solve_batch (host x0 in, xopt / distance out). reference_semantics: every start runs to "
                       "its own convergence, everything fp64; frontier_rules: straggler cap 2x + sibling stop + TF32 Gram matrix (gradient, objective and "
                       "outputs fp64), as GpuParallelizer runs it".format(n))
        if cpu_solve is not None:
            out["cpu"] = cpu_solve
            out["speedup_vs_cpu"] = out["reference_semantics"]["solves_per_s"] / cpu_solve["solves_per_s"]
        return out
    finally:
        L.use_library(prev)


def load_ncu_digest(n):
    """Hot-kernel ncu figures of the newest round that has them (profiles/r*/hot_kernel_ncu.json, written from one
    `ncu --set full` capture by tools/ncu_digest.py --json)."""
    import glob
    for path in sorted(glob.glob(os.path.join(ROOT, "profiles", "r*", "hot_kernel_ncu.json")), reverse=True):
        try:
            d = json.load(open(path))
            if str(n) in d:
                e = dict(d[str(n)])
                e["file"] = os.path.relpath(path, ROOT)
                return e
        except Exception:
            pass
    return None


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
                             "sample": "{} LM-form evals per step over {} forked workers (BLAS threads = 1). {}".format(per_step, procs, QSRS_CAVEAT)},
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
               "sample": "{} LM-form evals of the same workload in {:.1f} s on {} forked workers, oracle/qs_oracle.py. {}".format(
                   n_evals, dt, procs, QSRS_CAVEAT)}
    cpu_solve = None
    if rank == 0 and world == 1 and not args.no_cpu and not args.no_solve:
        cpu_solve = cpu_solve_arm(n, os.cpu_count() or 1)
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
    for _ in range(max(3, args.warmup)):  # page faults of the pinned output buffers, device buffers, host threads
        lib.eval_residuals_jac(gps, poi[sel], flat[sel], T, out_r, out_J)
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

    sf = None
    if not args.no_solve:
        sf = solve_form(lib, n, dist, world, cpu_solve)
    if rank != 0:
        return
    F = flops_lm(n, G, P)
    peaks = load_peaks()
    hbm_peak = peaks.get("hbm_gbs", 6650.0)
    evals_per_s_gpu = n_inst / (kern_ms * 1e-3)  # one GPU
    algorithmic_tf = evals_per_s_gpu * F / 1e12
    # Two roofs are close for this kernel: the HBM write of (r, J_r) and the FP64 pipe (DMMA + DFMA share
    # it).  `roofline` reports the BINDING one = the larger executed fraction; the other and the algorithmic-flop figure of
    # SURVEY 8d (which the kernel does not execute: rank-one derivative products, structured updates) are side keys.
    hbm_achieved = out_bytes / (kern_ms * 1e-3) / 1e9
    hbm_frac = hbm_achieved / hbm_peak
    ncu = load_ncu_digest(n)
    # executed FP64 flops per eval: DMMA (512 flop each) + DFMA (64) + DMUL/DADD (32), instruction counts from the ncu capture;
    # without a digest fall back to the tensor-pipe part the source fixes: (4 P + 8) N^3
    if ncu and ncu.get("fp64_flop_per_eval"):
        exec_flop, exec_src = float(ncu["fp64_flop_per_eval"]), ncu["file"]
    else:
        exec_flop, exec_src = (4.0 * P + 8.0) * (N ** 3), "source: (4 P + 8) N^3 DMMA flop per eval (no ncu digest found)"
    fp64_exec_tf = evals_per_s_gpu * exec_flop / 1e12
    fp64_peak_best = max(fp64_peak, dmma_peak)
    fp64_frac = fp64_exec_tf / fp64_peak_best if fp64_peak_best else 0.0
    if hbm_frac >= fp64_frac:
        roof = {"bound": "hbm", "achieved": hbm_achieved, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_frac,
                "peak_source": "MEASURED_PEAKS.json hbm_gbs (copy bandwidth, driver-measured)" if "hbm_gbs" in peaks else "fallback 6650 GB/s (B200_PROFILING.md)"}
    else:
        roof = {"bound": "tensor", "achieved": fp64_exec_tf, "peak": fp64_peak_best, "unit": "TFLOP/s", "frac": fp64_frac,
                "peak_source": "builder-measured FP64 pipe peak (qsb_measure_fp64_peak in this run; FP64 is not in MEASURED_PEAKS.json)"}
    roof.update({
        "traffic": int(ncu["dram_bytes_per_eval"] * n_inst) if ncu and ncu.get("dram_bytes_per_eval") else None,
        "algorithmic_bytes": int(out_bytes + n_inst * P * 8),
        "hbm": {"achieved": hbm_achieved, "peak": hbm_peak, "unit": "GB/s", "frac": hbm_frac, "bytes_per_eval": out_bytes / n_inst},
        "fp64_executed": {"achieved": fp64_exec_tf, "peak": fp64_peak_best, "unit": "TFLOP/s", "frac": fp64_frac,
                          "flop_per_eval": exec_flop, "source": exec_src, "dfma_peak": fp64_peak, "dmma_peak": dmma_peak,
                          "peak_source": "builder-measured in this run (qsb_measure_fp64_peak); FP64 is not in MEASURED_PEAKS.json"},
        "algorithmic": {"tflops": algorithmic_tf, "ratio_to_fp64_peak": algorithmic_tf / fp64_peak_best if fp64_peak_best else None,
                        "flop_per_eval": F,
                        "note": "SURVEY 8d dense prefix/suffix flop count; NOT a hardware fraction: the kernel executes "
                                "rank-one half-size products and 2x2-structured updates instead (parity-tested), so this exceeds 1"},
        "ncu": ncu,
    })
    line = {
        "metric": "objective+Jacobian evals/sec", "value": value, "unit": "evals/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": kern_ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": workload_config(args, n, G, P),
        "clocks": clocks, "gpu_launches": int(launches), "wall_s_timed_region": t_wall,
        "e2e": {"value": e2e_value, "unit": "evals/s", "h2d_bytes_per_step": int(e_inst * P * 8 + 2 * N * N * 8),
                "d2h_bytes_per_step": int(e_inst * (P + 1) * 2 * N * N * 8),
                "note": "qsb_eval_residuals_jac with host buffers, {} instances per step (r and J_r copied back: "
                        "PCIe-bound by the {:.2f} MB of Jacobian per eval); the call the search itself makes is solve_form".format(
                            e_inst, (P + 1) * 2 * N * N * 8 / 1e6)},
        "roofline": roof,
        "hs_form": {"value": world * n_inst / (hs_ms * 1e-3), "unit": "evals/s", "ms_per_step": hs_ms,
                    "e2e_value": world * n_inst / t_hs_e2e, "algorithmic_flop_per_eval": flops_hs(n, G, P)},
        "lm_system_form": {"value": world * n_sys / (sys_ms * 1e-3), "unit": "evals/s", "instances": n_sys,
                           "what": "r'r, J_r'r, J_r'J_r per instance (QSB_RUN_LMSYS, body-frame form for the unitary target: forward-only "
                                   "sweep + Gram product over N^2-real rows; nothing of size N^2 P leaves the SM's L1/L2)"},
        "parity_max_abs_err": worst,
    }
    if sf is not None:
        line["solve_form"] = sf
    if cpu is not None:
        line["cpu_baseline"] = cpu
        line["vs_cpu_baseline"] = {"device_resident": value / cpu["value"], "e2e": e2e_value / cpu["value"], "note": QSRS_CAVEAT}
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
    ap.add_argument("--no-solve", action="store_true", help="skip the solve_form key (frontier-shaped batched solves)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    if args.cpu_evals <= 0:  # 10-15 s of CPU work: ~120 (n=4) / ~75 (n=5) LM-form evals/s/core in NumPy
        args.cpu_evals = (1500 if args.qubits <= 4 else 900) * max(1, (os.cpu_count() or 1))
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
