#!/usr/bin/env python
"""Latency / throughput of the batched on-device solvers on frontier-shaped batches.

    python tools/solve_bench.py [--method lm|lbfgs]

Prints, per case: wall ms of one qsb_solve_* call (host buffers), total optimiser iterations, microseconds per
iteration of the slowest instance (latency view) and instance-iterations per second (throughput view).
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from qsearch_b200 import _lib, gates, gatesets, lowering, unitaries  # noqa: E402

YORKTOWN = [(0, 1), (0, 2), (1, 2), (2, 3), (2, 4), (3, 4)]


def case(n, depth, n_children, starts, seed, gateset):
    rng = np.random.default_rng(seed)
    layers = gateset.search_layers(n)
    base = gates.ProductGate(gateset.initial_layer(n))
    for _ in range(depth - 1):
        base = base.appending(layers[int(rng.integers(len(layers)))][0])
    children = [base.appending(layers[i % len(layers)][0]) for i in range(n_children)]
    gps = [lowering.lower(c) for c in children]
    poi = np.repeat(np.arange(n_children, dtype="int32"), starts)
    x0 = np.concatenate([rng.random(gps[i].n_params) * 2 * np.pi for i in poi])
    return gps, poi, x0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--method", default="lm")
    ap.add_argument("--only", default=None, help="run only the case with this name (for ncu captures)")
    ap.add_argument("--reps", type=int, default=1, help="identical repetitions per case (run-to-run spread)")
    ap.add_argument("--straggler", type=float, default=0.0, help="qsb_solver_opts.straggler_factor")
    ap.add_argument("--sibling", type=float, default=0.0, help="qsb_solver_opts.sibling_stop")
    ap.add_argument("--tf32", type=float, default=0.0, help="qsb_solver_opts.gram_tf32")
    a = ap.parse_args()
    lib = _lib.get_library()
    cases = [("n3 d6 2x4", 3, 6, 2, 4, gatesets.QubitCNOTLinear()),
             ("n4 d8 3x4", 4, 8, 3, 4, gatesets.QubitCNOTLinear()),
             ("n4 d16 3x4", 4, 16, 3, 4, gatesets.QubitCNOTLinear()),
             ("n4 d10 4x250", 4, 10, 4, 250, gatesets.QubitCNOTRing()),
             ("n5 d12 6x4", 5, 12, 6, 4, gatesets.QubitCNOTAdjacencyList(YORKTOWN)),
             ("n5 d30 6x4", 5, 30, 6, 4, gatesets.QubitCNOTAdjacencyList(YORKTOWN)),
             ("n5 d30 6x24", 5, 30, 6, 24, gatesets.QubitCNOTAdjacencyList(YORKTOWN))]
    for name, n, depth, nc, starts, gs in cases:
        if a.only and name != a.only:
            continue
        gps, poi, x0 = case(n, depth, nc, starts, 5, gs)
        T = unitaries.qft(1 << n)
        lib.solve(a.method, gps[:1], poi[:1] * 0, x0[:gps[0].n_params], T)  # warm-up (module load, allocations)
        kw = dict(straggler_factor=a.straggler, sibling_stop=a.sibling, gram_tf32=a.tf32) if a.method == "lm" else {}
        walls = []
        for _ in range(a.reps):
            t0 = time.perf_counter()
            x, f, iters, status = lib.solve(a.method, gps, poi, x0, T, **kw)
            walls.append(time.perf_counter() - t0)
        dt = min(walls)
        print(json.dumps({"case": name, "method": a.method, "kernel": os.environ.get("QSB_LM_KERNEL", "auto"),
                          "P": gps[0].n_params, "instances": len(poi), "wall_ms": round(dt * 1e3, 2),
                          "wall_ms_all": [round(w * 1e3, 2) for w in walls],
                          "status_counts": {int(k): int(v) for k, v in zip(*np.unique(status, return_counts=True))},
                          "iters_max": int(iters.max()), "iters_sum": int(iters.sum()),
                          "us_per_iter_latency": round(dt * 1e6 / max(1, int(iters.max())), 1),
                          "inst_iters_per_s": round(iters.sum() / dt, 1), "best_f": float(f.min())}), flush=True)


if __name__ == "__main__":
    main()
