#!/usr/bin/env python
"""Latency of the literal drop-in: one parameter vector per call, as the reference's scipy solvers call the backend.

    python tools/latency_bench.py

Prints microseconds per CudaCircuit.matrix(v) / mat_jac(v) (n = 3, 4, 5; depth-10 ring circuits) and the wall time of one
reference-style solve: scipy.optimize.least_squares(method='lm') driving the objective closures of
qsearch_b200.objectives (the call qsearch/solvers.py:155-157 makes), against the batched on-device solve of the same problem.
"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qsearch_b200 as q  # noqa: E402
from qsearch_b200 import backends, objectives, solvers, unitaries, workloads  # noqa: E402


def main():
    import scipy.optimize
    rng = np.random.default_rng(0)
    for n in (3, 4, 5):
        circ = workloads.microbench_circuits(n, 1, 10)[0][0]
        cc = backends.CudaBackend().prepare_circuit(circ)
        v = rng.random(cc.num_inputs) * 2 * np.pi
        for _ in range(20):
            cc.matrix(v); cc.mat_jac(v)
        reps = 300
        t0 = time.perf_counter()
        for _ in range(reps):
            cc.matrix(v)
        t1 = time.perf_counter()
        for _ in range(reps):
            cc.mat_jac(v)
        t2 = time.perf_counter()
        print(json.dumps({"n": n, "P": cc.num_inputs, "matrix_us": round(1e6 * (t1 - t0) / reps, 1),
                          "mat_jac_us": round(1e6 * (t2 - t1) / reps, 1)}), flush=True)
    # one reference-style solve of a reachable 3-qubit problem (QFT3 at its known depth 7, linear topology)
    g = q.gatesets.QubitCNOTLinear()
    layers = g.search_layers(3)
    circ = q.gates.ProductGate(g.initial_layer(3))
    for k in range(7):
        circ = circ.appending(layers[k % 2][0])
    T = unitaries.qft(8)
    opts = q.Options(target=T)
    obj = objectives.MatrixDistanceObjective()
    res_f, jac_f, eval_f = obj.gen_error_residuals(circ, opts), obj.gen_error_residuals_jac(circ, opts), obj.gen_eval_func(circ, opts)
    x0 = rng.random(circ.num_inputs) * 2 * np.pi
    t0 = time.perf_counter()
    r = scipy.optimize.least_squares(res_f, x0, jac_f, method="lm")
    t1 = time.perf_counter()
    print(json.dumps({"case": "scipy LM driving CudaBackend closures, 3q depth 7 (P = 44)", "wall_ms": round(1e3 * (t1 - t0), 1),
                      "nfev": int(r.nfev), "njev": int(r.njev), "us_per_call": round(1e6 * (t1 - t0) / (r.nfev + r.njev), 1),
                      "distance": float(eval_f(r.x))}), flush=True)
    s = solvers.LeastSquares_Jac_Solver()
    s.solve_batch([circ], opts, [x0], want_matrices=False)
    t0 = time.perf_counter()
    res = s.solve_batch([circ], opts, [x0], want_matrices=False)
    t1 = time.perf_counter()
    print(json.dumps({"case": "same problem, batched on-device LM (one instance)", "wall_ms": round(1e3 * (t1 - t0), 2),
                      "iterations": int(res.iterations[0]), "distance": float(res.distance[0])}), flush=True)


if __name__ == "__main__":
    main()
