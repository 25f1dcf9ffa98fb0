#!/usr/bin/env python
"""Host-buffer (end-to-end) calls of the C ABI, timed one by one: qsb_eval_residuals_jac into page-locked buffers and
qsb_eval_hs_grad, on the bench's 4-qubit workload.  QSB_LIB_PATH selects the library build (A/B runs).

    python tools/e2e_case.py [--qubits 4] [--instances 16384] [--reps 6]
"""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from qsearch_b200 import _lib, lowering, workloads  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--qubits", type=int, default=4)
    ap.add_argument("--instances", type=int, default=16384)
    ap.add_argument("--reps", type=int, default=6)
    a = ap.parse_args()
    n_circ = 100
    n_vec = (a.instances + n_circ - 1) // n_circ
    circuits, params, targets = workloads.microbench_workload(a.qubits, n_circ, n_vec)
    gps = [lowering.lower(c) for c in circuits]
    poi = np.repeat(np.arange(n_circ, dtype="int32"), n_vec)
    flat = params.reshape(len(poi), -1)
    lib = _lib.Library(path=os.environ.get("QSB_LIB_PATH"), devices=[0])
    N, P = 1 << a.qubits, gps[0].n_params
    out_r = lib.pinned_empty((len(poi), 2 * N * N))
    out_J = lib.pinned_empty((len(poi) * P, 2 * N * N))
    T = targets["qft"]
    lm, hs = [], []
    for _ in range(a.reps + 1):
        t0 = time.perf_counter()
        lib.eval_residuals_jac(gps, poi, flat, T, out_r, out_J)
        lm.append(time.perf_counter() - t0)
    for _ in range(a.reps + 1):
        t0 = time.perf_counter()
        lib.eval_hs_grad(gps, poi, flat, T)
        hs.append(time.perf_counter() - t0)
    print("n={} instances={} LM e2e evals/s: {}".format(a.qubits, len(poi), ["%.3e" % (len(poi) / t) for t in lm]))
    print("n={} instances={} HS e2e evals/s: {}".format(a.qubits, len(poi), ["%.3e" % (len(poi) / t) for t in hs]))


if __name__ == "__main__":
    main()
