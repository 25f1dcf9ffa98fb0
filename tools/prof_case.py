#!/usr/bin/env python
"""Small fixed case for ncu: a few launches of one hot-path kernel on a modest resident batch.

    python tools/prof_case.py --what resid --qubits 4 --instances 4736 --reps 3
"""
import argparse
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from qsearch_b200 import _lib, lowering, workloads  # noqa: E402

WHAT = {"matrix": _lib.RUN_MATRIX, "matjac": _lib.RUN_MATJAC, "hs": _lib.RUN_HS, "resid": _lib.RUN_RESID,
        "dist": _lib.RUN_DIST, "lmsys": _lib.RUN_LMSYS}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--what", default="resid")
    ap.add_argument("--qubits", type=int, default=4)
    ap.add_argument("--instances", type=int, default=4736)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--circuits", type=int, default=16)
    a = ap.parse_args()
    n_circ = a.circuits
    n_vec = (a.instances + n_circ - 1) // n_circ
    circuits, params, targets = workloads.microbench_workload(a.qubits, n_circ, n_vec)
    gps = [lowering.lower(c) for c in circuits]
    poi = np.repeat(np.arange(n_circ, dtype="int32"), n_vec)
    lib = _lib.Library(path=os.environ.get("QSB_LIB_PATH"), devices=[0])  # QSB_LIB_PATH: A/B builds of the library
    b = lib.batch(gps, poi)
    b.set_params(params.reshape(len(poi), -1))
    b.set_target(targets["qft"])
    ms = [b.run(WHAT[a.what], timed=True) for _ in range(a.reps)]
    print("what={} n={} instances={} ms={}".format(a.what, a.qubits, len(poi), ["%.3f" % m for m in ms]))
    print("evals/s = {:.3e}".format(len(poi) / (min(ms) * 1e-3)))
    b.close()


if __name__ == "__main__":
    main()
