#!/usr/bin/env python
"""End-to-end synthesis timings on the GPU box (BASELINE.json configs 0,1,2,4): wall seconds, CNOT counts, distances.

    python tools/synth_bench.py [--cases qft3,toffoli,...] [--reps 3] [--starts 4] [--solver lm|lbfgs]
"""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import qsearch_b200 as q  # noqa: E402
from qsearch_b200 import gates, gatesets, solvers, unitaries, utils, _lib  # noqa: E402

YORKTOWN = [(0, 1), (0, 2), (1, 2), (2, 3), (2, 4), (3, 4)]  # examples/qft5_ibm_yorktown.py:21

CASES = {
    "qft3": (lambda: unitaries.qft(8), gatesets.QubitCNOTLinear, "search"),
    "toffoli": (lambda: unitaries.toffoli, gatesets.QubitCNOTLinear, "search"),
    "fredkin": (lambda: unitaries.fredkin, gatesets.QubitCNOTLinear, "search"),
    "peres": (lambda: unitaries.peres, gatesets.QubitCNOTLinear, "search"),
    "or": (lambda: unitaries.logical_or, gatesets.QubitCNOTLinear, "search"),
    "toffoli_ring": (lambda: unitaries.toffoli, gatesets.QubitCNOTRing, "search"),
    "qft3_ring": (lambda: unitaries.qft(8), gatesets.QubitCNOTRing, "search"),
    "fredkin_ring": (lambda: unitaries.fredkin, gatesets.QubitCNOTRing, "search"),
    "peres_ring": (lambda: unitaries.peres, gatesets.QubitCNOTRing, "search"),
    "or_ring": (lambda: unitaries.logical_or, gatesets.QubitCNOTRing, "search"),
    "qft4_leap": (lambda: unitaries.qft(16), gatesets.QubitCNOTLinear, "leap"),
    "qft4_search": (lambda: unitaries.qft(16), gatesets.QubitCNOTLinear, "search"),
    "qft5_yorktown_leap": (lambda: unitaries.qft(32), lambda: gatesets.QubitCNOTAdjacencyList(YORKTOWN), "leap"),
    # two qutrits (examples/qutrit_csum.py): "cnots" counts the two-qutrit gates of the result
    "qt_csum": (lambda: gates.CSUMGate().const_matrix(), gatesets.QutritCNOTLinear, "search"),
    "qt_bswap": (lambda: utils.upgrade_qudits(unitaries.swap), gatesets.QutritCNOTLinear, "search"),
    "qt_tswap": (lambda: unitaries.general_swap(3), gatesets.QutritCNOTLinear, "search"),
    "qt_qft2": (lambda: unitaries.qft(9), gatesets.QutritCNOTLinear, "search"),
    "qt_csum_cpi": (lambda: gates.CSUMGate().const_matrix(), gatesets.QutritCPIPhaseLinear, "search"),
}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", default="qft3,toffoli,qft4_leap")
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--starts", type=int, default=1)
    ap.add_argument("--solver", default="lm")
    ap.add_argument("--min-depth", type=int, default=3)
    ap.add_argument("--timeout", type=float, default=600)
    ap.add_argument("--beams", type=int, default=-1)
    ap.add_argument("--fstarts", type=int, default=0, help="options.frontier_starts (0: sized to the GPUs by the parallelizer)")
    ap.add_argument("--straggler", type=float, default=-1.0, help="LM straggler_factor for frontier batches (-1: solver default)")
    ap.add_argument("--seed0", type=int, default=1000)
    ap.add_argument("--tf32", type=int, default=-1, help="LM Gram product on the TF32 path for frontier batches (-1: solver default = on, 0: off)")
    ap.add_argument("--reopt", type=int, default=0, help="run LEAPReoptimizing_PostProcessor with this window size on LEAP results")
    a = ap.parse_args()
    lib = _lib.get_library()
    print("library:", lib.version(), "devices:", lib.num_devices, flush=True)
    for name in a.cases.split(","):
        mk_t, mk_gs, kind = CASES[name]
        T = mk_t()
        for rep in range(a.reps):
            if a.solver == "lm":
                solver = solvers.LeastSquares_Jac_Solver(starts=a.starts, straggler_factor=None if a.straggler < 0 else a.straggler,
                                                         gram_tf32=None if a.tf32 < 0 else bool(a.tf32))
            else:
                solver = solvers.BFGS_Jac_Solver(starts=a.starts)
            opts = q.Options(target=T, gateset=mk_gs(), solver=solver, seed=a.seed0 + rep, timeout=a.timeout)
            if a.fstarts > 0:
                opts.frontier_starts = a.fstarts
            if a.beams > 0:
                opts.beams = a.beams
            if kind == "leap":
                opts.min_depth = a.min_depth
            comp = q.LeapCompiler() if kind == "leap" else q.SearchCompiler()
            l0 = lib.launch_count()
            t0 = time.perf_counter()
            res = comp.compile(opts)
            dt = time.perf_counter() - t0
            extra = {}
            if kind == "leap" and a.reopt > 0:
                t1 = time.perf_counter()
                post = q.post_processing.LEAPReoptimizing_PostProcessor().post_process_circuit(
                    res, opts.updated(reoptimize_size=a.reopt, timeout=a.timeout))
                extra = {"reopt_window": a.reopt, "reopt_wall_s": round(time.perf_counter() - t1, 3),
                         "cnots_before": repr(res["structure"]).count("CNOTGate"), "reopt_stats": post.get("stats")}
                res = dict(res, structure=post["structure"], parameters=post["parameters"], distance=post.get("distance", res["distance"]))
            print(json.dumps({"case": name, "rep": rep, "solver": a.solver, "starts": a.starts, "fstarts": a.fstarts, "straggler": a.straggler, "tf32": a.tf32, "beams": a.beams, "devices": lib.num_devices, "wall_s": round(dt, 3),
                              "cnots": repr(res["structure"]).count("CNOTGate") + repr(res["structure"]).count("CPIPhaseGate"), "distance": float(res["distance"]),
                              "stats": res["stats"], "launches": lib.launch_count() - l0, **extra}), flush=True)


if __name__ == "__main__":
    main()
