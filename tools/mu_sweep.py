#!/usr/bin/env python
"""LM damping-schedule sweep for qsb_k_lm3: Jacobian evaluations, factorisation attempts, final distances
and wall time of frontier-shaped batches under the QSB_LM_MU_* knobs.  Run with QSB_LM_PROFILE=1 so that the library prints its
attempt counters (stderr) after every solve.

    QSB_LM_PROFILE=1 python tools/mu_sweep.py 2>&1 | grep -v "^\\[qsb lm profile\\] instances=1 "
"""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from qsearch_b200 import _lib, gatesets, unitaries  # noqa: E402
from tools.solve_bench import case, YORKTOWN  # noqa: E402

VARIANTS = [{"QSB_LM_MU_DOWN": "0.3333333333333333", "QSB_LM_MU_STREAK": "0"}, {}, {"QSB_LM_MU_FAST": "3"}, {"QSB_LM_MU_FAST": "4"},
            {"QSB_LM_MU_FAST": "6"}, {"QSB_LM_MU_FAST": "8"}, {"QSB_LM_MU_FAST": "4", "QSB_LM_MU_DOWN": "0.6"}]
KEYS = ("QSB_LM_MU_DOWN", "QSB_LM_MU_UP", "QSB_LM_MU_STREAK", "QSB_LM_MU_FAST")


def main():
    lib = _lib.get_library()
    cases = [("n3 d6 4x74", 3, 6, 4, 74, gatesets.QubitCNOTLinear(), 5), ("n4 d10 4x74", 4, 10, 4, 74, gatesets.QubitCNOTRing(), 5),
             ("n4 d8 3x98 reachable", 4, 8, 3, 98, gatesets.QubitCNOTLinear(), 6),
             ("n5 d12 6x24", 5, 12, 6, 24, gatesets.QubitCNOTAdjacencyList(YORKTOWN), 5)]
    from qsearch_b200 import workloads, lowering
    children, x0s, tg = workloads.frontier_workload(4)
    fw = ([lowering.lower(c) for c in children], np.repeat(np.arange(len(children), dtype="int32"), 4), np.concatenate([np.ravel(x) for x in x0s]))
    cases.append(("bench solve_form n4 48x4 unreachable", 4, 0, 0, 0, None, 0))
    for name, n, depth, nc, starts, gs, seed in cases:
        gps, poi, x0 = fw if gs is None else case(n, depth, nc, starts, seed, gs)
        if gs is None:
            T = tg["unreachable"]
        elif "reachable" in name:
            from qsearch_b200 import gates, lowering as _low  # a target the first child reaches exactly
            rng = np.random.default_rng(1)
            layers = gs.search_layers(n)
            c = gates.ProductGate(gs.initial_layer(n))
            rng2 = np.random.default_rng(seed)
            for _ in range(depth - 1):
                c = c.appending(layers[int(rng2.integers(len(layers)))][0])
            c = c.appending(layers[0][0])
            T = lib.eval_matrix([_low.lower(c)], [0], rng.random(c.num_inputs) * 2 * np.pi)[0]
        else:
            T = unitaries.qft(1 << n)
        lib.solve("lm", gps[:1], poi[:1] * 0, x0[:gps[0].n_params], T)
        for v in VARIANTS:
            for k in KEYS:
                os.environ.pop(k, None)
            os.environ.update(v)
            sys.stdout.flush()
            t0 = time.perf_counter()
            x, f, iters, status = lib.solve("lm", gps, poi, x0, T)
            dt = time.perf_counter() - t0
            print(json.dumps({"case": name, "variant": v, "wall_ms": round(dt * 1e3, 1), "jac_sum": int(iters.sum()), "jac_max": int(iters.max()),
                              "f_min": float(f.min()), "f_mean": float(f.mean()), "n_below_1e-10": int((f < 1e-10).sum()),
                              "status": {int(s): int((status == s).sum()) for s in np.unique(status)}}), flush=True)


if __name__ == "__main__":
    main()
