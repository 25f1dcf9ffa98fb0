#!/usr/bin/env python
"""Throughput of the radix-3 interpreter kernels on the GPU box: objective+Jacobian evals/s (LM form and HS form) for
QutritCNOTLinear circuits (gatesets.py:295-305), inputs resident in HBM, timed with CUDA events by qsb_batch_run.

    python tools/qutrit_bench.py [--instances 20000] [--reps 5]
"""
import argparse
import json
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from qsearch_b200 import gates, gatesets, lowering, workloads, _lib  # noqa: E402
from qsearch_b200._lib import RUN_RESID, RUN_HS  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--instances", type=int, default=20000)
    ap.add_argument("--reps", type=int, default=5)
    a = ap.parse_args()
    lib = _lib.get_library()
    rng = np.random.default_rng(1)
    for n, depth in ((2, 3), (2, 8), (3, 4)):
        gs = gatesets.QutritCNOTLinear()
        layers = gs.search_layers(n)
        c = gates.ProductGate(gs.initial_layer(n))
        for k in range(depth):
            c = c.appending(layers[k % len(layers)][0])
        gp = lowering.lower(c)
        K = a.instances if n == 2 else a.instances // 4
        b = lib.batch([gp], np.zeros(K, dtype="int32"))
        try:
            b.set_params(rng.random((K, gp.n_params)) * 2 * np.pi)
            b.set_target(workloads.haar_unitary(gp.dim, 5))
            out = {"qutrits": n, "layers": depth, "N": gp.dim, "P": gp.n_params, "instances": K}
            for what, name in ((RUN_RESID, "lm_form"), (RUN_HS, "hs_form")):
                for _ in range(2):
                    b.run(what)
                b.sync()
                ms = [b.run(what, timed=True) for _ in range(a.reps)]
                out[name + "_evals_per_s"] = K / (np.median(ms) * 1e-3)
                out[name + "_ms"] = float(np.median(ms))
            out["lm_form_GBps_written"] = out["lm_form_evals_per_s"] * (gp.n_params + 1) * 2 * gp.dim ** 2 * 8 / 1e9
            print(json.dumps(out), flush=True)
        finally:
            b.close()


if __name__ == "__main__":
    main()
