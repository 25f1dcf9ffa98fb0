"""Generate the state-preparation fixture FROM THE REFERENCE ITSELF (run in the build container only).

    python tests/golden/make_golden_stateprep.py

Evaluates the reference's StateprepObjective closures (qsearch/objectives.py:63-79 ->
comparison.distance_with_initial_state / _jac, comparison.py:114-123) on seeded inputs: the three target states of
the reference's own tests/test_stateprep.py:35-39 plus random states, with the default initial state |0..0> and target
matrix I (defaults.py:75-81) and with a random initial state / Haar target matrix.  Written to stateprep.npz.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

import qsearch  # noqa: E402  (the reference)
from qsearch import gates as rg, gatesets as rgs, objectives as ro, defaults as rd  # noqa: E402
from qsearch_b200 import workloads  # noqa: E402


def rand_state(rng, N):
    v = rng.standard_normal(N) + 1j * rng.standard_normal(N)
    return v / np.linalg.norm(v)


def main():
    rng = np.random.default_rng(777)
    states = [np.array([1 / np.sqrt(2), -1 / np.sqrt(2)], dtype="complex128"),
              np.array([0.5, 0.5, 0.5, 0, 0, 0, 0, 0.5], dtype="complex128"),
              np.array([0, 0.5, 0.5j, 0, -0.5, 0, 0, 0, -0.5j, 0, 0, 0, 0, 0, 0, 0], dtype="complex128"),
              rand_state(rng, 4), rand_state(rng, 8), rand_state(rng, 32)]
    obj = ro.StateprepObjective()
    blob, i = {}, 0
    for target_state in states:
        N = target_state.shape[0]
        n = int(np.log2(N))
        gs = rgs.QubitCNOTLinear() if n < 5 else rgs.QubitCNOTRing()
        circ = rg.ProductGate(gs.initial_layer(n))
        layers = gs.search_layers(n)
        for _ in range(0 if n == 1 else 3):
            circ = circ.appending(layers[int(rng.integers(len(layers)))][0])
        for variant in ("default", "general"):
            opts = qsearch.Options(target_state=target_state, defaults=rd.stateprep_defaults,
                                   smart_defaults=rd.stateprep_smart_defaults)
            if variant == "general":
                opts.initial_state = rand_state(rng, N)
                opts.target = workloads.haar_unitary(N, 4000 + N)
            xs = rng.random((2, circ.num_inputs)) * 2 * np.pi
            pre = "c{}_".format(i)
            blob[pre + "repr"] = np.array(repr(circ))
            blob[pre + "x"] = xs
            blob[pre + "target_state"] = target_state
            blob[pre + "initial_state"] = np.asarray(opts.initial_state)
            blob[pre + "target"] = np.asarray(opts.target)
            assert obj.gen_error_residuals(circ, opts) is None and obj.gen_error_residuals_jac(circ, opts) is None
            for s, x in enumerate(xs):
                blob["{}s{}_dist".format(pre, s)] = np.float64(obj.gen_error_func(circ, opts)(x))
                f, g = obj.gen_error_jac(circ, opts)(x)
                blob["{}s{}_f".format(pre, s)] = np.float64(f)
                blob["{}s{}_g".format(pre, s)] = np.asarray(g, dtype="float64")
            i += 1
    blob["n_cases"] = np.array(i)
    path = os.path.join(HERE, "stateprep.npz")
    np.savez_compressed(path, **blob)
    print(path, os.path.getsize(path) // 1024, "KiB,", i, "cases")


if __name__ == "__main__":
    main()
