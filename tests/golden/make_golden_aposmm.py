"""Golden vectors for the APOSMM start selection of MultiStart_Solver (qsearch/multistart_solvers.py:66-88), generated
from the reference's own functions (qsearch/persistent_aposmm.py: initialize_APOSMM :559, add_to_local_H :262,
update_history_dist :299, decide_where_to_start_localopt :421, calc_rk :544).  persistent_aposmm imports without qsrs.
Runs only in the build container; writes tests/golden/aposmm.npz.
"""
import os
import sys

import numpy as np

sys.path.insert(0, "/root/reference")
from qsearch.persistent_aposmm import (initialize_APOSMM, add_to_local_H, update_history_dist,  # noqa: E402
                                       decide_where_to_start_localopt)

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "aposmm.npz")


def reference_start_indices(sample, f):
    """multistart_solvers.py:66-88 verbatim, with the function values supplied."""
    n_s, n = sample.shape
    specs = {'lb': np.zeros(n), 'ub': np.ones(n), 'standalone': True, 'initial_sample_size': n_s}
    _, _, rk_const, ld, mu, nu, _, H = initialize_APOSMM([], specs, None)
    add_to_local_H(H, sample, specs, on_cube=True)
    for i in range(n_s):
        H['f'][i] = f[i]
    H[['returned']] = True
    update_history_dist(H, n)
    return np.asarray(decide_where_to_start_localopt(H, n, n_s, rk_const, ld, mu, nu), dtype="int64")


def main():
    rng = np.random.default_rng(20261105)
    out = {}
    k = 0
    for n in (1, 2, 3, 5, 9, 24, 62, 165):
        for rep in range(3 if n < 100 else 1):
            n_s = 100 if rep < 2 else 37
            sample = rng.uniform(0, 1, (n_s, n))
            if rep == 1:  # a multimodal landscape on the cube, so that the r_k rule excludes points in low dimension
                f = np.sum(np.sin(5 * sample) ** 2, axis=1) + 0.1 * rng.random(n_s)
            else:
                f = rng.random(n_s)
            if rep == 2:
                f[5] = f[9]            # a tie
                sample[3, 0] = 5e-5    # closer to the boundary than mu = 1e-4
            out["x_%d" % k] = sample
            out["f_%d" % k] = f
            out["start_%d" % k] = reference_start_indices(sample.copy(), f)
            k += 1
    out["n_cases"] = np.array(k)
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, k, "cases;", [len(out["start_%d" % i]) for i in range(k)])


if __name__ == "__main__":
    main()
