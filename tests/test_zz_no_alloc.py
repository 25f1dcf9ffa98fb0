"""GPU test of qsb_eval_lm_system_static, the allocation-free variant of QSB_RUN_LMSYS (include/qsb200.h: no counterpart in
the reference).  Collected last on purpose: a
failure here hides no other test.  It calls the new entry point through the C ABI and checks it against the existing entry
points (qsb_batch_run(QSB_RUN_LMSYS) + qsb_batch_fetch) on two calls in a row - the static scratch array keeps the first
call's contents - within the tolerance of the existing LMSYS GPU test (1e-11), and against the oracle."""
import numpy as np
import pytest

from oracle import qs_oracle as O
from qsearch_b200 import _lib, lowering, workloads

pytestmark = pytest.mark.gpu

TOL = 1e-11


def _existing(lib, gp, V, T):
    b = lib.batch([gp], np.zeros(len(V), dtype="int32"))
    b.set_params(V)
    b.set_target(T)
    b.run(_lib.RUN_LMSYS)
    rr, g, A = b.fetch()
    b.close()
    P = gp.n_params
    return rr, g.reshape(len(V), P), A.reshape(len(V), P, P)


@pytest.mark.parametrize("n", [3, 4, 5])
def test_static_entry_point_matches_the_existing_one_twice_in_a_row(n):
    lib = _lib.get_library()
    circuits, params, targets = workloads.microbench_workload(n, 2, 6, depth=6)
    T = targets["haar"]
    for rep, ci in enumerate((0, 1)):  # second call: another program and other vectors over the same array
        gp = lowering.lower(circuits[ci])
        V = np.ascontiguousarray(params[ci])
        rr, g, A = lib.eval_lm_system_static(gp, V, T)
        rr0, g0, A0 = _existing(lib, gp, V, T)
        assert np.abs(rr - rr0).max() < TOL and np.abs(g - g0).max() < TOL and np.abs(A - A0).max() < TOL, (n, rep)
        U, J = O.mat_jac(circuits[ci], V[0])
        r, Jr = O.matrix_residuals(T, U), O.matrix_residuals_jac(T, U, J)
        assert abs(rr[0] - r @ r) < TOL and np.abs(g[0] - Jr.T @ r).max() < TOL and np.abs(A[0] - Jr.T @ Jr).max() < TOL
    # the same call again gives the same bits (nothing left over from the previous call leaks in)
    rr2, g2, A2 = lib.eval_lm_system_static(gp, V, T)
    assert np.array_equal(rr, rr2) and np.array_equal(g, g2) and np.array_equal(A, A2)


def test_static_entry_point_rejects_what_does_not_fit():
    lib = _lib.get_library()
    circuits, params, targets = workloads.microbench_workload(4, 1, 1)
    gp = lowering.lower(circuits[0])  # P = 62: ~32 KB of staging per instance, 16 MiB limit
    with pytest.raises(_lib.QsbError, match="static limit"):
        lib.eval_lm_system_static(gp, np.zeros((1000, gp.n_params)), targets["qft"])
    with pytest.raises(_lib.QsbError, match="unitary"):
        lib.eval_lm_system_static(gp, np.zeros((2, gp.n_params)), 0.5 * targets["qft"])
