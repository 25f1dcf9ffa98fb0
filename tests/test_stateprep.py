"""State-preparation objective (reference: qsearch/objectives.py:63-101, comparison.py:114-123).

CPU: the oracle's restatement against the reference-generated fixture tests/golden/stateprep.npz, and the product's
StateprepObjective closures on the host emulation of the kernels against the same fixture.
GPU (-m gpu): the same fixture through libqsb200.so, a batched parity check against the oracle, and synthesis of the
three target states of the reference's tests/test_stateprep.py:35-39 (SearchCompiler and LeapCompiler).
Tolerance: north_star's 1e-12 max-abs."""
import os

import numpy as np
import pytest

import qsearch_b200 as q
from oracle import qs_oracle as O
from qsearch_b200 import _lib, defaults, objectives, solvers
from tests import _golden

TOL = 1e-12

STATES = [np.array(s, dtype="complex128") for s in (
    [1 / np.sqrt(2), -1 / np.sqrt(2)],
    [0.5, 0.5, 0.5, 0, 0, 0, 0, 0.5],
    [0, 0.5, 0.5j, 0, -0.5, 0, 0, 0, -0.5j, 0, 0, 0, 0, 0, 0, 0])]


def _cases():
    blob = np.load(os.path.join(_golden.GOLDEN_DIR, "stateprep.npz"))
    for i in range(int(blob["n_cases"])):
        pre = "c{}_".format(i)
        yield (_golden.rebuild(str(blob[pre + "repr"])), blob[pre + "x"], blob[pre + "target_state"],
               blob[pre + "initial_state"], blob[pre + "target"],
               [{k: blob["{}s{}_{}".format(pre, s, k)] for k in ("dist", "f", "g")} for s in range(len(blob[pre + "x"]))])


def test_oracle_matches_reference_fixture():
    n = 0
    for circ, xs, a, b, A, gold in _cases():
        for x, gd in zip(xs, gold):
            U, J = O.mat_jac(circ, x)
            assert abs(O.distance_with_initial_state(a, b, A, O.matrix(circ, x)) - gd["dist"]) < 1e-13
            f, g = O.distance_with_initial_state_jac(a, b, A, U, J)
            assert abs(f - gd["f"]) < 1e-13 and np.abs(g - gd["g"]).max() < 1e-13
            n += 1
    assert n == 24


def _check_closures_against_fixture(max_qubits):
    obj = objectives.StateprepObjective()
    n = 0
    for circ, xs, a, b, A, gold in _cases():
        if circ.qudits > max_qubits:
            continue
        opts = q.Options(target_state=a, initial_state=b, target=A, objective=obj)
        err, jac = obj.gen_error_func(circ, opts), obj.gen_error_jac(circ, opts)
        assert obj.gen_error_residuals(circ, opts) is None and obj.gen_error_residuals_jac(circ, opts) is None
        for x, gd in zip(xs, gold):
            assert abs(err(x) - gd["dist"]) < TOL
            f, g = jac(x)
            assert abs(f - gd["f"]) < TOL and np.abs(g - gd["g"]).max() < TOL
            n += 1
    return n


@pytest.fixture()
def emu():
    from tests.emu import emulated_library
    prev = _lib.use_library(emulated_library())
    yield
    _lib.use_library(prev)


def test_emulated_closures_match_reference_fixture(emu):
    assert _check_closures_against_fixture(max_qubits=4) == 20


def test_defaults_and_solver_choice():
    opts = q.Options(target_state=STATES[1], defaults=defaults.stateprep_defaults,
                     smart_defaults=defaults.stateprep_smart_defaults)
    assert np.array_equal(opts.target, np.eye(8)) and opts.initial_state[0] == 1 and not opts.initial_state[1:].any()
    assert isinstance(solvers.default_solver(opts), solvers.BFGS_Jac_Solver)
    assert isinstance(solvers.default_solver(q.Options(target=np.eye(4), objective=objectives.MatrixDistanceObjective())),
                      solvers.LeastSquares_Jac_Solver)
    with pytest.raises(ValueError):  # no residual form: the least-squares solver refuses instead of ignoring the objective
        solvers.LeastSquares_Jac_Solver().solve_batch([q.gates.U3Gate()], opts.updated(target_state=STATES[0]))


def _stateprep_options(state, **kw):
    o = q.Options(target_state=state, defaults=defaults.stateprep_defaults,
                  smart_defaults=defaults.stateprep_smart_defaults, **kw)
    return o


def _overlap(result, state):
    U = O.matrix(result["structure"], result["parameters"])
    return abs(np.vdot(state, U[:, 0]))


def test_emulated_stateprep_search(emu):
    res = q.SearchCompiler().compile(_stateprep_options(STATES[0], seed=1))
    assert abs(_overlap(res, STATES[0]) - 1) < 1e-9
    bell = np.array([1, 0, 0, 1], dtype="complex128") / np.sqrt(2)
    res = q.SearchCompiler().compile(_stateprep_options(bell, seed=2))
    assert abs(_overlap(res, bell) - 1) < 1e-9 and repr(res["structure"]).count("CNOTGate") == 1


# ---------------------------------------------------------------------------------------------- B200
@pytest.mark.gpu
def test_gpu_closures_match_reference_fixture():
    assert _check_closures_against_fixture(max_qubits=5) == 24


@pytest.mark.gpu
def test_gpu_batched_stateprep_matches_oracle():
    """Ragged batch (three structures, many parameter vectors) of (f, grad f) against the oracle."""
    from qsearch_b200 import gatesets, lowering, workloads
    rng = np.random.default_rng(31)
    lib = _lib.get_library()
    for n in (3, 4, 5):
        N = 1 << n
        g = gatesets.QubitCNOTRing()
        layers = g.search_layers(n)
        circs = []
        for depth in (2, 5, 9):
            c = q.gates.ProductGate(g.initial_layer(n))
            for _ in range(depth):
                c = c.appending(layers[int(rng.integers(len(layers)))][0])
            circs.append(c)
        a = rng.standard_normal(N) + 1j * rng.standard_normal(N)
        a /= np.linalg.norm(a)
        b = rng.standard_normal(N) + 1j * rng.standard_normal(N)
        b /= np.linalg.norm(b)
        A = workloads.haar_unitary(N, 77 + n)
        opts = q.Options(target_state=a, initial_state=b, target=A)
        T = objectives.StateprepObjective().device_target(opts)
        gps = [lowering.lower(c) for c in circs]
        poi = np.array([0, 1, 1, 2, 0, 2, 2], dtype="int32")
        xs = [rng.random(gps[p].n_params) * 2 * np.pi for p in poi]
        f, grad = lib.eval_hs_grad(gps, poi, np.concatenate(xs), T)
        d = lib.eval_distance(gps, poi, np.concatenate(xs), T)
        off = 0
        for k, p in enumerate(poi):
            U, J = O.mat_jac(circs[p], xs[k])
            fo, go = O.distance_with_initial_state_jac(a, b, A, U, J)
            assert abs(f[k] - fo) < TOL and np.abs(grad[off:off + len(go)] - go).max() < TOL
            assert abs(d[k] - O.distance_with_initial_state(a, b, A, U)) < TOL
            off += len(go)


@pytest.mark.gpu
@pytest.mark.parametrize("which", [0, 1, 2])
@pytest.mark.parametrize("leap", [False, True])
def test_gpu_stateprep_synthesis(which, leap):
    """tests/test_stateprep.py:41-67 of the reference without the Qiskit round trip: the synthesised circuit applied to
    |0..0> must be the target state up to a global phase."""
    state = STATES[which]
    cls = q.LeapCompiler if leap else q.SearchCompiler
    res = cls().compile(_stateprep_options(state, seed=3 + which, timeout=120))
    assert res["distance"] < 1e-10
    assert abs(_overlap(res, state) - 1) < 1e-9
