"""GPU end-to-end gate (-m gpu): BASELINE.json configs[0..2] - synthesis through SearchCompiler / LeapCompiler with the
batched on-device solvers.  Acceptance as in the reference (tests/conftest.py:10-18, tests/test_solvers.py:70,78):
final distance < 1e-10; CNOT counts against README.md:20-31 (best-of-N, because start vectors are random)."""
import numpy as np
import pytest

import qsearch_b200 as q
from oracle import qs_oracle as O
from qsearch_b200 import gates, gatesets, lowering, solvers, unitaries, _lib

pytestmark = pytest.mark.gpu


def cnots(circ):
    return repr(circ).count("CNOTGate")


def check(result, target):
    d = O.matrix_distance_squared(target, O.matrix(result["structure"], result["parameters"]))
    assert d < 1e-10, d
    return cnots(result["structure"])


@pytest.mark.parametrize("method", ["lm", "lbfgs"])
def test_batched_solvers_on_reachable_targets(method):
    lib = _lib.get_library()
    rng = np.random.default_rng(42)
    for n, depth in [(3, 4), (4, 6), (5, 5)]:
        g = gatesets.QubitCNOTRing()
        layers = g.search_layers(n)
        c = gates.ProductGate(g.initial_layer(n))
        for _ in range(depth):
            c = c.appending(layers[int(rng.integers(len(layers)))][0])
        T = O.matrix(c, rng.random(c.num_inputs) * 2 * np.pi)
        gp = lowering.lower(c)
        K = 16
        x0 = rng.random((K, gp.n_params)) * 2 * np.pi
        x, f, iters, status = lib.solve(method, [gp], np.zeros(K, dtype="int32"), x0, T)
        assert f.min() < 1e-10, (n, method, f)
        k = int(np.argmin(f))
        assert abs(O.matrix_distance_squared(T, O.matrix(c, x.reshape(K, -1)[k])) - f[k]) < 1e-12


def test_qft3_linear_search_compiler():
    """configs[0]: 3-qubit QFT, linear topology, CNOT+U3 gateset (README.md:20: 7 CNOTs, best of 10, sometimes +1)."""
    T = unitaries.qft(8)
    best = min(check(q.SearchCompiler().compile(q.Options(target=T, seed=s)), T) for s in range(3))
    assert best <= 8


@pytest.mark.parametrize("name,lin,ring", [("toffoli", 8, 6), ("fredkin", 8, 7), ("peres", 7, 6), ("logical_or", 8, 6)])
def test_readme_three_qubit_set(name, lin, ring):
    """configs[1]: README.md:21-25 CNOT counts on linear and ring topologies."""
    T = getattr(unitaries, name)
    for gs, expect in ((gatesets.QubitCNOTLinear(), lin), (gatesets.QubitCNOTRing(), ring)):
        got = check(q.SearchCompiler().compile(q.Options(target=T, gateset=gs, seed=7)), T)
        assert got <= expect + 1, (name, type(gs).__name__, got, expect)


def test_multistart_solver_in_the_search():
    T = unitaries.qft(8)
    opts = q.Options(target=T, solver=solvers.MultiStart_Solver(4), seed=5)
    assert check(q.SearchCompiler().compile(opts), T) <= 8


def test_qft4_leap():
    """configs[2]: examples/leap_qft4.py - 4-qubit QFT through LeapCompiler with LeastSquares_Jac_Solver + multistart."""
    T = unitaries.qft(16)
    opts = q.Options(target=T, gateset=gatesets.QubitCNOTLinear(), solver=solvers.LeastSquares_Jac_Solver(starts=4),
                     min_depth=3, seed=11)
    res = q.LeapCompiler().compile(opts)
    n = check(res, T)
    assert res["cut_depths"][-1] == n and n <= 30  # README.md:31: 13 with the plain search; LEAP trades length for time


def test_leap_reoptimizing_post_processor_on_the_device():
    """post_processing.py:108-295 on the batched device solver: a padded 3-qubit LEAP-style result (Toffoli solved on a
    12-layer circuit; the README count is 8) gets shorter and still implements the target."""
    from qsearch_b200 import post_processing
    T = unitaries.toffoli
    g = gatesets.QubitCNOTLinear()
    layers = g.search_layers(3)
    circ = gates.ProductGate(g.initial_layer(3))
    for k in range(12):
        circ = circ.appending(layers[k % len(layers)][0])
    opts = q.Options(target=T, seed=4, gateset=g)
    opts.set_defaults(**q.defaults.standard_defaults)
    opts.set_smart_defaults(**q.defaults.standard_smart_defaults)
    res = solvers.LeastSquares_Jac_Solver(starts=16).solve_batch([circ], opts)
    assert res.distance[0] < 1e-10
    padded = {"structure": circ, "parameters": res.x[0], "cut_depths": [0, 4, 8, 12]}
    out = post_processing.LEAPReoptimizing_PostProcessor().post_process_circuit(
        padded, q.Options(target=T, seed=4, gateset=g, reoptimize_size=5))
    assert cnots(out["structure"]) < 12
    assert O.matrix_distance_squared(T, O.matrix(out["structure"], out["parameters"])) < 1e-10
