"""CPU suite: the batched on-device solvers and the search host code, run on the host emulation of the kernels."""
import pickle

import numpy as np
import pytest

import qsearch_b200 as q
from oracle import qs_oracle as O
from qsearch_b200 import _lib, gates, gatesets, lowering, solvers, unitaries
from tests.emu import emulated_library


@pytest.fixture(scope="module")
def emu():
    lib = emulated_library()
    prev = _lib.use_library(lib)
    yield lib
    _lib.use_library(prev)


def _reachable_problem(n, depth, seed):
    rng = np.random.default_rng(seed)
    g = gatesets.QubitCNOTLinear()
    c = gates.ProductGate(g.initial_layer(n))
    layers = g.search_layers(n)
    for _ in range(depth):
        c = c.appending(layers[int(rng.integers(len(layers)))][0])
    return c, O.matrix(c, rng.random(c.num_inputs) * 2 * np.pi), rng


@pytest.mark.parametrize("method", ["lm", "lbfgs"])
def test_solvers_reach_a_reachable_target(emu, method):
    c, T, rng = _reachable_problem(2, 3, 11)
    gp = lowering.lower(c)
    x0 = rng.random((4, gp.n_params)) * 2 * np.pi
    x, f, iters, status = emu.solve(method, [gp], np.zeros(4, dtype="int32"), x0, T)
    assert f.min() < 1e-12
    for k, xk in enumerate(x.reshape(4, -1)):  # the reported distance is the distance at the returned parameters
        assert abs(O.matrix_distance_squared(T, O.matrix(c, xk)) - f[k]) < 1e-12
    assert (iters > 0).all()


def test_solver_interface_and_pickle(emu):
    c, T, _ = _reachable_problem(2, 2, 5)
    opts = q.Options(target=T, seed=3)
    opts.set_defaults(**q.defaults.standard_defaults)
    for s in (solvers.LeastSquares_Jac_Solver(starts=3), solvers.BFGS_Jac_Solver(starts=3),
              solvers.NaiveMultiStart_Solver(3), solvers.MultiStart_Solver(3)):
        s2 = pickle.loads(pickle.dumps(s))
        assert s2 == s
        U, x = s2.solve_for_unitary(q.backends.CudaBackend().prepare_circuit(c), opts)
        assert U.shape == T.shape and x.shape == (c.num_inputs,)
        assert O.matrix_distance_squared(T, U) < 1e-10
        assert np.abs(U - O.matrix(c, x)).max() < 1e-12
    assert solvers.LeastSquares_Jac_Solver().distance_metric == "Residuals"
    assert solvers.BFGS_Jac_Solver().distance_metric == "Frobenius"
    b = pickle.loads(pickle.dumps(q.backends.CudaBackend()))
    circ = pickle.loads(pickle.dumps(b.prepare_circuit(c)))
    assert circ.num_inputs == c.num_inputs


def test_objective_closures_match_oracle(emu):
    c, T, rng = _reachable_problem(3, 2, 9)
    opts = q.Options(target=T)
    obj = q.objectives.MatrixDistanceObjective()
    x = rng.random(c.num_inputs) * 2 * np.pi
    U, J = O.mat_jac(c, x)
    assert abs(obj.gen_error_func(c, opts)(x) - O.matrix_distance_squared(T, U)) < 1e-12
    f, g = obj.gen_error_jac(c, opts)(x)
    fo, go = O.matrix_distance_squared_jac(T, U, J)
    assert abs(f - fo) < 1e-12 and np.abs(g - go).max() < 1e-12
    assert np.abs(obj.gen_error_residuals(c, opts)(x) - O.matrix_residuals(T, U)).max() < 1e-12
    Jr = obj.gen_error_residuals_jac(c, opts)(x)
    assert Jr.shape == (2 * 64, c.num_inputs) and np.abs(Jr - O.matrix_residuals_jac(T, U, J)).max() < 1e-12


def test_search_compiler_qft2(emu):
    T = unitaries.qft(4)
    res = q.SearchCompiler().compile(q.Options(target=T, seed=1))
    assert repr(res["structure"]).count("CNOTGate") == 3  # tests/test_solvers.py compiles qft(4); reference finds 3 CNOTs
    assert O.matrix_distance_squared(T, O.matrix(res["structure"], res["parameters"])) < 1e-10


def test_leap_compiler_qft2(emu):
    T = unitaries.qft(4)
    res = q.LeapCompiler().compile(q.Options(target=T, seed=2, min_depth=1))
    assert O.matrix_distance_squared(T, O.matrix(res["structure"], res["parameters"])) < 1e-10
    assert res["cut_depths"][0] == 0 and res["cut_depths"][-1] == repr(res["structure"]).count("CNOTGate")


def test_options_semantics():
    o = q.Options(a=1)
    o.set_defaults(b=2, c=3)
    o.set_smart_defaults(c=lambda opts: opts.a + opts.b)
    assert (o.a, o.b, o.c) == (1, 2, 3) and "c" in o and "zzz" not in o
    o2 = o.updated(a=10)
    assert o2.c == 12 and o.c == 3
    with pytest.raises(AttributeError):
        o.make_required("target")
    assert pickle.loads(pickle.dumps(q.Options(a=1, b=[1, 2]))).b == [1, 2]


@pytest.mark.reference
def test_drop_in_for_the_reference_compiler(emu):
    """The GPU Backend / Solver / Parallelizer plugged into the reference's own SearchCompiler and Options."""
    import sys
    sys.path.insert(0, "/root/reference")
    import qsearch
    from qsearch import unitaries as ru
    opts = qsearch.Options(target=ru.qft(4), backend=q.backends.CudaBackend(), solver=solvers.LeastSquares_Jac_Solver(),
                           parallelizer=q.parallelizers.GpuParallelizer, verbosity=0, stdout_enabled=False)
    res = qsearch.SearchCompiler().compile(opts)
    assert repr(res["structure"]).count("CNOTGate") == 3
    assert qsearch.comparison.matrix_distance_squared(ru.qft(4), res["structure"].matrix(res["parameters"])) < 1e-10
    # and through the reference's evaluate_step with its sequential parallelizer
    opts2 = opts.updated(parallelizer=qsearch.parallelizers.SequentialParallelizer)
    res2 = qsearch.SearchCompiler().compile(opts2)
    assert qsearch.comparison.matrix_distance_squared(ru.qft(4), res2["structure"].matrix(res2["parameters"])) < 1e-10


def test_tensor_core_lm_on_three_qubits(emu):
    """qsb_k_lm2 (segment-parallel DMMA Jacobian + DMMA SYRK) in emulation: reaches a reachable 3-qubit target and
    reports the distance at the returned parameters."""
    c, T, rng = _reachable_problem(3, 3, 21)
    gp = lowering.lower(c)
    x0 = rng.random((3, gp.n_params)) * 2 * np.pi
    x, f, iters, status = emu.solve("lm", [gp], np.zeros(3, dtype="int32"), x0, T)
    assert f.min() < 1e-12
    for k, xk in enumerate(x.reshape(3, -1)):
        assert abs(O.matrix_distance_squared(T, O.matrix(c, xk)) - f[k]) < 1e-12


def test_leap_reoptimizing_post_processor(emu):
    """post_processing.py:108-295 on the batched solver: a deliberately padded 2-qubit LEAP-style result gets shorter
    and still implements the target."""
    from qsearch_b200 import post_processing
    T = unitaries.qft(4)
    g = gatesets.QubitCNOTLinear()
    circ = gates.ProductGate(g.initial_layer(2))
    for _ in range(6):  # qft(4) needs 3 CNOTs; build a 6-layer circuit and solve it
        circ = circ.appending(g.search_layers(2)[0][0])
    opts = q.Options(target=T, seed=4)
    opts.set_defaults(**q.defaults.standard_defaults)
    opts.set_smart_defaults(**q.defaults.standard_smart_defaults)
    res = solvers.LeastSquares_Jac_Solver(starts=4).solve_batch([circ], opts)
    assert res.distance[0] < 1e-10
    padded = {"structure": circ, "parameters": res.x[0], "cut_depths": [0, 3, 6]}
    out = post_processing.LEAPReoptimizing_PostProcessor().post_process_circuit(padded, q.Options(target=T, seed=4, reoptimize_size=4))
    n_after = repr(out["structure"]).count("CNOTGate")
    assert n_after < 6
    assert O.matrix_distance_squared(T, O.matrix(out["structure"], out["parameters"])) < 1e-10


def test_frontier_sizing_and_stop_rule_plumbing(emu):
    """GpuParallelizer sizes one expansion to the SMs of the context's devices (role of num_tasks = worker count,
    qsearch/parallelizers.py:74, compiler.py:89-93) and only frontier batches get the batch-level stop rules."""
    from qsearch_b200 import parallelizers
    T = unitaries.qft(8)
    opts = q.Options(target=T, gateset=gatesets.QubitCNOTLinear())
    opts.set_defaults(**q.defaults.standard_defaults)
    opts.set_smart_defaults(**q.defaults.standard_smart_defaults)
    ndev = emu.num_devices
    assert parallelizers.default_num_tasks(opts) == 2 * parallelizers.MAX_BEAMS  # branching 2 on a 3-qubit line, beams capped
    assert parallelizers.frontier_starts(opts, 16) == min(parallelizers.MAX_STARTS, 148 * ndev // 16)
    assert parallelizers.frontier_starts(opts, 100000) == 1
    assert parallelizers.frontier_starts(opts.updated(frontier_starts=5), 16) == 5
    assert parallelizers.frontier_starts(opts.updated(solver=solvers.LeastSquares_Jac_Solver(starts=4)), 16) is None
    s = solvers.LeastSquares_Jac_Solver()
    plain, front = s._solve_opts(opts), s._solve_opts(opts, frontier=True)
    assert plain["straggler_factor"] == 0.0 and plain["sibling_stop"] == 0.0
    assert front["straggler_factor"] == s.FRONTIER_STRAGGLER and front["sibling_stop"] == opts.threshold
    assert plain["gram_tf32"] == 0.0 and front["gram_tf32"] == 1.0
    assert solvers.LeastSquares_Jac_Solver(gram_tf32=False)._solve_opts(opts, frontier=True)["gram_tf32"] == 0.0
    assert solvers.LeastSquares_Jac_Solver(straggler_factor=0.0)._solve_opts(opts, frontier=True)["straggler_factor"] == 0.0


def test_qutrit_search_and_multistart(emu):
    """d = 3 through the host code of the search (gatesets.py:283-305, compiler.py:70-78): the qudit count comes from
    log_d of the target's size, the gateset's layers lower to radix-3 programs, SearchCompiler finds a one-layer two-qutrit
    target with one two-qutrit gate, and MultiStart_Solver solves a single-qutrit gate."""
    g = gatesets.QutritCNOTLinear()
    assert g.d == 3 and gatesets.DefaultQutrit is gatesets.QutritCPIPhaseLinear
    c = gates.ProductGate(g.initial_layer(2)).appending(g.search_layers(2)[0][0])
    T = O.matrix(c, np.random.default_rng(8).random(c.num_inputs) * 2 * np.pi)
    res = q.SearchCompiler().compile(q.Options(target=T, gateset=g, seed=5, timeout=300))
    assert O.matrix_distance_squared(T, O.matrix(res["structure"], res["parameters"])) < 1e-10
    assert repr(res["structure"]).count("UpgradedConstantGate") == 1
    with pytest.raises(ValueError):  # an 8 x 8 target is not a whole number of qutrits (compiler.py:72-73)
        q.SearchCompiler().compile(q.Options(target=unitaries.qft(8), gateset=g))
    sq = gates.SingleQutritGate()
    T1 = O.matrix(sq, np.random.default_rng(1).random(8) * 6)
    opts = q.Options(target=T1, gateset=g, solver=solvers.MultiStart_Solver(2), seed=3)
    U, x = opts.solver.solve_for_unitary(sq, opts)
    assert O.matrix_distance_squared(T1, O.matrix(sq, x)) < 1e-10 and np.abs(U - O.matrix(sq, x)).max() < 1e-12
