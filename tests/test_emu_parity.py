"""CPU suite: the CUDA kernel sources, compiled for the host by tests/emu (threads + barriers), against the
reference-generated fixtures.  This checks kernel indexing / synchronisation logic without a GPU; the real parity
gate is tests/test_gpu_parity.py on the B200."""
import numpy as np
import pytest

from tests import _golden, _parity
from tests.emu import emulated_library


@pytest.mark.parametrize("fname,stride", [("gates_small", 1), ("gateset_layers", 3), ("microbench", 4), ("leap_long", 2), ("qutrit", 1)])
def test_emulated_kernels_match_reference_fixtures(fname, stride):
    lib = emulated_library()
    cases = _golden.load(fname)
    for case in cases[::stride]:
        if case.circuit.qudits == 5 and len(case.circuit._subgates) > 12:
            continue  # 512 emulated threads x long programs: left to the GPU suite
        _parity.check_fixture_case(lib, case)


def test_emulated_ragged_batch_matches_oracle():
    from qsearch_b200 import gatesets, unitaries
    lib = emulated_library()
    rng = np.random.default_rng(5)
    g = gatesets.QubitCNOTLinear()
    c0 = gatesets.ProductGate(g.initial_layer(3))
    c1 = g.successors(c0)[0][0]
    c2 = g.successors(c1)[1][0]
    circuits = [c0, c1, c2]
    xs = [rng.random((k, c.num_inputs)) * 2 * np.pi for k, c in zip((2, 1, 3), circuits)]
    assert _parity.check_against_oracle(lib, circuits, xs, unitaries.qft(8)) == 6


def test_bad_programs_are_rejected():
    from qsearch_b200 import lowering, gates, _lib
    lib = emulated_library()
    gp = lowering.lower(gates.KroneckerGate(gates.U3Gate(), gates.U3Gate()))
    bad = lowering.GateProgram(gp.n_qubits, gp.n_params, gp.ops.copy(), gp.consts)
    bad.ops["q0"][1] = 7
    with pytest.raises(_lib.QsbError):
        lib.eval_matrix([bad], [0], np.zeros(6))
    with pytest.raises(ValueError):  # qubit and qutrit leaves in one tree
        lowering.lower(gates.KroneckerGate(gates.SingleQutritGate(), gates.U3Gate()))
    with pytest.raises(ValueError):  # more qutrits than the interpreter's tiles hold
        lowering.lower(gates.KroneckerGate(*[gates.SingleQutritGate()] * 4))
    q3 = lowering.lower(gates.SingleQutritGate())
    assert (q3.radix, q3.dim, q3.n_params) == (3, 3, 8)
    as_qubits = lowering.GateProgram(q3.n_qubits, q3.n_params, q3.ops, q3.consts)  # radix 2: QUTRIT1 is not a qubit op
    with pytest.raises(_lib.QsbError):
        lib.eval_matrix([as_qubits], [0], np.zeros(8))
    with pytest.raises(_lib.QsbError):  # one batch, two radices
        lib.eval_matrix([gp, q3], [0, 1], np.zeros(14))
    with pytest.raises(ValueError):
        lib.eval_matrix([gp], [0], np.zeros(5))


def test_chunked_pipelined_host_calls(monkeypatch):
    """Host-buffer calls stream through several device chunks (two buffer sets, copy overlapping the next kernel):
    force tiny chunks and compare with the single-chunk result."""
    from qsearch_b200 import gatesets, lowering, unitaries
    lib = emulated_library()
    rng = np.random.default_rng(8)
    g = gatesets.QubitCNOTRing()
    c = gatesets.ProductGate(g.initial_layer(3)).appending(g.search_layers(3)[1][0]).appending(g.search_layers(3)[2][0])
    gp = lowering.lower(c)
    xs = rng.random((7, c.num_inputs)) * 2 * np.pi
    poi = np.zeros(7, dtype="int32")
    T = unitaries.qft(8)
    ref = lib.eval_residuals_jac([gp], poi, xs, T) + lib.eval_mat_jac([gp], poi, xs) + lib.eval_hs_grad([gp], poi, xs, T)
    monkeypatch.setenv("QSB_CHUNK_BYTES", str(2 * 64 * 8 * (c.num_inputs + 1) * 2 + 8))  # two instances per chunk
    got = lib.eval_residuals_jac([gp], poi, xs, T) + lib.eval_mat_jac([gp], poi, xs) + lib.eval_hs_grad([gp], poi, xs, T)
    for a, b in zip(ref, got):
        assert np.array_equal(a, b)
    pinned = lib.pinned_empty((7, 2 * 64))
    r, _ = lib.eval_residuals_jac([gp], poi, xs, T, out_r=pinned)
    assert r is pinned and np.array_equal(pinned, ref[0])


def test_emulated_warp_kernels_match_interpreter():
    """Register/DMMA warp kernels (mat_jac, residual Jacobian, HS gradient) against the shared-memory interpreter on
    the same emulated device batch, n = 3 and 4."""
    from qsearch_b200 import _lib, lowering, workloads
    lib = emulated_library()
    for n in (3, 4):
        circuits, params, targets = workloads.microbench_workload(n, 2, 2, depth=4)
        gps = [lowering.lower(c) for c in circuits]
        poi = np.repeat(np.arange(2, dtype="int32"), 2)
        out = {}
        for kind in (0, 1):
            b = lib.batch(gps, poi)
            b.set_kernel(kind)
            b.set_params(params.reshape(len(poi), -1))
            b.set_target(targets["haar"])
            res = []
            for what in (_lib.RUN_MATJAC, _lib.RUN_RESID, _lib.RUN_HS):
                b.run(what)
                res.extend(x for x in b.fetch() if x is not None)
            out[kind] = res
            b.close()
        for a, g in zip(out[0], out[1]):
            assert a.shape == g.shape and np.abs(a - g).max() < 1e-13
        assert any(not np.array_equal(a, g) for a, g in zip(out[0], out[1]))  # two different code paths did run


def test_edge_cases_constants_empty_and_mixed_batches():
    """Programs without parameters (constants + CNOT) through every entry point including the solvers, an empty
    batch, and one ragged batch mixing three structures (one of them with a constant leaf and a ZXZXZ gate)."""
    from oracle import qs_oracle as O
    from qsearch_b200 import gates, gatesets, lowering, unitaries
    lib = emulated_library()
    c = gates.KroneckerGate(gates.SXGate(), gates.CNOTGate())
    gp = lowering.lower(c)
    assert gp.n_params == 0
    Uo, T = O.matrix(c, []), unitaries.qft(8)
    none = np.zeros(0)
    assert np.abs(lib.eval_matrix([gp], [0], none)[0] - Uo).max() < 1e-15
    U2, J = lib.eval_mat_jac([gp], [0], none)
    assert np.abs(U2[0] - Uo).max() < 1e-15 and J.shape == (0, 8, 8)
    r, Jr = lib.eval_residuals_jac([gp], [0], none, T)
    f, g = lib.eval_hs_grad([gp], [0], none, T)
    assert np.abs(r[0] - O.matrix_residuals(T, Uo)).max() < 1e-15 and Jr.shape == (0, 128) and g.shape == (0,)
    assert abs(f[0] - O.matrix_distance_squared(T, Uo)) < 1e-15
    for method in ("lm", "lbfgs"):
        x, ff, it, st = lib.solve(method, [gp], [0], none, T)
        assert x.size == 0 and abs(ff[0] - f[0]) < 1e-15
    assert lib.eval_matrix([gp], np.zeros(0, dtype="int32"), none).shape == (0, 8, 8)
    g4 = gatesets.QubitCNOTRing()
    c1 = gates.ProductGate(g4.initial_layer(4))
    c2 = c1.appending(g4.search_layers(4)[3][0])
    c3 = gates.KroneckerGate(gates.SXGate(), gates.U3Gate(), gates.IdentityGate(), gates.ZXZXZGate())
    rng = np.random.default_rng(0)
    circuits, counts = [c1, c2, c3], (1, 2, 2)
    assert _parity.check_against_oracle(lib, circuits, [rng.random((k, z.num_inputs)) * 2 * np.pi for k, z in zip(counts, circuits)],
                                        unitaries.qft(16)) == 5


@pytest.mark.parametrize("n", [3, 4, 5])
def test_lm_system_run_kind(n):
    """QSB_RUN_LMSYS (r'r, J_r'r, J_r'J_r per instance, J_r never leaving the device) against the oracle: the body-frame
    form of qsb_lm3.cuh (unitary target) in all three row-per-lane geometries."""
    from oracle import qs_oracle as O
    from qsearch_b200 import _lib, lowering, workloads
    lib = emulated_library()
    circuits, params, targets = workloads.microbench_workload(n, 2, 2, depth=3)
    gps = [lowering.lower(c) for c in circuits]
    poi = np.repeat(np.arange(2, dtype="int32"), 2)
    T = targets["haar"]
    b = lib.batch(gps, poi)
    b.set_params(params.reshape(len(poi), -1))
    b.set_target(T)
    b.run(_lib.RUN_LMSYS)
    rr, g, A = b.fetch()
    b.close()
    P = gps[0].n_params
    for i in range(len(poi)):
        x = params.reshape(len(poi), -1)[i]
        U, J = O.mat_jac(circuits[poi[i]], x)
        r, Jr = O.matrix_residuals(T, U), O.matrix_residuals_jac(T, U, J)
        assert abs(rr[i] - r @ r) < 1e-12
        assert np.abs(g[i * P:(i + 1) * P] - Jr.T @ r).max() < 1e-12
        assert np.abs(A[i * P * P:(i + 1) * P * P].reshape(P, P) - Jr.T @ Jr).max() < 1e-12


def test_every_leaf_type_through_the_warp_kernels():
    """Rank-one derivative forms (generator-type and phase-type parameters) of all leaves against the oracle, n = 3, 4."""
    lib = emulated_library()
    for n in (3, 4):
        _parity.check_every_leaf_type(lib, n)


def test_program_cache_eviction_keeps_referenced_programs():
    """the device-program cache must never destroy a program a live Batch (or a call in flight) references.
    A tiny cache limit forces evictions while a batch is held; the batch must still evaluate correctly afterwards."""
    from oracle import qs_oracle as O
    from qsearch_b200 import _lib, lowering, gates, gatesets
    lib = _lib.Library(emulated_library().path)
    lib.PROGRAM_CACHE_LIMIT = 8
    rng = np.random.default_rng(5)
    g = gatesets.QubitCNOTLinear()
    base = gates.ProductGate(g.initial_layer(2)).appending(g.search_layers(2)[0][0])
    gp0 = lowering.lower(base)
    gp0.key = "held"
    x = rng.random((2, gp0.n_params)) * 2 * np.pi
    held = lib.batch([gp0], np.zeros(2, dtype="int32"))
    held.set_params(x)
    for i in range(40):  # 40 distinct programs through a cache of 8: evictions happen, `held` stays referenced
        gpi = lowering.lower(base)
        gpi.key = "tmp%d" % i
        lib.eval_matrix([gpi], [0], x[0])
    assert len(lib._programs) <= 9 and "held" in lib._programs and lib._programs["held"][2] == 1
    # a single call with more distinct programs than the limit keeps all of ITS programs alive
    many = []
    for i in range(12):
        gpi = lowering.lower(base)
        gpi.key = "many%d" % i
        many.append(gpi)
    U = lib.eval_matrix(many, np.arange(12, dtype="int32"), np.tile(x[0], 12))
    assert np.abs(U - O.matrix(base, x[0])[None]).max() < 1e-12
    held.run(_lib.RUN_MATRIX)
    Uh, _, _ = held.fetch()
    for k in range(2):
        assert np.abs(Uh[k] - O.matrix(base, x[k])).max() < 1e-12
    held.close()
    assert lib._programs["held"][2] == 0


def test_captured_constants_are_keyed_by_their_exact_bytes():
    """two UGates that differ by 1e-10 (below the 8 digits of NumPy's repr) must not share a device program."""
    from oracle import qs_oracle as O
    from qsearch_b200 import backends, gates
    lib = emulated_library()
    prev = __import__("qsearch_b200")._lib.use_library(lib)
    try:
        rng = np.random.default_rng(9)
        A = np.linalg.qr(rng.normal(size=(4, 4)) + 1j * rng.normal(size=(4, 4)))[0]
        B = A.copy()
        B[1, 2] += 1e-10
        ca = gates.ProductGate(gates.KroneckerGate(gates.U3Gate(), gates.U3Gate()), gates.UGate(A))
        cb = gates.ProductGate(gates.KroneckerGate(gates.U3Gate(), gates.U3Gate()), gates.UGate(B))
        assert repr(ca) == repr(cb)
        x = rng.random(6) * 2 * np.pi
        Ua = backends.CudaBackend().prepare_circuit(ca).matrix(x)
        Ub = backends.CudaBackend().prepare_circuit(cb).matrix(x)
        assert np.abs(Ua - O.matrix(ca, x)).max() < 1e-12 and np.abs(Ub - O.matrix(cb, x)).max() < 1e-12
        assert np.abs(Ua - Ub).max() > 1e-11
    finally:
        __import__("qsearch_b200")._lib.use_library(prev)


def test_zero_overlap_gradient_convention():
    lib = emulated_library()
    for n in (2, 3):
        _parity.check_zero_overlap(lib, n)


def test_general_lm_kernel_for_a_non_unitary_target():
    """The body-frame LM needs a unitary target; anything else (nothing in the reference forbids it) must take the general
    kernel and still minimise |U T^dagger - I|: here T = 0.8 x a reachable unitary, whose minimiser is the same circuit."""
    from oracle import qs_oracle as O
    from qsearch_b200 import gates, gatesets, lowering
    lib = emulated_library()
    rng = np.random.default_rng(31)
    g = gatesets.QubitCNOTLinear()
    c = gates.ProductGate(g.initial_layer(3)).appending(g.search_layers(3)[0][0]).appending(g.search_layers(3)[1][0])
    x_star = rng.random(c.num_inputs) * 2 * np.pi
    T = 0.8 * O.matrix(c, x_star)
    gp = lowering.lower(c)
    x0 = x_star[None, :] + 0.05 * rng.standard_normal((2, c.num_inputs))
    x, f, iters, status = lib.solve("lm", [gp], np.zeros(2, dtype="int32"), x0, T)
    for k, xk in enumerate(x.reshape(2, -1)):
        assert abs(O.matrix_distance_squared(T, O.matrix(c, xk)) - f[k]) < 1e-12
        r = O.matrix_residuals(T, O.matrix(c, xk))
        assert abs(0.5 * r @ r - 0.5 * 8 * 0.2 ** 2) < 1e-9  # U T^dagger = 0.8 I at the optimum: |0.8 I - I|^2 / 2


@pytest.mark.parametrize("kernel", ["lm3", "lm2"])
def test_batch_stop_rules(kernel, monkeypatch):
    """qsb_solver_opts.straggler_factor / sibling_stop: off by default (every start runs to its own convergence); with a
    sibling threshold the other starts of a solved circuit stop (status 7) and the best result is unchanged.  Both the
    body-frame kernel and the general one (circuits too long for the former's shared-memory factor) implement them."""
    from oracle import qs_oracle as O
    from qsearch_b200 import gates, gatesets, lowering
    lib = emulated_library()
    monkeypatch.setenv("QSB_LM_KERNEL", kernel)
    rng = np.random.default_rng(17)
    g = gatesets.QubitCNOTLinear()
    c = gates.ProductGate(g.initial_layer(3)).appending(g.search_layers(3)[0][0]).appending(g.search_layers(3)[1][0])
    T = O.matrix(c, rng.random(c.num_inputs) * 2 * np.pi)
    gp = lowering.lower(c)
    K = 24
    x0 = rng.random((K, c.num_inputs)) * 2 * np.pi
    poi = np.zeros(K, dtype="int32")
    x, f, iters, status = lib.solve("lm", [gp], poi, x0, T)
    assert set(np.unique(status)) <= {1, 2, 3, 4, 5} and f.min() < 1e-10
    x2, f2, iters2, status2 = lib.solve("lm", [gp], poi, x0, T, sibling_stop=1e-10, straggler_factor=2.0)
    assert f2.min() < 1e-10
    assert (status2 == 7).sum() + (status2 == 6).sum() > 0 and iters2.sum() < iters.sum()
    for k in range(K):  # whatever stopped an instance, the reported distance is the distance at the returned parameters
        assert abs(O.matrix_distance_squared(T, O.matrix(c, x2.reshape(K, -1)[k])) - f2[k]) < 1e-12


def test_two_qubit_rotations_against_oracle():
    """the native backend's RXX / RYY / RZZ leaves through every run kind of the interpreter."""
    lib = emulated_library()
    assert _parity.check_two_qubit_rotations(lib, 2) == 4
    assert _parity.check_two_qubit_rotations(lib, 3) == 4


def test_solvers_on_a_circuit_with_two_qubit_rotations():
    from oracle import qs_oracle as O
    from qsearch_b200 import lowering
    lib = emulated_library()
    c = _parity.two_qubit_rotation_circuits(2)[0]
    rng = np.random.default_rng(3)
    T = O.matrix(c, rng.random(c.num_inputs) * 2 * np.pi)
    gp = lowering.lower(c)
    x0 = rng.random((4, gp.n_params)) * 2 * np.pi
    for method in ("lm", "lbfgs"):
        x, f, iters, status = lib.solve(method, [gp], np.zeros(4, dtype="int32"), x0, T)
        assert f.min() < 1e-10, (method, f)
        k = int(np.argmin(f))
        assert abs(O.matrix_distance_squared(T, O.matrix(c, x.reshape(4, -1)[k])) - f[k]) < 1e-12


@pytest.mark.parametrize("n", [3, 4])
def test_tf32_gram_option_of_the_lm_kernel(n):
    """qsb_solver_opts.gram_tf32: only J_r^T J_r comes from TF32-rounded rows (mma.sync.m16n8k8.tf32; emulated here); the
    gradient, the objective and the returned distance stay fp64, so a reachable target is still met to rounding and the
    reported distance is exact at the returned parameters."""
    from oracle import qs_oracle as O
    from qsearch_b200 import gates, gatesets, lowering
    lib = emulated_library()
    rng = np.random.default_rng(50 + n)
    g = gatesets.QubitCNOTLinear()
    c = gates.ProductGate(g.initial_layer(n))
    for k in range(3):
        c = c.appending(g.search_layers(n)[k % (n - 1)][0])
    T = O.matrix(c, rng.random(c.num_inputs) * 2 * np.pi)
    gp = lowering.lower(c)
    x0 = rng.random((4, gp.n_params)) * 2 * np.pi
    x, f, iters, status = lib.solve("lm", [gp], np.zeros(4, dtype="int32"), x0, T, gram_tf32=1.0)
    assert f.min() < 1e-12
    for k, xk in enumerate(x.reshape(4, -1)):
        assert abs(O.matrix_distance_squared(T, O.matrix(c, xk)) - f[k]) < 1e-12


def test_solvers_on_qutrit_circuits():
    """SURVEY 8f rank 4 (qutrit leaves, gates.py:411-602): both solvers on a reachable two-qutrit target of the
    QutritCNOTLinear gateset (gatesets.py:295-305), through the radix-3 interpreter kernels."""
    from oracle import qs_oracle as O
    from qsearch_b200 import gatesets, lowering
    lib = emulated_library()
    gs = gatesets.QutritCNOTLinear()
    c = gatesets.ProductGate(gs.initial_layer(2)).appending(gs.search_layers(2)[0][0])
    rng = np.random.default_rng(11)
    x_star = rng.random(c.num_inputs) * 2 * np.pi
    T = O.matrix(c, x_star)
    gp = lowering.lower(c)
    assert (gp.radix, gp.dim, gp.n_params) == (3, 9, 32)
    x0 = x_star + 0.2 * rng.standard_normal((3, gp.n_params))
    for method in ("lm", "lbfgs"):
        x, f, iters, status = lib.solve(method, [gp], np.zeros(3, dtype="int32"), x0, T, max_iters=400 if method == "lbfgs" else 0)
        assert f.min() < (1e-10 if method == "lm" else 1e-6), (method, f)
        k = int(np.argmin(f))
        assert abs(O.matrix_distance_squared(T, O.matrix(c, x.reshape(3, -1)[k])) - f[k]) < 1e-12
    with pytest.raises(Exception):  # the LM-system run kind belongs to the qubit warp kernels
        b = lib.batch([gp], [0])
        try:
            b.set_params(x0[0]); b.set_target(T); b.run(6)
        finally:
            b.close()


@pytest.mark.parametrize("n", [3, 4])
def test_four_warp_shape_of_the_body_frame_lm(n, monkeypatch):
    """qsb_k_lm3<n, tf32, 4>: the body-frame LM as 4-warp CTAs (two per SM; chosen for batches of more than one instance per
    SM, forced here with QSB_LM_WARPS).  Same algorithm, different split of the work over warps: same end points."""
    from oracle import qs_oracle as O
    from qsearch_b200 import gates, gatesets, lowering
    lib = emulated_library()
    rng = np.random.default_rng(23 + n)
    g = gatesets.QubitCNOTLinear()
    c = gates.ProductGate(g.initial_layer(n)).appending(g.search_layers(n)[0][0]).appending(g.search_layers(n)[n - 2][0])
    T = O.matrix(c, rng.random(c.num_inputs) * 2 * np.pi)
    gp = lowering.lower(c)
    K = 6
    x0 = rng.random((K, c.num_inputs)) * 2 * np.pi
    poi = np.zeros(K, dtype="int32")
    out = {}
    for warps in ("8", "4"):
        monkeypatch.setenv("QSB_LM_WARPS", warps)
        for tf32 in (0, 1):
            out[warps, tf32] = lib.solve("lm", [gp], poi, x0, T, gram_tf32=tf32)
    for tf32 in (0, 1):
        (x8, f8, it8, st8), (x4, f4, it4, st4) = out["8", tf32], out["4", tf32]
        assert f4.min() < 1e-10 and f8.min() < 1e-10
        assert np.abs(np.sort(f4) - np.sort(f8)).max() < 1e-5 and abs(int(it4.sum()) - int(it8.sum())) <= 0.2 * it8.sum() + 4
        for k in range(K):
            assert abs(O.matrix_distance_squared(T, O.matrix(c, x4.reshape(K, -1)[k])) - f4[k]) < 1e-12
