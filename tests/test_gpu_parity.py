"""GPU parity gate (-m gpu): the CUDA path through the C ABI against (i) the reference-generated fixtures,
(ii) the NumPy oracle on seeded inputs, (iii) size-independent properties at the benchmark's full size.
Tolerance: 1e-12 max-abs (north_star).  Modelled on the reference's tests/test_circuit.py and tests/test_gradients.py."""
import numpy as np
import pytest

from oracle import qs_oracle as O
from qsearch_b200 import _lib, gates, gatesets, lowering, unitaries, workloads
from qsearch_b200.backends import CudaBackend, CudaCircuit
from tests import _golden, _parity

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    return _lib.get_library()


@pytest.mark.parametrize("fname", _golden.ALL_FILES)
def test_fixtures(lib, fname):
    for case in _golden.load(fname):
        _parity.check_fixture_case(lib, case)


def _reference_gate_list():
    u, xzxz, zxzxz, cnot = gates.U3Gate(), gates.XZXZGate(), gates.ZXZXZGate(), gates.CNOTGate()
    return [gates.XGate(), gates.YGate(), gates.ZGate(), xzxz, zxzxz, u, gates.U2Gate(), gates.U1Gate(),
            gates.ProductGate(u, xzxz, u), gates.KroneckerGate(u, xzxz, u), gates.ProductGate(u, zxzxz, u),
            gates.KroneckerGate(u, zxzxz, u), gates.KroneckerGate(u, u), gates.KroneckerGate(u, cnot, xzxz),
            gates.ProductGate(gates.KroneckerGate(u, u, u),
                              gates.KroneckerGate(gates.ProductGate(cnot, gates.KroneckerGate(xzxz, u)), gates.IdentityGate()))]


@pytest.mark.parametrize("gate", _reference_gate_list(), ids=lambda g: repr(g)[:40])
def test_backend_matches_oracle(lib, gate):
    """tests/test_circuit.py:8-23 with the CUDA backend in place of qsrs."""
    circ = CudaBackend().prepare_circuit(gate)
    assert not isinstance(circ, gates.Gate)  # tests/test_defaults.py:42
    assert CudaBackend().prepare_circuit(circ) is circ  # tests/test_circuit.py:60-62
    rng = np.random.default_rng(3)
    for _ in range(10):
        v = rng.random(gate.num_inputs) * 2 * np.pi
        U, J = circ.mat_jac(v)
        Uo, Jo = O.mat_jac(gate, v)
        assert np.abs(circ.matrix(v) - Uo).max() < 1e-12 and np.abs(U - Uo).max() < 1e-12
        assert len(J) == len(Jo) == circ.num_inputs
        for a, b in zip(J, Jo):
            assert np.abs(a - b).max() < 1e-12
        assert O.matrix_distance_squared(U, Uo) < 1e-10


@pytest.mark.parametrize("gate", _reference_gate_list(), ids=lambda g: repr(g)[:40])
def test_finite_difference(lib, gate):
    """tests/test_gradients.py:29-50: central differences of matrix() against mat_jac()."""
    circ = CudaCircuit(gate)
    rng = np.random.default_rng(4)
    eps = 1e-5
    P = gate.num_inputs
    V = rng.random((20, P)) * 2 * np.pi
    for v in V:
        _, J = circ.mat_jac(v)
        pts = np.concatenate([v + eps * np.eye(P), v - eps * np.eye(P)])
        Us = circ.matrix_batch(pts)
        fd = (Us[:P] - Us[P:]) / (2 * eps)
        assert np.abs(fd - J).sum(axis=(1, 2)).max() < 1e-7


@pytest.mark.parametrize("n", [3, 4, 5])
def test_every_leaf_type_through_the_warp_kernels(lib, n):
    """All parameterised leaves (generator-type and phase-type derivatives), a constant leaf and CNOTs in both
    directions through the register/DMMA kernels, including all-zero angles."""
    _parity.check_every_leaf_type(lib, n)


@pytest.mark.parametrize("n", [2, 3, 4, 5])
def test_ragged_batches_against_oracle(lib, n):
    rng = np.random.default_rng(100 + n)
    circuits = []
    for gs in (gatesets.QubitCNOTLinear(), gatesets.QubitCNOTRing(),
               gatesets.QubitCNOTAdjacencyList([(0, 1), (0, 2), (1, 2), (2, 3), (2, 4), (3, 4)])):
        layers = gs.search_layers(n)
        for depth in (0, 1, 3, 7):
            c = gates.ProductGate(gs.initial_layer(n))
            for _ in range(depth):
                c = c.appending(layers[int(rng.integers(len(layers)))][0])
            circuits.append(c)
    xs = [rng.random((int(rng.integers(1, 4)), c.num_inputs)) * 2 * np.pi for c in circuits]
    for T in (unitaries.qft(1 << n), workloads.haar_unitary(1 << n, 77)):
        _parity.check_against_oracle(lib, circuits, xs, T)


def test_empty_and_zero_parameter_inputs(lib):
    gp = lowering.lower(gates.KroneckerGate(gates.CNOTGate(), gates.IdentityGate()))
    U = lib.eval_matrix([gp], [0, 0], np.zeros(0))
    assert np.array_equal(U[0], np.kron(unitaries.cnot, np.eye(2))) and np.array_equal(U[0], U[1])
    U2, J = lib.eval_mat_jac([gp], [0], np.zeros(0))
    assert J.shape == (0, 8, 8)
    gp2 = lowering.lower(gates.U3Gate())
    assert lib.eval_matrix([gp2], np.zeros(0, dtype="int32"), np.zeros(0)).shape == (0, 2, 2)


@pytest.mark.parametrize("n", [4, 5])
def test_full_size_properties(lib, n):
    """At the benchmark's size the oracle is too slow; use properties that hold for every instance:
    unitarity, |r|^2 = 2N - 2 Re Tr(U T^dagger), J_r[k].r = -Re Tr(K_k), f from HS equals f from the residual trace,
    and bit-identical results for an instance evaluated alone or inside the big batch."""
    n_circ, n_vec = (100, 200) if n == 4 else (40, 50)
    circuits, params, targets = workloads.microbench_workload(n, n_circ, n_vec)
    n_circ, n_vec = params.shape[:2]
    gps = [lowering.lower(c) for c in circuits]
    P, N = circuits[0].num_inputs, 1 << n
    poi = np.repeat(np.arange(n_circ, dtype="int32"), n_vec)
    flat = params.reshape(-1, P)
    T = targets["haar"]
    U = lib.eval_matrix(gps, poi, flat)
    assert np.abs(U @ U.conj().transpose(0, 2, 1) - np.eye(N)).max() < 1e-12
    f, g = lib.eval_hs_grad(gps, poi, flat, T)
    d = lib.eval_distance(gps, poi, flat, T)
    tr = np.einsum("bij,ij->b", U, T.conj())  # Tr(U T^dagger)
    assert np.abs(f - (1 - np.abs(tr) / N)).max() < 1e-12 and np.abs(d - np.abs(f)).max() < 1e-12
    sel = np.arange(0, len(poi), max(1, len(poi) // 512))
    r, Jr = lib.eval_residuals_jac(gps, poi[sel], flat[sel], T)
    Jr = Jr.reshape(len(sel), P, 2 * N * N)
    assert np.abs((r * r).sum(1) - (2 * N - 2 * tr[sel].real)).max() < 1e-11
    diag = np.arange(N) * (N + 1)
    assert np.abs(np.einsum("bkm,bm->bk", Jr, r) + Jr[:, :, diag].sum(2)).max() < 1e-11
    # gradient of f from J_r: S_k = Tr(T^dagger J_k) = conj-symmetric pieces of K_k's diagonal
    trK = Jr[:, :, diag].sum(2) + 1j * Jr[:, :, N * N + diag].sum(2)  # Tr(K_k) = Tr(J_k T^dagger)
    S = tr[sel]
    g_from_J = -(S.real[:, None] * trK.real + S.imag[:, None] * trK.imag) / (N * np.abs(S)[:, None])
    assert np.abs(g.reshape(-1, P)[sel] - g_from_J).max() < 1e-12
    for i in (0, len(sel) // 2, len(sel) - 1):
        r1, J1 = lib.eval_residuals_jac(gps, poi[sel[i]:sel[i] + 1], flat[sel[i]:sel[i] + 1], T)
        assert np.array_equal(r1[0], r[i]) and np.array_equal(J1, Jr[i])
    # a slice against the oracle itself
    for i in sel[:3]:
        ro, Jo = O.eval_lm(circuits[poi[i]], T, flat[i])
        k = int(np.where(sel == i)[0][0])
        assert np.abs(r[k] - ro).max() < 1e-12 and np.abs(Jr[k].T - Jo).max() < 1e-12


def test_device_resident_batch_matches_host_calls(lib):
    circuits, params, targets = workloads.microbench_workload(4, 6, 5)
    gps = [lowering.lower(c) for c in circuits]
    poi = np.repeat(np.arange(6, dtype="int32"), 5)
    flat = params.reshape(30, -1)
    T = targets["qft"]
    b = lib.batch(gps, poi)
    b.set_params(flat)
    b.set_target(T)
    ms = b.run(_lib.RUN_RESID, first=7, count=11, timed=True)
    assert ms > 0
    r, Jr, _ = b.fetch()
    r2, Jr2 = lib.eval_residuals_jac(gps, poi[7:18], flat[7:18], T)
    assert np.array_equal(r, r2) and np.array_equal(Jr, Jr2)
    b.run(_lib.RUN_HS)
    f, g, _ = b.fetch()
    f2, g2 = lib.eval_hs_grad(gps, poi, flat, T)
    assert np.array_equal(f, f2) and np.array_equal(g, g2)
    b.close()


@pytest.mark.parametrize("n", [3, 4, 5])
def test_warp_kernels_match_interpreter(lib, n):
    """The register/DMMA warp kernels against the shared-memory interpreter on the same device-resident batch."""
    circuits, params, targets = workloads.microbench_workload(n, 6, 5)
    gps = [lowering.lower(c) for c in circuits]
    poi = np.repeat(np.arange(6, dtype="int32"), 5)
    out = {}
    for kind in (0, 1):
        b = lib.batch(gps, poi)
        b.set_kernel(kind)
        b.set_params(params.reshape(len(poi), -1))
        b.set_target(targets["haar"] if "haar" in targets else targets["qft"])
        res = []
        for what in (_lib.RUN_MATJAC, _lib.RUN_RESID, _lib.RUN_HS):
            b.run(what)
            res.extend(x for x in b.fetch() if x is not None)
        out[kind] = res
        b.close()
    for a, g in zip(out[0], out[1]):
        assert a.shape == g.shape and np.abs(a - g).max() < 1e-13


def test_tensor_core_lm_matches_interpreter_lm(lib, monkeypatch):
    """Same LM trajectory from both kernel families: identical iteration counts, same end point."""
    rng = np.random.default_rng(12)
    g = gatesets.QubitCNOTRing()
    c = gates.ProductGate(g.initial_layer(4))
    layers = g.search_layers(4)
    for _ in range(5):
        c = c.appending(layers[int(rng.integers(len(layers)))][0])
    T = O.matrix(c, rng.random(c.num_inputs) * 2 * np.pi)
    gp = lowering.lower(c)
    x0 = rng.random((8, gp.n_params)) * 2 * np.pi
    poi = np.zeros(8, dtype="int32")
    x1, f1, it1, st1 = lib.solve("lm", [gp], poi, x0, T)
    monkeypatch.setenv("QSB_KERNEL", "generic")
    lib2 = _lib.Library(devices=[0])
    x2, f2, it2, st2 = lib2.solve("lm", [gp], poi, x0, T)
    assert f1.min() < 1e-10 and f2.min() < 1e-10
    assert np.abs(f1 - f2).max() < 1e-9


def test_two_devices_shard_one_batch():
    """One context over two GPUs: the library splits the batch into contiguous shards, results are identical."""
    import ctypes
    one = _lib.Library(devices=[0])
    n = ctypes.c_int()
    one.check(one.dll.qsb_device_count(ctypes.byref(n)))
    if n.value < 2:
        pytest.skip("needs two GPUs")
    two = _lib.Library(devices=[0, 1])
    circuits, params, targets = workloads.microbench_workload(4, 3, 7)
    gps = [lowering.lower(c) for c in circuits]
    poi = np.repeat(np.arange(3, dtype="int32"), 7)
    flat = params.reshape(len(poi), -1)
    a = one.eval_residuals_jac(gps, poi, flat, targets["qft"]) + one.eval_hs_grad(gps, poi, flat, targets["qft"])
    b = two.eval_residuals_jac(gps, poi, flat, targets["qft"]) + two.eval_hs_grad(gps, poi, flat, targets["qft"])
    for x, y in zip(a, b):
        assert np.array_equal(x, y)
    for x, y in zip(one.solve("lm", gps, poi, flat, targets["qft"]), two.solve("lm", gps, poi, flat, targets["qft"])):
        assert np.allclose(x, y, atol=1e-9)


@pytest.mark.parametrize("n", [3, 4, 5])
def test_lm_system_run_kind(lib, n):
    """QSB_RUN_LMSYS on the device: r'r, J_r'r and J_r'J_r per instance against the oracle."""
    circuits, params, targets = workloads.microbench_workload(n, 3, 2, depth=5)
    gps = [lowering.lower(c) for c in circuits]
    poi = np.repeat(np.arange(3, dtype="int32"), 2)
    T = targets["haar"]
    flat = params.reshape(len(poi), -1)
    b = lib.batch(gps, poi)
    b.set_params(flat)
    b.set_target(T)
    b.run(_lib.RUN_LMSYS)
    rr, g, A = b.fetch()
    b.close()
    P = gps[0].n_params
    for i in range(len(poi)):
        U, J = O.mat_jac(circuits[poi[i]], flat[i])
        r, Jr = O.matrix_residuals(T, U), O.matrix_residuals_jac(T, U, J)
        assert abs(rr[i] - r @ r) < 1e-11
        assert np.abs(g[i * P:(i + 1) * P] - Jr.T @ r).max() < 1e-11
        assert np.abs(A[i * P * P:(i + 1) * P * P].reshape(P, P) - Jr.T @ Jr).max() < 1e-11


def test_zero_overlap_gradient_convention(lib):
    for n in (3, 4, 5):
        _parity.check_zero_overlap(lib, n)


def test_inserting_programs_against_oracle(lib):
    """ProductGate.inserting (gates.py:1022-1030): the circuits the re-optimising post-processor builds
    (post_processing.py:250) lower to programs whose parameter order shifts behind the insertion point."""
    rng = np.random.default_rng(8)
    g = gatesets.QubitCNOTLinear()
    base = gates.ProductGate(g.initial_layer(4))
    layers = g.search_layers(4)
    for k in range(5):
        base = base.appending(layers[k % len(layers)][0])
    circuits = [base.inserting(layers[(k + 1) % len(layers)][0], depth=k + 1) for k in range(4)]
    circuits.append(gates.ProductGate(*base._subgates[:2], *base._subgates[4:]))  # a window cut out (post_processing.py:247)
    xs = [rng.random((2, c.num_inputs)) * 2 * np.pi for c in circuits]
    assert _parity.check_against_oracle(lib, circuits, xs, unitaries.qft(16)) == 10


def test_body_frame_lm_agrees_with_the_general_kernel(lib, monkeypatch):
    """qsb_k_lm3 (body-frame normal equations, unitary target) and qsb_k_lm2 (explicit J_r) are the same Gauss-Newton model:
    same convergence on reachable targets, and on an unreachable one the same local minima from the same starts."""
    rng = np.random.default_rng(77)
    for n, depth in ((3, 4), (4, 5), (5, 4)):
        g = gatesets.QubitCNOTRing()
        c = gates.ProductGate(g.initial_layer(n))
        layers = g.search_layers(n)
        for _ in range(depth):
            c = c.appending(layers[int(rng.integers(len(layers)))][0])
        gp = lowering.lower(c)
        K = 12
        x0 = rng.random((K, gp.n_params)) * 2 * np.pi
        poi = np.zeros(K, dtype="int32")
        for T in (O.matrix(c, rng.random(c.num_inputs) * 2 * np.pi), unitaries.qft(1 << n)):
            monkeypatch.setenv("QSB_LM_KERNEL", "lm3")
            x3, f3, it3, st3 = lib.solve("lm", [gp], poi, x0, T)
            monkeypatch.setenv("QSB_LM_KERNEL", "lm2")
            x2, f2, it2, st2 = lib.solve("lm", [gp], poi, x0, T)
            monkeypatch.delenv("QSB_LM_KERNEL")
            # ftol = xtol = 1e-8 (scipy's defaults, solvers.py:155) stop an unreachable minimum a few 1e-7 short in f; a reachable
            # target is hit to rounding by both
            assert abs(f3.min() - f2.min()) < 1e-5, (n, f3.min(), f2.min())
            assert (f3.min() < 1e-10) == (f2.min() < 1e-10)
            assert np.median(np.abs(f3 - f2)) < 1e-5  # most starts end in the same minimum (trajectories differ in rounding only)
            for k in range(K):
                assert abs(O.matrix_distance_squared(T, O.matrix(c, x3.reshape(K, -1)[k])) - f3[k]) < 1e-12


@pytest.mark.parametrize("kernel", ["lm3", "lm2"])
def test_batch_stop_rules(lib, kernel, monkeypatch):
    """Straggler cap and sibling stop (qsb_solver_opts): fewer iterations, same best value, consistent outputs; in the
    body-frame kernel and in the general one that takes over for circuits too long for it."""
    monkeypatch.setenv("QSB_LM_KERNEL", kernel)
    rng = np.random.default_rng(5)
    g = gatesets.QubitCNOTLinear()
    layers = g.search_layers(4)
    base = gates.ProductGate(g.initial_layer(4))
    for _ in range(5):
        base = base.appending(layers[int(rng.integers(len(layers)))][0])
    children = [base.appending(l) for l, _ in layers]
    T = O.matrix(children[1], rng.random(children[1].num_inputs) * 2 * np.pi)  # child 1 is reachable
    gps = [lowering.lower(c) for c in children]
    K = 32
    poi = np.repeat(np.arange(len(gps), dtype="int32"), K)
    x0 = np.concatenate([rng.random(gps[i].n_params) * 2 * np.pi for i in poi])
    x, f, it, st = lib.solve("lm", gps, poi, x0, T)
    x2, f2, it2, st2 = lib.solve("lm", gps, poi, x0, T, straggler_factor=2.0, sibling_stop=1e-10)
    assert set(np.unique(st)) <= {0, 1, 2, 3, 4, 5}
    assert f.reshape(3, K)[1].min() < 1e-10 and f2.reshape(3, K)[1].min() < 1e-10
    assert (st2 == 7).sum() > 0 and np.all(poi[st2 == 7] == 1)  # only starts of the solved child are stopped as siblings
    assert it2.sum() < it.sum() and it2.max() <= 2 * it.max()  # the capped batch also runs the gentler damping schedule: the tail is bounded by the cap, not by the plain run's
    off = np.concatenate([[0], np.cumsum([gps[i].n_params for i in poi])])
    for k in range(0, len(poi), 7):
        assert abs(O.matrix_distance_squared(T, O.matrix(children[poi[k]], x2[off[k]:off[k + 1]])) - f2[k]) < 1e-12


@pytest.mark.parametrize("n", [2, 3, 4, 5])
def test_two_qubit_rotations_against_oracle(lib, n):
    """RXX / RYY / RZZ (qsrs/src/circuits.rs:457-633) on the device: every run kind against the oracle, 1e-12."""
    assert _parity.check_two_qubit_rotations(lib, n) == 4


def test_solvers_on_a_circuit_with_two_qubit_rotations(lib):
    c = _parity.two_qubit_rotation_circuits(3)[0]
    rng = np.random.default_rng(3)
    T = O.matrix(c, rng.random(c.num_inputs) * 2 * np.pi)
    gp = lowering.lower(c)
    x0 = rng.random((16, gp.n_params)) * 2 * np.pi
    for method in ("lm", "lbfgs"):
        x, f, iters, status = lib.solve(method, [gp], np.zeros(16, dtype="int32"), x0, T)
        assert f.min() < 1e-10, (method, f)
        k = int(np.argmin(f))
        assert abs(O.matrix_distance_squared(T, O.matrix(c, x.reshape(16, -1)[k])) - f[k]) < 1e-12


@pytest.mark.parametrize("n", [3, 4, 5])
def test_tf32_gram_option_of_the_lm_kernel(lib, n):
    """gram_tf32 on the device: same convergence on reachable targets (distance exact at the returned x: fp64 everywhere except
    the Gauss-Newton matrix), and on an unreachable one minima as good as the fp64 kernel's."""
    rng = np.random.default_rng(60 + n)
    g = gatesets.QubitCNOTRing()
    layers = g.search_layers(n)
    c = gates.ProductGate(g.initial_layer(n))
    for _ in range(5):
        c = c.appending(layers[int(rng.integers(len(layers)))][0])
    gp = lowering.lower(c)
    K = 16
    x0 = rng.random((K, gp.n_params)) * 2 * np.pi
    poi = np.zeros(K, dtype="int32")
    T = O.matrix(c, rng.random(c.num_inputs) * 2 * np.pi)
    x, f, it, st = lib.solve("lm", [gp], poi, x0, T, gram_tf32=1.0)
    assert f.min() < 1e-10
    for k in range(K):
        assert abs(O.matrix_distance_squared(T, O.matrix(c, x.reshape(K, -1)[k])) - f[k]) < 1e-12
    Tq = unitaries.qft(1 << n)
    f64 = lib.solve("lm", [gp], poi, x0, Tq)[1]
    f32 = lib.solve("lm", [gp], poi, x0, Tq, gram_tf32=1.0)[1]
    assert f32.min() < f64.min() + 1e-4 and np.median(f32) < np.median(f64) + 1e-2


# ---------------------------------------------------------------------------- qutrits (gates.py:411-602)
def _qutrit_gate_list():
    sq, i3 = gates.SingleQutritGate(), gates.IdentityGate(d=3)
    ucnot = gates.UpgradedConstantGate(gates.CNOTGate())
    return [sq, gates.ProductGate(sq, sq), gates.KroneckerGate(sq, sq), gates.KroneckerGate(sq, gates.CPIGate()),
            gates.ProductGate(gates.KroneckerGate(sq, sq), gates.CSUMGate(), gates.KroneckerGate(sq, sq)),
            gates.ProductGate(gates.KroneckerGate(sq, sq, sq), gates.KroneckerGate(gates.ProductGate(ucnot, gates.KroneckerGate(sq, sq)), i3))]


@pytest.mark.parametrize("gate", _qutrit_gate_list(), ids=lambda g: repr(g)[:40])
def test_qutrit_backend_matches_oracle_and_finite_differences(lib, gate):
    """The d = 3 entries of the reference's backend tests (tests/test_circuit.py:36,41, tests/test_gradients.py:25) through the
    radix-3 interpreter kernels: matrix / mat_jac against the oracle, mat_jac against central differences."""
    circ = CudaBackend().prepare_circuit(gate)
    assert circ.program.radix == 3 and circ.program.dim == 3 ** gate.qudits
    rng = np.random.default_rng(5)
    P = gate.num_inputs
    for _ in range(5):
        v = rng.random(P) * 2 * np.pi
        U, J = circ.mat_jac(v)
        Uo, Jo = O.mat_jac(gate, v)
        assert np.abs(circ.matrix(v) - Uo).max() < 1e-12 and np.abs(U - Uo).max() < 1e-12
        assert len(J) == len(Jo) == P
        assert np.abs(J - np.array(Jo)).max() < 1e-12
        eps = 1e-5
        Us = circ.matrix_batch(np.concatenate([v + eps * np.eye(P), v - eps * np.eye(P)]))
        assert np.abs((Us[:P] - Us[P:]) / (2 * eps) - J).sum(axis=(1, 2)).max() < 1e-7


@pytest.mark.parametrize("n", [1, 2, 3])
def test_qutrit_ragged_batches_against_oracle(lib, n):
    """Several structures of different depth in one batch, every run kind, against the oracle; then a batch large enough to
    fill the GPU several times over, checked through unitarity and agreement with the small batch."""
    rng = np.random.default_rng(60 + n)
    sq = gates.SingleQutritGate()
    if n == 1:
        circuits = [sq, gates.ProductGate(sq, sq), gates.ProductGate(sq, sq, sq)]
    else:
        gs = gatesets.QutritCNOTLinear()
        layers = gs.search_layers(n)
        c0 = gates.ProductGate(gs.initial_layer(n))
        circuits = [c0]
        for k in range(3):
            circuits.append(circuits[-1].appending(layers[k % len(layers)][0]))
        np.random.seed(9)
        circuits.append(c0.appending(gatesets.QutritCPIPhaseLinear().search_layers(n)[-1][0]))
    N = 3 ** n
    T = workloads.haar_unitary(N, 77 + n)
    xs = [rng.random((2, c.num_inputs)) * 2 * np.pi for c in circuits]
    xs[0][1, :] = 0.0
    assert _parity.check_against_oracle(lib, circuits, xs, T) == 2 * len(circuits)
    c = circuits[-1]
    gp = lowering.lower(c)
    K = 3000
    V = rng.random((K, gp.n_params)) * 2 * np.pi
    V[:2] = xs[-1]
    U = lib.eval_matrix([gp], np.zeros(K, dtype="int32"), V)
    assert np.abs(U @ U.conj().transpose(0, 2, 1) - np.eye(N)).max() < 1e-12
    f, g = lib.eval_hs_grad([gp], np.zeros(K, dtype="int32"), V, T)
    f2, g2 = lib.eval_hs_grad([gp], np.zeros(2, dtype="int32"), V[:2], T)
    assert np.array_equal(f[:2], f2) and np.array_equal(g.reshape(K, -1)[:2], g2.reshape(2, -1))
    assert np.abs(f - (1 - np.abs(np.einsum("kij,ij->k", U, T.conj())) / N)).max() < 1e-12
