"""Shared parity checks: a Library (CUDA build on the B200, or the host emulation of the same kernel sources in the
CPU suite) against the reference-generated fixtures and the NumPy oracle.  Tolerance = north_star's 1e-12 max-abs."""
import numpy as np

from oracle import qs_oracle as O
from qsearch_b200 import lowering
from tests import _golden

TOL = 1e-12


def check_fixture_case(lib, case, tol=TOL):
    gp = lowering.lower(case.circuit)
    n = len(case.xs)
    poi = np.zeros(n, dtype="int32")
    xs = np.asarray(case.xs, dtype="float64")
    U = lib.eval_matrix([gp], poi, xs)
    U2, J = lib.eval_mat_jac([gp], poi, xs)
    P, N = gp.n_params, gp.dim
    J = J.reshape(n, P, N, N)
    for s, gold in enumerate(case.samples):
        assert np.abs(U[s] - gold["U"]).max() < tol
        assert np.abs(U2[s] - gold["U"]).max() < tol
        if "J" in gold:
            assert J[s].shape == gold["J"].shape
            if P:
                assert np.abs(J[s] - gold["J"]).max() < tol
        else:
            w = _golden.proj_weights(N)
            assert np.abs(J[s] @ w - gold["J_proj"]).max() < 10 * tol
    for tname, T in case.targets.items():
        d = lib.eval_distance([gp], poi, xs, T)
        for s, gold in enumerate(case.samples):
            assert abs(d[s] - gold["dist_" + tname]) < tol
        if not P:
            continue
        f, g = lib.eval_hs_grad([gp], poi, xs, T)
        r, Jr = lib.eval_residuals_jac([gp], poi, xs, T)
        g, Jr = g.reshape(n, P), Jr.reshape(n, P, 2 * N * N)
        for s, gold in enumerate(case.samples):
            assert abs(f[s] - gold["f_" + tname]) < tol
            assert np.abs(g[s] - gold["g_" + tname]).max() < tol
            assert np.abs(r[s] - gold["r_" + tname]).max() < tol
            if "Jr_" + tname in gold:
                assert np.abs(Jr[s].T - gold["Jr_" + tname]).max() < tol  # reference layout: (2N^2, P)
            else:
                w = _golden.proj_weights(2 * N * N)
                assert np.abs(Jr[s] @ w - gold["Jr_" + tname + "_proj"]).max() < 50 * tol


def check_against_oracle(lib, circuits, xs_per_circuit, T, tol=TOL):
    """One ragged batch (several structures, several vectors each) against the NumPy oracle."""
    gps = [lowering.lower(c) for c in circuits]
    poi, params = [], []
    for ci, xs in enumerate(xs_per_circuit):
        for x in xs:
            poi.append(ci)
            params.append(np.asarray(x, dtype="float64"))
    flat = np.concatenate(params) if params else np.zeros(0)
    N = gps[0].dim
    U = lib.eval_matrix(gps, poi, flat)
    U2, J = lib.eval_mat_jac(gps, poi, flat)
    f, g = lib.eval_hs_grad(gps, poi, flat, T)
    d = lib.eval_distance(gps, poi, flat, T)
    r, Jr = lib.eval_residuals_jac(gps, poi, flat, T)
    k = 0
    for i, (ci, x) in enumerate(zip(poi, params)):
        c, P = circuits[ci], gps[ci].n_params
        Uo, Jo = O.mat_jac(c, x)
        Jo = np.array(Jo).reshape(P, N, N)
        assert np.abs(U[i] - Uo).max() < tol and np.abs(U2[i] - Uo).max() < tol
        assert np.abs(J[k:k + P] - Jo).max() < tol
        fo, go = O.matrix_distance_squared_jac(T, Uo, list(Jo))
        assert abs(f[i] - fo) < tol and np.abs(g[k:k + P] - go).max() < tol
        assert abs(d[i] - O.matrix_distance_squared(T, Uo)) < tol
        assert np.abs(r[i] - O.matrix_residuals(T, Uo)).max() < tol
        assert np.abs(Jr[k:k + P].T - O.matrix_residuals_jac(T, Uo, list(Jo))).max() < tol
        k += P
    return len(poi)


def all_leaf_circuit(n):
    """Every parameterised one-qubit leaf, a constant leaf and CNOTs in both directions on n qubits."""
    from qsearch_b200 import gates
    leaves = [gates.XGate(), gates.YGate(), gates.ZGate(), gates.U1Gate(), gates.U2Gate(), gates.U3Gate(), gates.XZXZGate(),
              gates.ZXZXZGate(), gates.SXGate()]
    layers = []
    for s0 in range(0, len(leaves), n):
        row = leaves[s0:s0 + n]
        row += [gates.IdentityGate()] * (n - len(row))
        layers.append(gates.KroneckerGate(*row))
        q = (s0 // n) % (n - 1)
        layers.append(gates.NonadjacentCNOTGate(n, q, q + 1) if s0 % 2 == 0 else gates.NonadjacentCNOTGate(n, q + 1, q))
    layers.append(gates.KroneckerGate(*leaves[:n][::-1]))
    return gates.ProductGate(*layers)


def check_every_leaf_type(lib, n, tol=TOL):
    """mat_jac and the residual Jacobian of all_leaf_circuit(n) against the oracle, including all-zero angles."""
    from qsearch_b200 import unitaries
    rng = np.random.default_rng(11)
    c = all_leaf_circuit(n)
    gp = lowering.lower(c)
    P = c.num_inputs
    xs = rng.random((3, P)) * 2 * np.pi
    xs[2, :] = 0.0  # degenerate angles: sin = 0 in every gate
    T = unitaries.qft(1 << n)
    poi = np.zeros(3, dtype="int32")
    U, J = lib.eval_mat_jac([gp], poi, xs)
    r, Jr = lib.eval_residuals_jac([gp], poi, xs, T)
    for i in range(3):
        Uo, Jo = O.mat_jac(c, xs[i])
        assert np.abs(U[i] - Uo).max() < tol
        assert np.abs(J[i * P:(i + 1) * P] - np.array(Jo)).max() < tol
        assert np.abs(r[i] - O.matrix_residuals(T, Uo)).max() < tol
        assert np.abs(Jr[i * P:(i + 1) * P].T - O.matrix_residuals_jac(T, Uo, Jo)).max() < tol


def check_zero_overlap(lib, n):
    """S = Tr(T^dagger U) == 0 exactly: all-zero angles make the U3 layer the identity and T = X (x) I
    is traceless.  f = 1 and the gradient is +inf in every component, the native backend's convention
    (qsrs/src/utils.rs:144-146; the Python backend divides by zero, comparison.py:53).  Both HS kernel families."""
    from qsearch_b200 import gates, _lib
    c = gates.ProductGate(gates.KroneckerGate(*[gates.U3Gate() for _ in range(n)]))
    gp = lowering.lower(c)
    X = np.array([[0, 1], [1, 0]], dtype="complex128")
    T = np.kron(X, np.eye(1 << (n - 1)))
    x = np.zeros((1, c.num_inputs))
    for kind in (0, 1):
        b = lib.batch([gp], np.zeros(1, dtype="int32"))
        b.set_kernel(kind)
        b.set_params(x)
        b.set_target(T)
        b.run(_lib.RUN_HS)
        f, g, _ = b.fetch()
        b.close()
        assert f[0] == 1.0
        assert np.all(np.isposinf(g)), g
    assert lib.eval_distance([gp], [0], x, T)[0] == 1.0


def two_qubit_rotation_circuits(n):
    """Circuits mixing RXX / RYY / RZZ (qsrs/src/circuits.rs:457-633) with one-qubit leaves and CNOTs on n qubits, every
    adjacent qubit pair used by every rotation kind."""
    from qsearch_b200 import gates
    I, u3 = gates.IdentityGate(), gates.U3Gate
    def at(g2, q):  # two-qubit gate on (q, q + 1)
        return gates.KroneckerGate(*([I] * q + [g2] + [I] * (n - q - 2)))
    layers = [gates.KroneckerGate(*[u3() for _ in range(n)])]
    kinds = [gates.RXXGate, gates.RYYGate, gates.RZZGate]
    for q in range(n - 1):
        for k, cls in enumerate(kinds):
            layers.append(at(cls(), q))
            if (q + k) % 2 == 0:
                layers.append(gates.KroneckerGate(*[gates.XZXZGate() if i == (q + k) % n else I for i in range(n)]))
        layers.append(at(gates.CNOTGate(), q))
    full = gates.ProductGate(*layers)
    short = gates.ProductGate(layers[0], at(gates.RYYGate(), n - 2), at(gates.RZZGate(), 0), at(gates.RXXGate(), 0))
    return [full, short]


def check_two_qubit_rotations(lib, n, tol=TOL):
    from qsearch_b200 import unitaries
    rng = np.random.default_rng(40 + n)
    circuits = two_qubit_rotation_circuits(n)
    xs = [rng.random((2, c.num_inputs)) * 2 * np.pi for c in circuits]
    xs[1][1, :] = 0.0
    return check_against_oracle(lib, circuits, xs, unitaries.qft(1 << n), tol)
