"""Generate golden fixtures FROM THE REFERENCE ITSELF (run in the build container only).

    python tests/golden/make_golden.py

Imports /root/reference/qsearch (NumPy backend; qsrs is not buildable here), evaluates the hot-path
functions on seeded inputs and writes small .npz fixtures next to this script.  Circuits are stored as their
repr() string, which qsearch_b200.gates reproduces class-for-class, so tests rebuild the same tree with
eval() in the mirror namespace.  The GPU box has no /root/reference: tests only read the .npz files.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, "/root/reference")

import qsearch  # noqa: E402  (the reference)
from qsearch import gates as rg, gatesets as rgs, comparison as rc, unitaries as ru, advanced_unitaries as rau  # noqa: E402
from qsearch_b200 import workloads, unitaries as mu  # noqa: E402


def ref_eval(circ, x, targets):
    """Everything the hot path produces for one (circuit, x), straight from the reference."""
    U = np.array(circ.matrix(x))
    U2, J = circ.mat_jac(x)
    J = np.array([np.array(j) for j in J]) if len(J) else np.zeros((0,) + U.shape, dtype="complex128")
    out = {"U": U, "J": J}
    assert np.abs(U - U2).max() < 1e-13
    for tname, T in targets.items():
        I = np.eye(T.shape[0])
        f, g = rc.matrix_distance_squared_jac(T, U2, list(J)) if len(J) else (rc.matrix_distance_squared(T, U), np.zeros(0))
        out["f_" + tname] = np.float64(f)
        out["dist_" + tname] = np.float64(rc.matrix_distance_squared(T, U))
        out["g_" + tname] = np.asarray(g, dtype="float64")
        out["r_" + tname] = rc.matrix_residuals(T, U, I)
        if len(J):
            out["Jr_" + tname] = np.ascontiguousarray(rc.matrix_residuals_jac(T, U2, list(J)))
    return out


def pack(cases, path, keep_full_J=True, full_J_cases=()):
    blob = {}
    for i, (circ, xs, targets) in enumerate(cases):
        blob["c{}_repr".format(i)] = np.array(repr(circ))
        blob["c{}_x".format(i)] = np.array(xs)
        for tname, T in targets.items():
            blob["c{}_T_{}".format(i, tname)] = T
        for s, x in enumerate(xs):
            res = ref_eval(circ, x, targets)
            for k, val in res.items():
                big = val.nbytes > (1100 * 1024 if (k == "J" and s == 0 and i in full_J_cases) else 48 * 1024)
                if (k == "J" or k.startswith("Jr_")) and (big or not keep_full_J):
                    # large tensors: keep a fixed random projection of the Jacobians instead of the full tensors
                    rs = np.random.default_rng(7)
                    w = rs.standard_normal(val.shape[-1] if k == "J" else val.shape[0])
                    val = val @ w if k == "J" else w @ val
                    k = k + "_proj"
                blob["c{}_s{}_{}".format(i, s, k)] = val
    blob["n_cases"] = np.array(len(cases))
    np.savez_compressed(path, **blob)
    print(path, os.path.getsize(path) // 1024, "KiB,", len(cases), "cases")


def main():
    rng = np.random.default_rng(12345)
    u, xzxz, zxzxz, cnot, ident = rg.U3Gate(), rg.XZXZGate(), rg.ZXZXZGate(), rg.CNOTGate(), rg.IdentityGate(1)

    # 1. the gate lists of the reference's own backend tests (tests/test_circuit.py:25-42, tests/test_gradients.py:11-27), d=2 only
    small = [
        rg.XGate(), rg.YGate(), rg.ZGate(), xzxz, zxzxz, u, rg.U2Gate(), rg.U1Gate(), ident, cnot,
        rg.ProductGate(u, xzxz, u), rg.KroneckerGate(u, xzxz, u), rg.ProductGate(u, zxzxz, u),
        rg.KroneckerGate(u, zxzxz, u), rg.KroneckerGate(u, u), rg.KroneckerGate(u, cnot, xzxz),
        rg.ProductGate(rg.KroneckerGate(u, u, u), rg.KroneckerGate(rg.ProductGate(cnot, rg.KroneckerGate(xzxz, u)), rg.IdentityGate())),
        rg.SXGate(), rg.CZGate(), rg.ISwapGate(), rg.XXGate(), rg.CNOTRootGate(), rg.NonadjacentCNOTGate(3, 2, 0),
        rg.KroneckerGate(rg.XGate(), rg.CZGate(), rg.YGate()), rg.ProductGate(rg.KroneckerGate(u, u), rg.ISwapGate(), rg.KroneckerGate(rg.ZGate(), rg.U2Gate())),
    ]
    cases = []
    for c in small:
        N = 2 ** c.qudits
        xs = rng.random((3, c.num_inputs)) * 2 * np.pi
        cases.append((c, xs, {"haar": workloads.haar_unitary(N, 900 + N)}))
    pack(cases, os.path.join(HERE, "gates_small.npz"))

    # 2. every layer of the standard gatesets (tests/test_circuit.py:48-58) + Yorktown adjacency (examples/qft5_ibm_yorktown.py:21)
    cases = []
    for gs, n in [(rgs.QubitCNOTLinear(), 4), (rgs.QubitCNOTRing(), 4), (rgs.QubitCNOTLinear(), 3), (rgs.QubitCNOTRing(), 3),
                  (rgs.QubitCNOTAdjacencyList([(0, 1), (0, 2), (1, 2), (2, 3), (2, 4), (3, 4)]), 5),
                  (rgs.U3CNOTLinear(), 3), (rgs.ZXZXZCNOTLinear(), 3), (rgs.QubitCZLinear(), 3), (rgs.QubitXXLinear(), 2)]:
        for layer in [gs.initial_layer(n)] + [l for l, _ in gs.search_layers(n)]:
            xs = rng.random((2, layer.num_inputs)) * 2 * np.pi
            cases.append((layer, xs, {"qft": ru.qft(2 ** n)}))
    pack(cases, os.path.join(HERE, "gateset_layers.npz"))

    # 3. the microbenchmark circuits (configs[3]); also proves the generator matches between reference and mirror classes
    cases = []
    for n in (3, 4, 5):
        ref_circs, _ = workloads.microbench_circuits(n, 3, 10, gatesets_module=rgs, gates_module=rg)
        mir_circs, mrng = workloads.microbench_circuits(n, 3, 10)
        assert [repr(c) for c in ref_circs] == [repr(c) for c in mir_circs]
        P = ref_circs[0].num_inputs
        params = workloads.microbench_params(mrng, 3, 2, P)
        targets = {"qft": ru.qft(2 ** n), "haar": workloads.haar_unitary(2 ** n, workloads.SEED_BASE + 100 + n)}
        assert np.array_equal(mu.qft(2 ** n), ru.qft(2 ** n))
        for ci, c in enumerate(ref_circs[: (3 if n < 5 else 2)]):
            cases.append((c, params[ci, : (2 if n < 5 else 1)], targets))
    pack(cases, os.path.join(HERE, "microbench.npz"), full_J_cases=(0, 3, 6))

    # 4. long LEAP-shaped circuits (P up to ~215) - Jacobians stored as projections to keep the fixture small
    cases = []
    gs = rgs.QubitCNOTAdjacencyList([(0, 1), (0, 2), (1, 2), (2, 3), (2, 4), (3, 4)])
    layers = gs.search_layers(5)
    c = rg.ProductGate(gs.initial_layer(5))
    for _ in range(40):
        c = c.appending(layers[int(rng.integers(len(layers)))][0])
    cases.append((c, rng.random((1, c.num_inputs)) * 2 * np.pi, {"qft": ru.qft(32)}))
    gs = rgs.QubitCNOTLinear()
    c = rg.ProductGate(gs.initial_layer(4))
    for _ in range(20):
        c = gs.successors(c)[int(rng.integers(2))][0]
    cases.append((c, rng.random((2, c.num_inputs)) * 2 * np.pi, {"qft": ru.qft(16)}))
    pack(cases, os.path.join(HERE, "leap_long.npz"), keep_full_J=False)

    # 5. constant targets of the README benchmark set
    np.savez_compressed(os.path.join(HERE, "targets.npz"), toffoli=ru.toffoli, fredkin=ru.fredkin, peres=ru.peres,
                        logical_or=ru.logical_or, full_adder=ru.full_adder, HHL=rau.HHL, mirogate=rau.mirogate,
                        qft8=ru.qft(8), qft16=ru.qft(16), qft32=ru.qft(32), sqrt_x=ru.sqrt_x, cnot=ru.cnot,
                        swap=ru.swap, sqrt_cnot=ru.sqrt_cnot,
                        **{"acnot_{}_{}_{}".format(n, c, t): ru.arbitrary_cnot(n, c, t)
                           for n in (2, 3, 4, 5) for c in range(n) for t in range(n) if c != t})
    print("done")


if __name__ == "__main__":
    main()
