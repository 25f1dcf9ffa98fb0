"""Pin the CPU oracle against outputs of the reference itself (tests/golden/*.npz). CPU only."""
import numpy as np
import pytest

from oracle import qs_oracle as O
from tests import _golden

TOL = 1e-13


def check_case(case, evaluate):
    """evaluate(circuit, x) -> (U, J[P,N,N]); objective quantities are derived with the oracle's comparison math."""
    for x, gold in zip(case.xs, case.samples):
        U, J = evaluate(case.circuit, x)
        J = np.array(J).reshape((-1,) + U.shape)
        assert np.abs(U - gold["U"]).max() < TOL
        if "J" in gold:
            assert J.shape == gold["J"].shape
            if J.size:
                assert np.abs(J - gold["J"]).max() < TOL
        else:
            w = _golden.proj_weights(U.shape[0])
            assert np.abs(J @ w - gold["J_proj"]).max() < 1e-12
        for tname, T in case.targets.items():
            f, g = O.matrix_distance_squared_jac(T, U, list(J))
            assert abs(f - gold["f_" + tname]) < TOL
            assert abs(O.matrix_distance_squared(T, U) - gold["dist_" + tname]) < TOL
            if J.size:
                assert np.abs(g - gold["g_" + tname]).max() < TOL
            assert np.abs(O.matrix_residuals(T, U) - gold["r_" + tname]).max() < TOL
            if J.size:
                Jr = O.matrix_residuals_jac(T, U, list(J))
                if "Jr_" + tname in gold:
                    assert np.abs(Jr - gold["Jr_" + tname]).max() < TOL
                else:
                    w = _golden.proj_weights(Jr.shape[0])
                    assert np.abs(w @ Jr - gold["Jr_" + tname + "_proj"]).max() < 1e-11


@pytest.mark.parametrize("fname", _golden.ALL_FILES)
def test_numpy_oracle_matches_reference_fixtures(fname):
    for case in _golden.load(fname):
        check_case(case, O.mat_jac)
        for x, gold in zip(case.xs, case.samples):
            assert np.abs(O.matrix(case.circuit, x) - gold["U"]).max() < TOL


def test_constant_targets_match_reference():
    import os
    from qsearch_b200 import unitaries as mu
    blob = np.load(os.path.join(_golden.GOLDEN_DIR, "targets.npz"))
    for name in ["toffoli", "fredkin", "peres", "logical_or", "HHL", "mirogate", "sqrt_x", "cnot", "swap", "sqrt_cnot"]:
        assert np.array_equal(getattr(mu, name), blob[name]), name
    for n in (8, 16, 32):
        assert np.array_equal(mu.qft(n), blob["qft{}".format(n)])
    for key in blob.files:
        if key.startswith("acnot_"):
            n, c, t = map(int, key.split("_")[1:])
            assert np.array_equal(mu.arbitrary_cnot(n, c, t), blob[key])


@pytest.mark.reference
def test_oracle_against_live_reference():
    """In the build container: compare the oracle with the live reference on fresh random inputs."""
    import sys
    sys.path.insert(0, "/root/reference")
    from qsearch import gatesets as rgs, gates as rg, comparison as rc, unitaries as ru
    rng = np.random.default_rng(99)
    for gs, n in [(rgs.QubitCNOTRing(), 4), (rgs.QubitCNOTLinear(), 3), (rgs.QubitCNOTRing(), 5)]:
        layers = gs.search_layers(n)
        c = rg.ProductGate(gs.initial_layer(n))
        for _ in range(6):
            c = c.appending(layers[int(rng.integers(len(layers)))][0])
        x = rng.random(c.num_inputs) * 2 * np.pi
        T = ru.qft(2 ** n)
        U, J = c.mat_jac(x)
        Uo, Jo = O.mat_jac(c, x)  # oracle walks the reference's own tree by class name
        assert np.abs(U - Uo).max() < TOL and max(np.abs(a - b).max() for a, b in zip(J, Jo)) < TOL
        assert np.abs(rc.matrix_residuals_jac(T, U, J) - O.matrix_residuals_jac(T, Uo, Jo)).max() < TOL
        f, g = rc.matrix_distance_squared_jac(T, U, J)
        fo, go = O.matrix_distance_squared_jac(T, Uo, Jo)
        assert abs(f - fo) < TOL and np.abs(g - go).max() < TOL


@pytest.mark.reference
def test_oracle_against_live_reference_on_qutrits():
    """The same on qutrit circuits (gates.py:411-602, gatesets.py:283-305): fresh random inputs, the reference's own trees."""
    import sys
    sys.path.insert(0, "/root/reference")
    from qsearch import gatesets as rgs, gates as rg, comparison as rc
    from qsearch_b200 import workloads
    rng = np.random.default_rng(101)
    np.random.seed(3)  # CPIPhaseGate draws its phases when it is made (gates.py:590); the oracle reads them off the object
    for gs, n in [(rgs.QutritCNOTLinear(), 2), (rgs.QutritCPIPhaseLinear(), 2), (rgs.QutritCNOTLinear(), 3)]:
        layers = gs.search_layers(n)
        c = rg.ProductGate(gs.initial_layer(n))
        for _ in range(3):
            c = c.appending(layers[int(rng.integers(len(layers)))][0])
        x = rng.random(c.num_inputs) * 2 * np.pi
        T = workloads.haar_unitary(3 ** n, 5 + n)
        U, J = c.mat_jac(x)
        Uo, Jo = O.mat_jac(c, x)
        assert np.abs(U - Uo).max() < TOL and max(np.abs(a - b).max() for a, b in zip(J, Jo)) < TOL
        assert np.abs(rc.matrix_residuals_jac(T, U, J) - O.matrix_residuals_jac(T, Uo, Jo)).max() < TOL
        f, g = rc.matrix_distance_squared_jac(T, U, J)
        fo, go = O.matrix_distance_squared_jac(T, Uo, Jo)
        assert abs(f - fo) < TOL and np.abs(g - go).max() < TOL


def test_aposmm_start_selection_matches_the_reference():
    """MultiStart_Solver's start points (qsearch/multistart_solvers.py:66-88): `solvers.aposmm_start_indices` against indices
    produced by the reference's own persistent_aposmm functions (tests/golden/make_golden_aposmm.py): 22 seeded samples,
    dimensions 1..165, including a tie for the minimum and a point closer to the boundary than mu."""
    import os
    from qsearch_b200.solvers import aposmm_start_indices
    z = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "aposmm.npz"))
    for k in range(int(z["n_cases"])):
        assert np.array_equal(aposmm_start_indices(z["x_%d" % k], z["f_%d" % k]), z["start_%d" % k]), k


def test_two_qubit_rotations_known_answers():
    """RXX / RYY / RZZ exist only in the reference's Rust backend (qsrs/src/circuits.rs:457-633), which cannot be built here,
    so the oracle's restatement is pinned by known answers instead of reference-generated fixtures: exp(-i x/2 P(x)P) by
    scipy's matrix exponential, and the derivative by central differences."""
    import scipy.linalg
    from oracle import qs_oracle as O
    from qsearch_b200 import gates
    paulis = {"RXXGate": np.array([[0, 1], [1, 0]]), "RYYGate": np.array([[0, -1j], [1j, 0]]), "RZZGate": np.diag([1.0, -1.0])}
    for name, Pm in paulis.items():
        g = getattr(gates, name)()
        PP = np.kron(Pm, Pm)
        for x in (0.0, 0.3, 2.1, 5.9):
            U, J = O.mat_jac(g, [x])
            assert np.abs(U - scipy.linalg.expm(-0.5j * x * PP)).max() < 1e-14
            assert np.abs(O.matrix(g, [x]) - U).max() == 0
            fd = (O.matrix(g, [x + 1e-6]) - O.matrix(g, [x - 1e-6])) / 2e-6
            assert len(J) == 1 and np.abs(J[0] - fd).max() < 1e-9
            assert np.abs(J[0] - (-0.5j) * PP @ U).max() < 1e-14
