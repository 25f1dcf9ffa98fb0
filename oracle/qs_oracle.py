"""CPU oracle (TEST INFRASTRUCTURE ONLY) for the qsearch hot path.

A NumPy restatement of the reference's NumPy backend for the path named in BASELINE.json:
circuit unitary (`matrix`), its parameter Jacobian (`mat_jac`), the Hilbert-Schmidt objective and
gradient, and the LM residual vector and residual Jacobian.  Every function cites the reference
file:line it restates (paths relative to /root/reference).

Pinning: tests/golden/*.npz hold outputs of the *reference itself* (generated in the build container
by tests/golden/make_golden.py, which imports /root/reference/qsearch); tests/test_oracle.py checks this
oracle against them on CPU.  The oracle is only ever imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  The product (qsearch_b200/) never imports it.

The oracle walks a circuit tree by class *name* (like qsrs' object_to_gate, qsrs/src/lib.rs:183-277), so it
accepts genuine qsearch.gates objects as well as qsearch_b200.gates objects.
"""
import numpy as np

SX = 0.5 * np.array([[1 + 1j, 1 - 1j], [1 - 1j, 1 + 1j]], dtype="complex128")  # unitaries.py:120-122


# ----------------------------------------------------------------------------- leaf gates
def _rz(t):  # unitaries.py:126-127
    return np.array([[np.exp(-0.5j * t), 0], [0, np.exp(0.5j * t)]], dtype="complex128")


def _rz_d(t):  # utils.py:113-115 (re_rot_z_jac), same as unitaries.py:129-130
    s, c = np.sin(t / 2), np.cos(t / 2)
    return np.array([[0.5 * (-s - 1j * c), 0], [0, 0.5 * (-s + 1j * c)]], dtype="complex128")


def _rx(t):  # unitaries.py:132-133
    c, s = np.cos(t / 2), np.sin(t / 2)
    return np.array([[c, -1j * s], [-1j * s, c]], dtype="complex128")


def _rx_d(t):  # unitaries.py:135-136
    c, s = np.cos(t / 2), np.sin(t / 2)
    return np.array([[-0.5 * s, -0.5j * c], [-0.5j * c, -0.5 * s]], dtype="complex128")


def _ry(t):  # unitaries.py:138-139
    c, s = np.cos(t / 2), np.sin(t / 2)
    return np.array([[c, -s], [s, c]], dtype="complex128")


def _ry_d(t):  # unitaries.py:141-142
    c, s = np.cos(t / 2), np.sin(t / 2)
    return np.array([[-0.5 * s, -0.5 * c], [0.5 * c, -0.5 * s]], dtype="complex128")


def _u3(v):  # gates.py:324-346
    ct, st = np.cos(v[0] / 2), np.sin(v[0] / 2)
    ep, el = np.cos(v[1]) + 1j * np.sin(v[1]), np.cos(v[2]) + 1j * np.sin(v[2])
    epl = ep * el
    U = np.array([[ct, -st * el], [st * ep, ct * epl]], dtype="complex128")
    J0 = np.array([[-0.5 * st, -0.5 * ct * el], [0.5 * ct * ep, -0.5 * st * epl]], dtype="complex128")
    J1 = np.array([[0, 0], [1j * st * ep, 1j * ct * epl]], dtype="complex128")
    J2 = np.array([[0, -1j * st * el], [0, 1j * ct * epl]], dtype="complex128")
    return U, [J0, J1, J2]


def _u2(v):  # gates.py:364-376
    k = 1 / np.sqrt(2)
    e0, e1 = np.exp(1j * v[0]), np.exp(1j * v[1])
    U = k * np.array([[1, -e1], [e0, e0 * e1]], dtype="complex128")
    J0 = k * np.array([[0, 0], [1j * e0, 1j * e0 * e1]], dtype="complex128")
    J1 = k * np.array([[0, -1j * e1], [0, 1j * e0 * e1]], dtype="complex128")
    return U, [J0, J1]


def _u1(v):  # gates.py:393-399 : exp(i v/2) * Rz(v) = diag(1, e^{iv})
    ph = np.exp(0.5j * v[0])
    U = ph * _rz(v[0])
    J = 0.5j * ph * _rz(v[0]) + ph * _rz_d(v[0])
    return U, [J]


def _xzxz(v):  # gates.py:282-304 : Rz(v1) SX Rz(v0) SX
    inner, inner_d = SX @ _rz(v[0]) @ SX, SX @ _rz_d(v[0]) @ SX
    return _rz(v[1]) @ inner, [_rz(v[1]) @ inner_d, _rz_d(v[1]) @ inner]


def _zxzxz(v):  # gates.py:219-254 : Rz(v2) SX Rz(v1) SX Rz(v0)
    def chain(a, b, c):
        return c @ SX @ b @ SX @ a
    z = [_rz(t) for t in v]
    d = [_rz_d(t) for t in v]
    return chain(*z), [chain(d[0], z[1], z[2]), chain(z[0], d[1], z[2]), chain(z[0], z[1], d[2])]


def _rxx(v):  # qsrs/src/circuits.rs:457-514 (GateRXX::mat / mat_jac; no Python twin in qsearch/gates.py)
    c, s = np.cos(v[0] / 2), -1j * np.sin(v[0] / 2)
    dc, ds = -np.sin(v[0] / 2) / 2, -1j * np.cos(v[0] / 2) / 2
    def m(c, s):
        return np.array([[c, 0, 0, s], [0, c, s, 0], [0, s, c, 0], [s, 0, 0, c]], dtype="complex128")
    return m(c, s), [m(dc, ds)]


def _ryy(v):  # qsrs/src/circuits.rs:516-578 (GateRYY)
    c, ns, ps = np.cos(v[0] / 2), -1j * np.sin(v[0] / 2), 1j * np.sin(v[0] / 2)
    dc, dns, dps = -np.sin(v[0] / 2) / 2, -1j * np.cos(v[0] / 2) / 2, 1j * np.cos(v[0] / 2) / 2
    def m(c, ns, ps):
        return np.array([[c, 0, 0, ps], [0, c, ns, 0], [0, ns, c, 0], [ps, 0, 0, c]], dtype="complex128")
    return m(c, ns, ps), [m(dc, dns, dps)]


def _rzz(v):  # qsrs/src/circuits.rs:580-633 (GateRZZ)
    pos, neg = np.exp(0.5j * v[0]), np.exp(-0.5j * v[0])
    return np.diag([neg, pos, pos, neg]).astype("complex128"), [np.diag([-0.5j * neg, 0.5j * pos, 0.5j * pos, -0.5j * neg]).astype("complex128")]


def _qutrit_entries(s, c, p, m):
    """The 3x3 of SingleQutritGate (gates.py:417-439) as a function of the sines s[0..2], cosines c[0..2] and the phase
    factors p[b] = e^{+i v[3+b]}, m[b] = e^{-i v[3+b]}, b = 0..4, all passed as independent variables."""
    return np.array([
        [c[0] * c[1] * p[0], s[0] * p[2], c[0] * s[1] * p[3]],
        [s[1] * s[2] * m[3] * m[4] - s[0] * c[1] * c[2] * p[0] * p[1] * m[2], c[0] * c[2] * p[1],
         -c[1] * s[2] * m[0] * m[4] - s[0] * s[1] * c[2] * p[1] * m[2] * p[3]],
        [-s[0] * c[1] * s[2] * p[0] * m[2] * p[4] - s[1] * c[2] * m[1] * m[3], c[0] * s[2] * p[4],
         c[1] * c[2] * m[0] * m[1] - s[0] * s[1] * s[2] * m[2] * p[3] * p[4]],
    ], dtype="complex128")


def _single_qutrit(v):  # gates.py:411-523 (SingleQutritGate.matrix / mat_jac)
    """The reference writes the eight Jacobian matrices out by hand (gates.py:441-517).  Here they follow from the matrix:
    every entry is affine in (sin v_a, cos v_a) and in (p_b, m_b), so with U(...) the entries as a function of those,
        dU/dv_a     = U(sin -> cos, cos -> -sin) - U(sin -> 0, cos -> 0)                       a = 0..2
        dU/dv_{3+b} = i [U(m_b -> 0) - U(p_b -> 0, m_b -> 0)] - i [U(p_b -> 0) - U(p_b -> 0, m_b -> 0)]   b = 0..4
    tests/golden/qutrit.npz pins both against the reference's hand-written matrices."""
    v = np.asarray(v, dtype="float64")
    s, c = np.sin(v[:3]), np.cos(v[:3])
    p, m = np.exp(1j * v[3:8]), np.exp(-1j * v[3:8])
    U = _qutrit_entries(s, c, p, m)
    J = []
    for a in range(3):
        s1, c1, s0, c0 = s.copy(), c.copy(), s.copy(), c.copy()
        s1[a], c1[a] = c[a], -s[a]
        s0[a] = c0[a] = 0.0
        J.append(_qutrit_entries(s1, c1, p, m) - _qutrit_entries(s0, c0, p, m))
    for b in range(5):
        p0, m0 = p.copy(), m.copy()
        p0[b] = m0[b] = 0.0
        base = _qutrit_entries(s, c, p0, m0)
        only_p, only_m = p0.copy(), m0.copy()
        only_p[b], only_m[b] = p[b], m[b]
        J.append(1j * (_qutrit_entries(s, c, only_p, m0) - base) - 1j * (_qutrit_entries(s, c, p0, only_m) - base))
    return U, J


_PARAM_LEAVES = {
    "SingleQutritGate": _single_qutrit,
    "XGate": lambda v: (_rx(v[0]), [_rx_d(v[0])]),
    "YGate": lambda v: (_ry(v[0]), [_ry_d(v[0])]),
    "ZGate": lambda v: (_rz(v[0]), [_rz_d(v[0])]),
    "U3Gate": _u3,
    "U2Gate": _u2,
    "U1Gate": _u1,
    "XZXZGate": _xzxz,
    "ZXZXZGate": _zxzxz,
    "RXXGate": _rxx,
    "RYYGate": _ryy,
    "RZZGate": _rzz,
}


def _const_of(gate):
    """Constant unitary of a zero-parameter leaf (gates.py:53-54 shortcut; lib.rs:259-268)."""
    if hasattr(gate, "const_matrix"):
        return np.asarray(gate.const_matrix(), dtype="complex128")
    return np.asarray(gate.matrix([]), dtype="complex128")


# ----------------------------------------------------------------------------- tree evaluation
def mat_jac(gate, v):
    """(U, [J_0..J_{P-1}]) for a circuit tree, same algorithm as the reference's NumPy backend."""
    v = np.asarray(v, dtype="float64")
    name = type(gate).__name__
    if name == "KroneckerGate":
        return _kron_mat_jac(gate, v)
    if name == "ProductGate":
        return _product_mat_jac(gate, v)
    if name in _PARAM_LEAVES:
        return _PARAM_LEAVES[name](v)
    if gate.num_inputs == 0:
        return _const_of(gate), []
    raise ValueError("Unknown gate {}".format(name))


def matrix(gate, v):
    """Circuit unitary (gates.py:856-868 Kronecker, :946-963 Product)."""
    v = np.asarray(v, dtype="float64")
    name = type(gate).__name__
    if name in ("KroneckerGate", "ProductGate"):
        subs, k, mats = gate._subgates, 0, []
        for g in subs:
            mats.append(matrix(g, v[k:k + g.num_inputs]))
            k += g.num_inputs
        U = mats[0]
        for M in mats[1:]:
            U = np.kron(U, M) if name == "KroneckerGate" else M @ U  # later factors multiply on the left (:959)
        return U
    return mat_jac(gate, v)[0]


def _sub_mat_jacs(gate, v):
    out, k = [], 0
    for g in gate._subgates:
        out.append(mat_jac(g, v[k:k + g.num_inputs]))
        k += g.num_inputs
    return out


def _kron_mat_jac(gate, v):  # gates.py:870-888
    U, jacs = None, []
    for M, Js in _sub_mat_jacs(gate, v):
        jacs = [np.kron(J, M) for J in jacs]
        jacs += [J if U is None else np.kron(U, J) for J in Js]
        U = M if U is None else np.kron(U, M)
    return U, jacs


def _product_mat_jac(gate, v):  # gates.py:965-1004
    subs = _sub_mat_jacs(gate, v)
    if len(subs) == 1:
        return subs[0]
    after = subs[0][0]
    for M, _ in subs[1:]:
        after = M @ after  # full product (:985-989)
    before = np.eye(after.shape[0], dtype="complex128")
    jacs = []
    for M, Js in subs:
        after = after @ M.conj().T  # peel the current factor off the "after" product (:992)
        jacs += [after @ (J @ before) for J in Js]  # (:993-995)
        before = M @ before  # (:996)
    return before, jacs


# ----------------------------------------------------------------------------- objective math
def matrix_distance_squared(A, B):  # comparison.py:8-22
    return np.abs(1 - np.abs(np.sum(A * np.conj(B))) / A.shape[0])


def matrix_distance_squared_jac(T, M, J):  # comparison.py:37-55
    S = np.sum(T * np.conj(M))
    f = 1 - np.abs(S) / T.shape[0]
    if len(J) == 0:
        return f, np.zeros(0)
    Sk = np.array([np.sum(T * np.conj(K)) for K in J])
    g = -(S.real * Sk.real + S.imag * Sk.imag) / (T.shape[0] * np.abs(S))
    return f, g


def matrix_residuals(T, U, I=None):  # comparison.py:57-64
    M = U @ T.conj().T
    return np.concatenate([(M.real - np.eye(T.shape[0])).ravel(), M.imag.ravel()])


def matrix_residuals_jac(T, M, J):  # comparison.py:66-72 ; returns shape (2N^2, P), a transposed view of [P][2N^2]
    Td = T.conj().T
    rows = np.array([np.concatenate([(K @ Td).real.ravel(), (K @ Td).imag.ravel()]) for K in J])
    return rows.T


# ----------------------------------------------------------------------------- state preparation objective
def distance_with_initial_state(stateA, stateB, A, B):  # comparison.py:114-115
    return np.abs(1 - np.abs(np.vdot(A @ stateA, B @ stateB)))


def distance_with_initial_state_jac(stateA, stateB, A, B, J):  # comparison.py:117-123
    si = A @ stateA
    s = np.vdot(si, B @ stateB)
    vu = np.array([np.vdot(si, K @ stateB) for K in J])
    g = -(s.real * vu.real + s.imag * vu.imag) / np.abs(s) if len(J) else np.zeros(0)
    return 1 - np.abs(s), g


# ----------------------------------------------------------------------------- the timed units
def eval_lm(gate, T, x):
    """One LM-form objective+Jacobian evaluation: r in R^{2N^2}, J_r in R^{2N^2 x P} (a8+a11+a12)."""
    # the reference's closures evaluate the circuit twice: matrix() for r, mat_jac() for J_r (objectives.py:50-61)
    return matrix_residuals(T, matrix(gate, x)), matrix_residuals_jac(T, *mat_jac(gate, x))


def eval_hs(gate, T, x):
    """One HS-form objective+Jacobian evaluation: f, grad f in R^P (a8+a10)."""
    U, J = mat_jac(gate, x)
    return matrix_distance_squared_jac(T, U, J)
