"""Lowering: circuit tree -> flat gate program (host side).

The reference hands its Python tree to Rust by walking it recursively and dispatching on the class *name*
(``object_to_gate``, qsrs/src/lib.rs:183-277); unknown zero-parameter gates are captured through their constant
matrix (:259-268) and unknown parameterised gates raise ``ValueError`` (:270-273).  This module does the same
walk but emits a *flat* program of elementary ops acting on the full register (include/qsb200.h, ``qsb_op``):

* ``ProductGate``  -> its sub-programs one after another (later subgates multiply on the left, gates.py:959);
* ``KroneckerGate`` -> its sub-programs at increasing qubit offsets (they commute, so order is free; depth-first
  order keeps the parameter order of gates.py:861-864);
* parameterised one-qubit leaves -> one op carrying (qubit, first parameter index); the native backend's two-qubit
  rotations ``RXXGate`` / ``RYYGate`` / ``RZZGate`` (qsrs/src/lib.rs:216-218) -> one op on (qubit, qubit + 1);
* ``CNOTGate`` / ``NonadjacentCNOTGate`` -> a permutation op (no constant needed);
* ``IdentityGate`` -> nothing;
* any other zero-parameter gate -> a constant op with its matrix in the program's constant pool;
* qutrit circuits (d = 3: ``SingleQutritGate`` leaves and 3^k x 3^k constants such as ``CSUMGate``, ``CPIGate``,
  ``CPIPhaseGate``, ``UpgradedConstantGate``, gates.py:411-602, 761-788) -> a radix-3 program of ``OP_QUTRIT1`` and
  ``OP_CONSTK`` ops.  The radix is read off the leaves; a tree that mixes qubit and qutrit leaves raises ``ValueError``.

It accepts genuine ``qsearch.gates`` trees and ``qsearch_b200.gates`` trees alike.
"""
import numpy as np

OP_DTYPE = np.dtype([("opcode", "<i4"), ("q0", "<i4"), ("q1", "<i4"), ("param_off", "<i4"),
                     ("const_off", "<i4"), ("reserved", "<i4")])

(OP_RX, OP_RY, OP_RZ, OP_U1, OP_U2, OP_U3, OP_XZXZ, OP_ZXZXZ, OP_CONST1, OP_CNOT, OP_CONSTK, OP_RXX, OP_RYY, OP_RZZ,
 OP_QUTRIT1) = range(1, 16)
MAX_QUBITS = 5
MAX_QUTRITS = 3

# two-qubit rotations of the native backend (qsrs/src/lib.rs:216-218): one op on qubits (q, q + 1), 1 parameter
_PARAM_LEAVES_2Q = {"RXXGate": OP_RXX, "RYYGate": OP_RYY, "RZZGate": OP_RZZ}

_PARAM_LEAVES = {
    "XGate": (OP_RX, 1), "YGate": (OP_RY, 1), "ZGate": (OP_RZ, 1), "U1Gate": (OP_U1, 1), "U2Gate": (OP_U2, 2),
    "U3Gate": (OP_U3, 3), "XZXZGate": (OP_XZXZ, 2), "ZXZXZGate": (OP_ZXZXZ, 3),
}


class GateProgram:
    """Flat program + constant pool for one circuit structure (host data only; picklable)."""

    __slots__ = ("n_qubits", "n_params", "ops", "consts", "key", "radix")

    def __init__(self, n_qubits, n_params, ops, consts, key=None, radix=2):
        self.n_qubits = n_qubits  # number of qudits of size `radix`
        self.n_params = n_params
        self.ops = ops
        self.consts = consts
        self.key = key
        self.radix = radix

    @property
    def n_ops(self):
        return len(self.ops)

    @property
    def dim(self):
        """N: the programs' matrices are N x N."""
        return self.radix ** self.n_qubits

    def __getstate__(self):
        return {s: getattr(self, s) for s in self.__slots__}

    def __setstate__(self, st):
        self.radix = 2  # programs pickled before the radix existed
        for k, v in st.items():
            setattr(self, k, v)

    def __repr__(self):
        return "GateProgram(n_qubits={}, n_params={}, n_ops={}{})".format(
            self.n_qubits, self.n_params, len(self.ops), "" if self.radix == 2 else ", radix={}".format(self.radix))


def _constant_matrix(gate):
    if hasattr(gate, "const_matrix"):
        return np.asarray(gate.const_matrix(), dtype="complex128")
    return np.asarray(gate.matrix([]), dtype="complex128")  # reference objects (lib.rs:259-268)


def _int_root(dim, q):
    """r with r ** q == dim, or None."""
    r = int(round(dim ** (1.0 / q))) if q > 0 else 0
    for c in (r - 1, r, r + 1):
        if c >= 2 and c ** q == dim:
            return c
    return None


def circuit_radix(gate):
    """Qudit size of a circuit tree: 3 if its leaves are qutrit gates, 2 if they are qubit gates (None: no leaf says,
    e.g. an empty product).  Raises ValueError for a mixture or any other size."""
    found = set()

    def visit(g):
        name = type(g).__name__
        if name in ("ProductGate", "KroneckerGate"):
            for sub in g._subgates:
                visit(sub)
        elif name == "SingleQutritGate":
            found.add(3)
        elif name in _PARAM_LEAVES or name in _PARAM_LEAVES_2Q or name in ("CNOTGate", "NonadjacentCNOTGate"):
            found.add(2)
        elif g.num_inputs == 0:
            M = _constant_matrix(g)
            r = _int_root(M.shape[0], g.qudits) if M.ndim == 2 and M.shape[0] == M.shape[1] else None
            if r is None:
                raise ValueError("Unsupported gate {}: a {} matrix on {} qudits".format(name, M.shape, g.qudits))
            found.add(r)
        # unknown parameterised gates are reported by the walk itself

    visit(gate)
    if not found:
        return None
    if len(found) > 1 or not found <= {2, 3}:
        raise ValueError("Unsupported circuit: qudit sizes {} (all leaves must be qubit gates, or all qutrit gates)".format(sorted(found)))
    return found.pop()


class _Builder:
    def __init__(self, radix=2):
        self.radix = radix
        self.ops = []
        self.consts = []
        self.const_len = 0
        self.const_index = {}

    def add_const(self, M):
        key = M.tobytes()
        off = self.const_index.get(key)
        if off is None:
            flat = np.ascontiguousarray(M, dtype="complex128").view("float64").ravel()
            off = self.const_len
            self.consts.append(flat)
            self.const_len += flat.size
            self.const_index[key] = off
        return off

    def walk(self, gate, q_off, p_off):
        name = type(gate).__name__
        if name == "ProductGate":
            for sub in gate._subgates:
                self.walk(sub, q_off, p_off)
                p_off += sub.num_inputs
            return
        if self.radix == 3 and name not in ("KroneckerGate",):
            if name == "SingleQutritGate":
                if gate.num_inputs != 8:
                    raise ValueError("Unknown gate {}: unexpected parameter count".format(name))
                self.ops.append((OP_QUTRIT1, q_off, 0, p_off, 0, 0))
                return
            if gate.num_inputs != 0:
                raise ValueError("Unknown gate {} with {} parameters".format(name, gate.num_inputs))
            M = _constant_matrix(gate)
            dim = M.shape[0]
            if M.ndim != 2 or M.shape[1] != dim or dim != 3 ** gate.qudits:
                raise ValueError("Unsupported gate {}: expected a {}-qutrit matrix".format(name, gate.qudits))
            if not np.array_equal(M, np.eye(dim)):
                self.ops.append((OP_CONSTK, q_off, gate.qudits, 0, self.add_const(M), 0))
            return
        if name == "KroneckerGate":
            for sub in gate._subgates:
                self.walk(sub, q_off, p_off)
                q_off += sub.qudits
                p_off += sub.num_inputs
            return
        if name in _PARAM_LEAVES:
            opcode, np_ = _PARAM_LEAVES[name]
            if gate.num_inputs != np_:
                raise ValueError("Unknown gate {}: unexpected parameter count".format(name))
            self.ops.append((opcode, q_off, 0, p_off, 0, 0))
            return
        if name in _PARAM_LEAVES_2Q:
            if gate.num_inputs != 1 or gate.qudits != 2:
                raise ValueError("Unknown gate {}: unexpected shape".format(name))
            self.ops.append((_PARAM_LEAVES_2Q[name], q_off, q_off + 1, p_off, 0, 0))
            return
        if name == "CNOTGate":
            self.ops.append((OP_CNOT, q_off, q_off + 1, 0, 0, 0))
            return
        if name == "NonadjacentCNOTGate":
            self.ops.append((OP_CNOT, q_off + gate.control, q_off + gate.target, 0, 0, 0))
            return
        if gate.num_inputs != 0:
            raise ValueError("Unknown gate {} with {} parameters".format(name, gate.num_inputs))
        M = _constant_matrix(gate)
        dim = M.shape[0]
        if M.ndim != 2 or M.shape[1] != dim or dim != (1 << gate.qudits):
            raise ValueError("Unsupported gate {}: only qubit (d = 2) gates are supported".format(name))
        if name == "IdentityGate" or (dim <= 32 and np.array_equal(M, np.eye(dim))):
            return
        if dim == 2:
            self.ops.append((OP_CONST1, q_off, 0, 0, self.add_const(M), 0))
        else:
            self.ops.append((OP_CONSTK, q_off, gate.qudits, 0, self.add_const(M), 0))


def lower(circuit):
    """Lower a circuit tree to a GateProgram.  Raises ValueError for unsupported gates (no CPU fallback)."""
    n = int(circuit.qudits)
    radix = circuit_radix(circuit) or 2
    if radix == 3 and (n < 1 or n > MAX_QUTRITS):
        raise ValueError("Unsupported circuit: {} qutrits (1..{} supported)".format(n, MAX_QUTRITS))
    if n < 1 or n > MAX_QUBITS:
        raise ValueError("Unsupported circuit: {} qubits (1..{} supported)".format(n, MAX_QUBITS))
    b = _Builder(radix)
    b.walk(circuit, 0, 0)
    ops = np.array(b.ops, dtype=OP_DTYPE) if b.ops else np.zeros(0, dtype=OP_DTYPE)
    consts = np.concatenate(b.consts) if b.consts else np.zeros(0, dtype="float64")
    return GateProgram(n, int(circuit.num_inputs), ops, np.ascontiguousarray(consts, dtype="float64"), radix=radix)
