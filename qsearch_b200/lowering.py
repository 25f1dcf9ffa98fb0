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
* any other zero-parameter gate -> a constant op with its matrix in the program's constant pool.

It accepts genuine ``qsearch.gates`` trees and ``qsearch_b200.gates`` trees alike.
"""
import numpy as np

OP_DTYPE = np.dtype([("opcode", "<i4"), ("q0", "<i4"), ("q1", "<i4"), ("param_off", "<i4"),
                     ("const_off", "<i4"), ("reserved", "<i4")])

OP_RX, OP_RY, OP_RZ, OP_U1, OP_U2, OP_U3, OP_XZXZ, OP_ZXZXZ, OP_CONST1, OP_CNOT, OP_CONSTK, OP_RXX, OP_RYY, OP_RZZ = range(1, 15)
MAX_QUBITS = 5

# two-qubit rotations of the native backend (qsrs/src/lib.rs:216-218): one op on qubits (q, q + 1), 1 parameter
_PARAM_LEAVES_2Q = {"RXXGate": OP_RXX, "RYYGate": OP_RYY, "RZZGate": OP_RZZ}

_PARAM_LEAVES = {
    "XGate": (OP_RX, 1), "YGate": (OP_RY, 1), "ZGate": (OP_RZ, 1), "U1Gate": (OP_U1, 1), "U2Gate": (OP_U2, 2),
    "U3Gate": (OP_U3, 3), "XZXZGate": (OP_XZXZ, 2), "ZXZXZGate": (OP_ZXZXZ, 3),
}


class GateProgram:
    """Flat program + constant pool for one circuit structure (host data only; picklable)."""

    __slots__ = ("n_qubits", "n_params", "ops", "consts", "key")

    def __init__(self, n_qubits, n_params, ops, consts, key=None):
        self.n_qubits = n_qubits
        self.n_params = n_params
        self.ops = ops
        self.consts = consts
        self.key = key

    @property
    def n_ops(self):
        return len(self.ops)

    def __repr__(self):
        return "GateProgram(n_qubits={}, n_params={}, n_ops={})".format(self.n_qubits, self.n_params, len(self.ops))


def _constant_matrix(gate):
    if hasattr(gate, "const_matrix"):
        return np.asarray(gate.const_matrix(), dtype="complex128")
    return np.asarray(gate.matrix([]), dtype="complex128")  # reference objects (lib.rs:259-268)


class _Builder:
    def __init__(self):
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
    if n < 1 or n > MAX_QUBITS:
        raise ValueError("Unsupported circuit: {} qubits (1..{} supported)".format(n, MAX_QUBITS))
    b = _Builder()
    b.walk(circuit, 0, 0)
    ops = np.array(b.ops, dtype=OP_DTYPE) if b.ops else np.zeros(0, dtype=OP_DTYPE)
    consts = np.concatenate(b.consts) if b.consts else np.zeros(0, dtype="float64")
    return GateProgram(n, int(circuit.num_inputs), ops, np.ascontiguousarray(consts, dtype="float64"))
