"""Host-side circuit tree: the structural mirror of the reference's Gate classes.

The reference (qsearch/gates.py) stores a circuit as a tree of ``ProductGate`` / ``KroneckerGate``
nodes over leaf gates and evaluates it in NumPy.  Here the tree is *structure only*: class names,
``num_inputs``, ``qudits``, ``_subgates`` and ``__repr__`` match the reference (so the lowering in
``lowering.py`` accepts either this tree or a genuine ``qsearch.gates`` tree, dispatching on the class
name exactly like qsrs' ``object_to_gate``, qsrs/src/lib.rs:183-277), but no node contains matrix
arithmetic.  ``matrix(v)`` / ``mat_jac(v)`` lower the tree to a flat gate program and run it on the
GPU through ``CudaBackend``; there is no CPU evaluation path and a missing CUDA library raises.

Constant gates carry their (tiny) constant matrix as data for the lowering's constant pool.
"""
import numpy as np
from hashlib import md5

from . import unitaries, utils


class Gate:
    """Base protocol (reference: qsearch/gates.py:16-104)."""

    num_inputs = 0
    qudits = 1

    # -- evaluation goes through the CUDA backend -------------------------------------------
    def _prepared(self):
        circ = self.__dict__.get("_qsb_prepared")
        if circ is None:
            from .backends import CudaBackend

            circ = CudaBackend().prepare_circuit(self)
            self.__dict__["_qsb_prepared"] = circ
        return circ

    def matrix(self, v):
        """complex128[N,N] unitary for parameter vector v (reference contract: gates.py:26-37)."""
        return self._prepared().matrix(v)

    def mat_jac(self, v):
        """(U, [dU/dv_k]) (reference contract: gates.py:39-55)."""
        return self._prepared().mat_jac(v)

    # -- identity semantics: gates are immutable values compared by repr (gates.py:82-98) ---
    def __eq__(self, other):
        return repr(self) == repr(other)

    def __hash__(self):
        return int(md5(repr(self).encode()).hexdigest(), 16)

    def copy(self):
        return self

    def __copy__(self):
        return self

    def __deepcopy__(self, memo):
        return self

    def __getstate__(self):
        # device handles are never pickled (cf. PyGateWrapper.__reduce__, qsrs/src/lib.rs:356-363)
        d = dict(self.__dict__)
        d.pop("_qsb_prepared", None)
        return d

    def _parts(self):
        return [self]

    def validate_structure(self):
        return True

    def __repr__(self):
        return "{}()".format(type(self).__name__)


class _Param1Q(Gate):
    qudits = 1


class XGate(_Param1Q):
    """Rx(v0) (reference: gates.py:132-149)."""
    num_inputs = 1


class YGate(_Param1Q):
    """Ry(v0) (reference: gates.py:151-168)."""
    num_inputs = 1


class ZGate(_Param1Q):
    """Rz(v0) (reference: gates.py:170-187)."""
    num_inputs = 1


class ZXZXZGate(_Param1Q):
    """Rz(v2) SX Rz(v1) SX Rz(v0) (reference: gates.py:208-268)."""
    num_inputs = 3


class XZXZGate(_Param1Q):
    """Rz(v1) SX Rz(v0) SX (reference: gates.py:270-316)."""
    num_inputs = 2


class U3Gate(_Param1Q):
    """IBM U3(theta, phi, lambda) (reference: gates.py:318-356)."""
    num_inputs = 3


class U2Gate(_Param1Q):
    """IBM U2(phi, lambda) (reference: gates.py:358-385)."""
    num_inputs = 2


class U1Gate(_Param1Q):
    """IBM U1(lambda) = diag(1, e^{i lambda}) (reference: gates.py:387-409)."""
    num_inputs = 1


class SingleQutritGate(Gate):
    """Arbitrary single-qutrit gate: three angles and five phases (reference: gates.py:411-523).  Lowered to one
    QSB_OP_QUTRIT1 op of a radix-3 program."""
    num_inputs = 8
    qudits = 1


# ---------------------------------------------------------------------------- constant gates
class _Const(Gate):
    num_inputs = 0
    _const = None

    def const_matrix(self):
        """The constant unitary as host data (used only to fill the program's constant pool)."""
        return self._const


class IdentityGate(_Const):
    """Identity on `qudits` qudits (reference: gates.py:106-129)."""

    def __init__(self, qudits=1, d=2):
        self.qudits = qudits
        self._d = d
        self._const = np.array(np.eye(d ** qudits), dtype="complex128")

    def __repr__(self):
        if self.qudits == 1 and self._d == 2:
            return "IdentityGate()"
        return "IdentityGate(qudits={}, d={})".format(self.qudits, self._d)


class SXGate(_Const):
    _const = unitaries.sqrt_x


# ---------------------------------------------------------------------------- two-qutrit constants (d = 3)
def _perm9(images):
    """9x9 permutation matrix with M[images[j], j] = 1."""
    M = np.zeros((9, 9), dtype="complex128")
    for j, i in enumerate(images):
        M[i, j] = 1.0
    return M


class CSUMGate(_Const):
    """The reference's CSUM constant (gates.py:525-549): control 1 cycles the target b -> b + 1 mod 3, control 2 exchanges
    its levels 0 and 1 (the matrix as the reference defines it, not the textbook |a, a + b>)."""
    qudits = 2
    _const = _perm9([0, 1, 2, 4, 5, 3, 7, 6, 8])


class CPIGate(_Const):
    """Exchanges |1,0> and |1,1> (reference: gates.py:551-575)."""
    qudits = 2
    _const = _perm9([0, 1, 2, 4, 3, 5, 6, 7, 8])


class CPIPhaseGate(_Const):
    """CPI with a sign on |1,0> -> |1,1> and five random phases on the last five basis states, drawn when the gate is
    made (reference: gates.py:577-602; like there, every instance has its own phases and ``repr`` does not show them)."""
    qudits = 2

    def __init__(self):
        M = _perm9([0, 1, 2, 4, 3, 5, 6, 7, 8])
        M[3, 4] = -1.0
        phases = np.concatenate([np.ones(4), np.exp(2j * np.pi * np.random.random(5))])
        self._const = M @ np.diag(phases)


class UpgradedConstantGate(_Const):
    """A constant gate of a smaller qudit size acting on the low levels of larger qudits, identity elsewhere
    (reference: gates.py:761-788; the embedding is utils.upgrade_qudits)."""

    def __init__(self, other, df=3):
        if other.num_inputs > 0:
            raise AttributeError("UpgradedConstantGate is designed for only constant gates")
        OU = np.asarray(other.const_matrix() if hasattr(other, "const_matrix") else other.matrix([]), dtype="complex128")
        di = int(round(OU.shape[0] ** (1.0 / other.qudits)))
        if df <= di:
            raise AttributeError("Gate cannot be upgraded because it is already of an equal or higher dit level")
        self.df = df
        self.qudits = other.qudits
        self.subgate = other
        self._const = utils.upgrade_qudits(OU, di, df)

    def __repr__(self):
        return "UpgradedConstantGate({}, df={})".format(repr(self.subgate), self.df)


class _Param2Q(Gate):
    qudits = 2
    num_inputs = 1


class RXXGate(_Param2Q):
    """exp(-i v0/2 X(x)X) on two adjacent qubits.  The reference's native backend knows this gate by the class NAME
    "RXXGate" (object_to_gate, qsrs/src/lib.rs:216; matrix and derivative qsrs/src/circuits.rs:457-514) but qsearch/gates.py
    ships no Python class of that name; this one supplies the structure so that such circuits can be built and lowered."""


class RYYGate(_Param2Q):
    """exp(-i v0/2 Y(x)Y) (qsrs/src/lib.rs:217, qsrs/src/circuits.rs:516-578)."""


class RZZGate(_Param2Q):
    """exp(-i v0/2 Z(x)Z) (qsrs/src/lib.rs:218, qsrs/src/circuits.rs:580-633)."""


class _Const2Q(_Const):
    qudits = 2


class CNOTGate(_Const2Q):
    _const = unitaries.cnot


class CZGate(_Const2Q):
    _const = unitaries.cz


class ISwapGate(_Const2Q):
    _const = unitaries.iswap


class XXGate(_Const2Q):
    _const = unitaries.xx


class CNOTRootGate(_Const2Q):
    _const = unitaries.sqrt_cnot


class NonadjacentCNOTGate(_Const):
    """CNOT(control, target) embedded in a `qudits`-qubit register (reference: gates.py:696-728)."""

    def __init__(self, qudits, control, target):
        self.qudits = qudits
        self.control = control
        self.target = target

    def const_matrix(self):
        return unitaries.arbitrary_cnot(self.qudits, self.control, self.target)

    def validate_structure(self):
        return self.control < self.qudits and self.target < self.qudits and self.control != self.target

    def __repr__(self):
        return "NonadjacentCNOTGate({}, {}, {})".format(self.qudits, self.control, self.target)


class UGate(_Const):
    """Arbitrary constant unitary (reference: gates.py:730-759)."""

    def __init__(self, U, d=2, gatename="CUSTOM", gateparams=(), gateindices=None):
        self.d = d
        self.U = np.asarray(U, dtype="complex128")
        self.qudits = int(round(np.log(self.U.shape[0]) / np.log(d)))
        self.gatename = gatename
        self.gateparams = gateparams
        self.gateindices = gateindices
        self._const = self.U

    def __repr__(self):
        return "UGate({}, {}, {}, {}, {})".format(
            repr(self.U), repr(self.d), repr(self.gatename), repr(self.gateparams), repr(self.gateindices))


class CUGate(_Const):
    """Controlled-U on two qubits (reference: gates.py:791-824); `flipped` puts the control on the second qubit."""

    def __init__(self, U, gatename="Name", gateparams=(), flipped=False):
        U = np.asarray(U, dtype="complex128")
        n = U.shape[0]
        cu = np.zeros((2 * n, 2 * n), dtype="complex128")
        cu[:n, :n] = U if flipped else np.eye(n)
        cu[n:, n:] = np.eye(n) if flipped else U
        self._U = U
        self._const = cu
        self.gatename = gatename
        self.gateparams = gateparams
        self.flipped = flipped
        self.qudits = 2

    def __repr__(self):
        return "CUGate({}, {}{})".format(repr(self._U), repr(self.gatename), ", flipped=True" if self.flipped else "")


# ---------------------------------------------------------------------------- composite gates
class KroneckerGate(Gate):
    """Parallel composition: subgate 0 acts on the most significant qubits (reference: gates.py:845-933)."""

    def __init__(self, *subgates):
        self._subgates = tuple(subgates)
        self.num_inputs = sum(g.num_inputs for g in subgates)
        self.qudits = sum(g.qudits for g in subgates)

    def appending(self, gate):
        return KroneckerGate(*self._subgates, gate)

    def _parts(self):
        return self._subgates

    def validate_structure(self):
        return all(g.validate_structure() for g in self._subgates)

    def __repr__(self):
        return "KroneckerGate({})".format(repr(self._subgates)[1:-1])


class ProductGate(Gate):
    """Sequential composition: later subgates multiply on the left (reference: gates.py:935-1051)."""

    def __init__(self, *subgates):
        self._subgates = list(subgates)
        self.num_inputs = sum(g.num_inputs for g in subgates)
        self.qudits = 0 if len(subgates) == 0 else subgates[0].qudits

    def appending(self, *gates):
        return ProductGate(*self._subgates, *gates)

    def inserting(self, *gates, depth=-1):
        # same slicing convention as the reference (gates.py:1022-1030)
        return ProductGate(*self._subgates[:depth], *gates, *self._subgates[depth:])

    def validate_structure(self):
        return all(g.qudits == self.qudits and g.validate_structure() for g in self._subgates)

    def __repr__(self):
        return "ProductGate({})".format(repr(self._subgates)[1:-1])
