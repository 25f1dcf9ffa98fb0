"""Backend boundary: ``Backend.prepare_circuit(circ, options) -> circuit-like`` (reference: qsearch/backends.py:21-52).

``CudaBackend`` plays the role of the reference's ``NativeBackend`` (backends.py:49-52 -> qsrs.native_from_object,
qsrs/src/lib.rs:543-554): it lowers the Python gate tree to a flat gate program and returns a ``CudaCircuit`` that
evaluates on the B200 through libqsb200.so.  Like ``PyGateWrapper`` (lib.rs:280-369) the result exposes
``matrix(v)``, ``mat_jac(v)``, ``num_inputs``, ``dits`` and is *not* a ``Gate``; unsupported parameterised gates
raise ``ValueError`` (lib.rs:270-273).  There is no silent fallback to a CPU circuit (contrast
``SmartDefaultBackend``, backends.py:37-41).
"""
import hashlib

import numpy as np

from . import _lib
from .lowering import lower, GateProgram


class Backend:
    """Interface (reference: qsearch/backends.py:21-32)."""

    def __init__(self, options=None):
        pass

    def prepare_circuit(self, circ, options=None):
        raise NotImplementedError

    def __eq__(self, other):
        return type(self) == type(other)

    def __hash__(self):
        return hash(type(self).__name__)


_PROGRAM_CACHE = {}


def program_for(circ):
    """GateProgram of a circuit structure.  Structures without captured constants are cached by repr() - the reference's
    own notion of gate identity (qsearch/gates.py:82-86).  A tree that captures constant matrices (UGate, CUGate, any
    unknown zero-parameter gate: lowering.py CONST1 / CONSTK) is lowered every time and keyed by the exact bytes of its
    program and constant pool: NumPy's repr of a matrix keeps 8 digits and elides large arrays, so two circuits whose
    constants differ below that would otherwise share one device program."""
    if isinstance(circ, CudaCircuit):
        return circ.program
    key = repr(circ)
    gp = _PROGRAM_CACHE.get(key)
    if gp is None:
        gp = lower(circ)
        if gp.consts.size:
            gp.key = ("const", gp.radix, gp.n_qubits, gp.n_params, hashlib.sha1(np.ascontiguousarray(gp.ops).tobytes()).hexdigest(),
                      hashlib.sha1(gp.consts.tobytes()).hexdigest())
            return gp
        gp.key = key
        if len(_PROGRAM_CACHE) > 8192:
            _PROGRAM_CACHE.clear()
        _PROGRAM_CACHE[key] = gp
    return gp


class CudaCircuit:
    """A prepared circuit: flat gate program + evaluation on the GPU (cf. PyGateWrapper, qsrs/src/lib.rs:280-369)."""

    def __init__(self, source, program=None):
        self._source = source
        self.program = program if program is not None else program_for(source)

    # -- reference protocol -------------------------------------------------------------------------------
    @property
    def num_inputs(self):
        return self.program.n_params

    @property
    def qudits(self):
        return self.program.n_qubits

    @property
    def dits(self):
        return self.program.n_qubits

    def matrix(self, v):
        """complex128[N,N] (Gate.matrix, qsearch/gates.py:946-963)."""
        self._check(v)
        return _lib.get_library().eval_matrix([self.program], [0], v)[0]

    def mat_jac(self, v):
        """(U, J[P,N,N]) with J[k] = dU/dv_k (Gate.mat_jac, qsearch/gates.py:965-1004)."""
        self._check(v)
        U, J = _lib.get_library().eval_mat_jac([self.program], [0], v)
        return U[0], J

    # -- batched forms (many parameter vectors of this one structure) ------------------------------------------
    def matrix_batch(self, V):
        V = np.atleast_2d(np.asarray(V, dtype="float64"))
        return _lib.get_library().eval_matrix([self.program], np.zeros(len(V), dtype="int32"), V)

    def as_python(self):
        return self._source

    def _check(self, v):
        if np.size(v) != self.program.n_params:
            raise ValueError("expected {} parameters, got {}".format(self.program.n_params, np.size(v)))

    # device handles are never pickled: rebuild from the Python tree (cf. lib.rs:356-363)
    def __reduce__(self):
        return (CudaCircuit, (self._source,))

    def __repr__(self):
        return "CudaCircuit({})".format(repr(self._source))


class CudaBackend(Backend):
    """B200 backend.  Picklable (no state); CUDA is touched only when a prepared circuit is evaluated."""

    def prepare_circuit(self, circ, options=None):
        if isinstance(circ, CudaCircuit):  # already prepared (lib.rs:254-258; tests/test_circuit.py:60-62)
            return circ
        return CudaCircuit(circ)


# the role SmartDefaultBackend plays in the reference's defaults (defaults.py:104), without its silent fallback
DefaultBackend = CudaBackend
