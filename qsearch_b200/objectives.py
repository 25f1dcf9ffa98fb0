"""Objective closures (reference: qsearch/objectives.py:37-61, MatrixDistanceObjective) evaluated on the GPU.

Each ``gen_*`` binds ``options.target`` and a circuit and returns ``f(parameters)`` exactly like the reference, but
the closure calls one fused device entry point (distance, HS value+gradient, residuals, residual Jacobian) instead of
``circuit.matrix`` / ``circuit.mat_jac`` followed by NumPy arithmetic (qsearch/comparison.py:8-72).  The circuit may be
a prepared ``CudaCircuit`` or any gate tree the lowering accepts.
"""
import numpy as np

from . import _lib
from .backends import program_for


class Objective:
    has_residuals = True  # False: gen_error_residuals* return None, so default_solver avoids least squares (solvers.py:51-52)
    # True: gen_eval_func is the Hilbert-Schmidt distance against device_target(options), i.e. exactly the value the
    # on-device solvers return with xopt, so the search may reuse it (parallelizers.solve_frontier).  A subclass with its
    # own evaluation function leaves this False and is evaluated through gen_eval_func like in the reference
    # (compiler.py:147, leap_compiler.py:213).
    device_distance = False

    def gen_eval_func(self, circuit, options):
        return self.gen_error_func(circuit, options)

    def device_target(self, options):
        """The N x N complex matrix the device kernels contract the circuit unitary with."""
        return np.ascontiguousarray(options.target, dtype="complex128")

    def __eq__(self, other):
        return type(self) == type(other)

    def __hash__(self):
        return hash(type(self).__name__)


class MatrixDistanceObjective(Objective):
    device_distance = True

    def gen_error_func(self, circuit, options):
        """f(x) = |1 - |Tr(T^dagger U)|/N|   (comparison.matrix_distance_squared, comparison.py:8-22)."""
        gp, T = program_for(circuit), np.asarray(options.target, dtype="complex128")

        def generated_error_func(parameters):
            return float(_lib.get_library().eval_distance([gp], [0], parameters, T)[0])
        return generated_error_func

    def gen_error_jac(self, circuit, options):
        """(f, grad f)   (comparison.matrix_distance_squared_jac, comparison.py:37-55)."""
        gp, T = program_for(circuit), np.asarray(options.target, dtype="complex128")

        def generated_error_jac(parameters):
            f, g = _lib.get_library().eval_hs_grad([gp], [0], parameters, T)
            return float(f[0]), g
        return generated_error_jac

    def gen_error_residuals(self, circuit, options):
        """r = [Re(U T^dagger - I); Im(U T^dagger)]   (comparison.matrix_residuals, comparison.py:57-64)."""
        gp, T = program_for(circuit), np.asarray(options.target, dtype="complex128")

        def generated_error_residuals(parameters):
            return _lib.get_library().eval_residuals_jac([gp], [0], parameters, T)[0][0]
        return generated_error_residuals

    def gen_error_residuals_jac(self, circuit, options):
        """J_r of shape (2N^2, P)   (comparison.matrix_residuals_jac, comparison.py:66-72)."""
        gp, T = program_for(circuit), np.asarray(options.target, dtype="complex128")

        def generated_error_residuals_jac(parameters):
            return _lib.get_library().eval_residuals_jac([gp], [0], parameters, T)[1].T
        return generated_error_residuals_jac

    # batched form used by the search loop: distances of many (circuit, parameters) pairs in one launch
    def eval_batch(self, circuits, parameter_vectors, options):
        gps = [program_for(c) for c in circuits]
        T = self.device_target(options)
        flat = np.concatenate([np.asarray(v, dtype="float64").ravel() for v in parameter_vectors]) if gps else np.zeros(0)
        return _lib.get_library().eval_distance(gps, np.arange(len(gps), dtype="int32"), flat, T)


class StateprepObjective(Objective):
    """State preparation (reference: qsearch/objectives.py:63-101; comparison.distance_with_initial_state / _jac,
    comparison.py:114-123): ``s = <A a | U b>`` with a = ``options.target_state``, b = ``options.initial_state``,
    A = ``options.target``; ``f = 1 - |s|`` and ``df/dv_k = -(Re s Re s_k + Im s Im s_k)/|s|``, ``s_k = <A a | J_k b>``.

    On the device this is the Hilbert-Schmidt kernel with a rank-one matrix in the target slot:
    ``conj(s) = sum_ij (A a)_i conj(b_j) conj(U_ij) = (1/N) sum_ij T'_ij conj(U_ij)`` with ``T' = N (A a) b^dagger``, so
    ``1 - |sum T' . conj(U)|/N = 1 - |s|`` and the gradient formula of comparison.py:37-55 applied to ``T'`` is the one
    above (the kernels never assume that the matrix in the target slot is unitary).  Like the reference it has no
    residual form (objectives.py:81-101 return None), so the default solver for it is BFGS_Jac_Solver."""
    has_residuals = False
    device_distance = True

    def device_target(self, options):
        a = np.asarray(options.target_state, dtype="complex128").ravel()
        b = np.asarray(options.initial_state, dtype="complex128").ravel()
        A = np.asarray(options.target, dtype="complex128")
        if A.shape != (a.size, a.size) or b.size != a.size:
            raise ValueError("target_state, initial_state and target must have matching sizes")
        return np.ascontiguousarray(a.size * np.outer(A @ a, np.conj(b)))

    def gen_error_func(self, circuit, options):
        gp, T = program_for(circuit), self.device_target(options)

        def generated_error_func(parameters):
            return float(_lib.get_library().eval_distance([gp], [0], parameters, T)[0])
        return generated_error_func

    def gen_error_jac(self, circuit, options):
        gp, T = program_for(circuit), self.device_target(options)

        def generated_error_jac(parameters):
            f, g = _lib.get_library().eval_hs_grad([gp], [0], parameters, T)
            return float(f[0]), g
        return generated_error_jac

    def gen_error_residuals(self, circuit, options):
        return None

    def gen_error_residuals_jac(self, circuit, options):
        return None

    def eval_batch(self, circuits, parameter_vectors, options):
        return MatrixDistanceObjective.eval_batch(self, circuits, parameter_vectors, options)
