"""Solvers behind the reference's ``Solver.solve_for_unitary(circuit, options, x0=None) -> (U, xopt)`` interface
(qsearch/solvers.py:88-163), running as batched on-device optimisers (csrc/qsb_solver.cuh, csrc/qsb_lbfgs.cuh).

* ``LeastSquares_Jac_Solver``: Levenberg-Marquardt on the residuals ``[Re(U T^dagger - I); Im(U T^dagger)]``
  (role of scipy ``least_squares(method='lm')``, solvers.py:146-163, and of Ceres LM, qsrs/src/solvers.rs:88-121).
* ``BFGS_Jac_Solver``: L-BFGS on the Hilbert-Schmidt distance (role of scipy BFGS, solvers.py:138-144, and of NLopt
  LD_LBFGS, qsrs/src/solvers.rs:33-74).
* ``MultiStart_Solver`` / ``NaiveMultiStart_Solver``: the multistart axis of qsearch/multistart_solvers.py:32-144 as a
  batch axis: sample points are scored in one launch and the local solves run as one batch.

Every solver also has ``solve_batch(circuits, options, x0s=None)``: one launch for a whole frontier expansion
(children x starts), which is what ``GpuParallelizer`` calls.  Start vectors are uniform in [0, 2 pi) unless given
(solvers.py:155,157).  Solvers are stateless and picklable (constructor arguments only).
"""
import math

import os
import sys
import time

import numpy as np

from . import _lib
from . import objectives as _objectives
from .backends import program_for


def _target_of(options):
    """Matrix for the target slot of the kernels: ``options.target``, or what the objective derives from it
    (StateprepObjective: the rank-one matrix N (A a) b^dagger, objectives.py)."""
    obj = options.objective if "objective" in options else None
    if obj is not None and hasattr(obj, "device_target"):
        return obj.device_target(options)
    return np.ascontiguousarray(options.target, dtype="complex128")


def _has_residuals(options):
    obj = options.objective if "objective" in options else None
    return getattr(obj, "has_residuals", True)


def rides_on_device(options):
    """True if the distance the device solvers return IS options.objective's evaluation function (see
    objectives.Objective.device_distance); otherwise callers must evaluate objective.gen_eval_func themselves."""
    obj = options.objective if "objective" in options else None
    if obj is None:
        return True
    return bool(getattr(obj, "device_distance", False)) and \
        getattr(type(obj), "gen_eval_func", None) is getattr(_objectives.Objective, "gen_eval_func")


def finite_or_inf(v):
    """NaN (a failed solve) compares false with everything and would corrupt heap order: treat it as +inf."""
    v = float(v)
    return v if v == v else float("inf")


def _tolerances(options):
    # solvers.py:154-155: tighter tolerances under max_quality_optimization
    if "max_quality_optimization" in options and options.max_quality_optimization:
        return dict(ftol=5e-16, xtol=5e-16, gtol=1e-15)
    return dict(ftol=1e-8, xtol=1e-8, gtol=1e-8)


class BatchResult:
    """Best start per circuit of one batched solve."""
    __slots__ = ("x", "distance", "iterations", "status", "matrices", "all_iterations", "all_status")

    def __init__(self, x, distance, iterations, status, matrices, all_iterations=None, all_status=None):
        self.x, self.distance, self.iterations, self.status, self.matrices = x, distance, iterations, status, matrices
        self.all_iterations, self.all_status = all_iterations, all_status  # every (circuit, start) instance of the launch


class Solver:
    """Interface (reference: qsearch/solvers.py:88-104)."""
    method = None
    starts = 1

    def solve_for_unitary(self, circuit, options, x0=None):
        res = self.solve_batch([circuit], options, None if x0 is None else [np.asarray(x0, dtype="float64")])
        return (res.matrices[0], res.x[0])

    def _start_points(self, gp, options, rng, starts=None):
        return rng.random((int(starts) if starts else self.starts, gp.n_params)) * (2 * np.pi)

    def _solve_opts(self, options, frontier=False):
        return _tolerances(options)

    def solve_batch(self, circuits, options, x0s=None, want_matrices=True, starts=None, frontier=False):
        """Solve every circuit from `starts` start vectors each, in one launch; keep the best start per circuit.
        x0s: optional list (one entry per circuit) of a vector or an array [k, P] of start vectors.
        starts: overrides the solver's own number of random starts per circuit (the parallelizer sizes it to the GPUs).
        frontier: the batch is one expansion of a search, which only waits for its slowest instance: the solver may
        apply its batch-level stop rules (LeastSquares_Jac_Solver: straggler cap, sibling stop)."""
        lib = _lib.get_library()
        T = _target_of(options)
        gps = [program_for(c) for c in circuits]
        rng = np.random.default_rng(options.seed) if "seed" in options and options.seed is not None else np.random.default_rng()
        poi, chunks, owner = [], [], []
        for ci, gp in enumerate(gps):
            if x0s is not None and x0s[ci] is not None:
                pts = np.atleast_2d(np.asarray(x0s[ci], dtype="float64"))
            else:
                pts = self._start_points(gp, options, rng, starts)
            for p in pts:
                if p.size != gp.n_params:
                    raise ValueError("start vector of length {} for a circuit with {} parameters".format(p.size, gp.n_params))
                poi.append(ci)
                chunks.append(p)
                owner.append(ci)
        flat = np.concatenate(chunks) if chunks else np.zeros(0)
        trace = os.environ.get("QSB_TRACE_SOLVES")
        t0 = time.perf_counter() if trace else 0.0
        x, f, iters, status = lib.solve(self.method, gps, poi, flat, T, **self._solve_opts(options, frontier))
        if trace:  # one line per batched solve: size, parameter counts, wall, iteration tail, how the instances ended
            st = np.asarray(status)
            print("[qsb solve] {} instances={} programs={} P={}..{} wall_ms={:.1f} c_call_ms={:.1f} iters median={} max={} status={} best_f={:.2e}".format(
                self.method, len(poi), len(gps), min(g.n_params for g in gps), max(g.n_params for g in gps),
                1e3 * (time.perf_counter() - t0), getattr(lib, "last_solve_call_ms", float("nan")), int(np.median(iters)), int(np.max(iters)),
                {int(k): int((st == k).sum()) for k in np.unique(st)}, float(np.min(f))), file=sys.stderr, flush=True)
        owner = np.asarray(owner)
        offs = np.concatenate([[0], np.cumsum([gps[c].n_params for c in owner])]).astype(np.int64)
        best_x, best_f, best_it, best_st = [], [], [], []
        for ci in range(len(gps)):
            idx = np.nonzero(owner == ci)[0]
            fi = np.where(np.isnan(f[idx]), np.inf, f[idx])
            b = idx[int(np.argmin(fi))]
            best_x.append(x[offs[b]:offs[b + 1]].copy())
            best_f.append(finite_or_inf(f[b]))
            best_it.append(int(iters[b]))
            best_st.append(int(status[b]))
        mats = None
        if want_matrices and gps:
            mats = lib.eval_matrix(gps, np.arange(len(gps), dtype="int32"), np.concatenate(best_x) if best_x else np.zeros(0))
        return BatchResult(best_x, best_f, best_it, best_st, mats, np.asarray(iters), np.asarray(status))

    def __eq__(self, other):
        return type(self) == type(other) and self.__dict__ == other.__dict__

    def __hash__(self):
        return hash(type(self).__name__)


class LeastSquares_Jac_Solver(Solver):
    """Batched Levenberg-Marquardt on the device (qsb_solve_lm)."""
    method = "lm"
    # batch-level stop rules used when the batch is a frontier expansion (qsb_solver_opts, include/qsb200.h): an
    # instance is stopped once it has used FRONTIER_STRAGGLER x the Jacobian evaluations of the slowest instance among the
    # half of the batch that finished first, or once another start of its circuit reached options.threshold
    FRONTIER_STRAGGLER = 2.0

    def __init__(self, starts=1, max_iters=0, straggler_factor=None, sibling_stop=None, gram_tf32=None):
        self.starts = int(starts)
        self.max_iters = int(max_iters)
        self.straggler_factor = straggler_factor  # None: off for plain solves, FRONTIER_STRAGGLER for frontier batches
        self.sibling_stop = sibling_stop          # None: off for plain solves, options.threshold for frontier batches
        # J_r^T J_r (which only steers the optimiser) from TF32-rounded rows on the TF32 tensor path; the gradient, the
        # objective and all returned values stay fp64.  None: off for plain solves, on for frontier batches
        self.gram_tf32 = gram_tf32

    def _solve_opts(self, options, frontier=False):
        if not _has_residuals(options):
            # the reference's least-squares solver silently ignores the objective here (solvers.py:150); fail loudly
            raise ValueError("LeastSquares_Jac_Solver needs an objective with a residual form; {} has none".format(
                type(options.objective).__name__))
        o = _tolerances(options)
        o["max_iters"] = self.max_iters
        thr = options.threshold if "threshold" in options else 1e-10
        sf, ss = self.straggler_factor, self.sibling_stop
        o["straggler_factor"] = float(sf) if sf is not None else (self.FRONTIER_STRAGGLER if frontier else 0.0)
        o["sibling_stop"] = float(ss) if ss is not None else (float(thr) if frontier else 0.0)
        o["gram_tf32"] = float(bool(self.gram_tf32)) if self.gram_tf32 is not None else (1.0 if frontier else 0.0)
        return o

    @property
    def distance_metric(self):
        return "Residuals"


class BFGS_Jac_Solver(Solver):
    """Batched L-BFGS on the device (qsb_solve_lbfgs); memory 10 like the native solver (qsrs/src/lib.rs:402-413)."""
    method = "lbfgs"

    def __init__(self, starts=1, memory_size=10, max_iters=0):
        self.starts = int(starts)
        self.memory_size = int(memory_size)
        self.max_iters = int(max_iters)

    def _solve_opts(self, options, frontier=False):
        if not rides_on_device(options):
            # the reference's BFGS minimises objective.gen_error_jac (solvers.py:142); the device kernel minimises the
            # Hilbert-Schmidt distance against objective.device_target - refuse anything else instead of ignoring it
            raise ValueError("BFGS_Jac_Solver on the device needs MatrixDistanceObjective or StateprepObjective, got {}".format(
                type(options.objective).__name__))
        thr = options.threshold if "threshold" in options else 1e-10
        strict = "max_quality_optimization" in options and options.max_quality_optimization
        # scipy BFGS: gtol = options.threshold (solvers.py:142); NLopt: stopval 1e-16 (solvers.rs:64)
        return dict(gtol=1e-15 if strict else thr, ftol=1e-15, stop_value=1e-16, lbfgs_memory=self.memory_size,
                    max_iters=self.max_iters)

    @property
    def distance_metric(self):
        return "Frobenius"


class NaiveMultiStart_Solver(Solver):
    """`num_threads` uniform random starts of the inner solver, best result kept
    (reference: qsearch/multistart_solvers.py:112-144)."""

    def __init__(self, num_threads, inner_solver=None):
        self.num_threads = int(num_threads)
        self.inner_solver = inner_solver

    def _inner(self, options):
        inner = self.inner_solver
        if inner is None:
            inner = options.inner_solver if "inner_solver" in options else LeastSquares_Jac_Solver()
        return inner

    @property
    def distance_metric(self):
        return (self.inner_solver or LeastSquares_Jac_Solver()).distance_metric

    def solve_batch(self, circuits, options, x0s=None, want_matrices=True, starts=None, frontier=False):
        inner = self._inner(options)
        gps = [program_for(c) for c in circuits]
        rng = np.random.default_rng(options.seed) if "seed" in options and options.seed is not None else np.random.default_rng()
        pts_per_circuit = []
        for ci, gp in enumerate(gps):
            pts = rng.random((self.num_threads, gp.n_params)) * (2 * np.pi)
            if x0s is not None and x0s[ci] is not None:
                pts[0] = np.asarray(x0s[ci], dtype="float64")
            pts_per_circuit.append(pts)
        return inner.solve_batch(circuits, options, pts_per_circuit, want_matrices, frontier=frontier)


def aposmm_start_indices(x_on_cube, f, mu=1e-4):
    """Start points of APOSMM for ONE batch of uniform samples with no local-optimisation history: the case
    MultiStart_Solver creates (qsearch/multistart_solvers.py:66-88).  Restates, vectorised over the sample,
    persistent_aposmm.update_history_dist (:299-358: for every point the distance to the nearest other point whose
    value is not worse), decide_where_to_start_localopt (:421-541: keep points with no such point within r_k and at least
    `mu` away from the cube's boundary, sorted by value) and calc_rk (:544-556) with the default rk_const of
    initialize_APOSMM (:568).  Returns sample indices, best first."""
    x = np.asarray(x_on_cube, dtype="float64")
    f = np.asarray(f, dtype="float64")
    n_s, n = x.shape
    ok = ~np.isnan(f)
    sq = np.einsum("ij,ij->i", x, x)
    d = np.sqrt(np.maximum(sq[:, None] + sq[None, :] - 2.0 * (x @ x.T), 0.0))
    cand = (f[None, :] <= f[:, None]) & ok[None, :] & ok[:, None]  # cand[i, j]: j is not worse than i
    np.fill_diagonal(cand, False)
    dist_to_better = np.where(cand, d, np.inf).min(axis=1)
    if n_s and not np.isinf(dist_to_better).any():
        # the minimum is not unique, so no point is "best": the first one with the minimal value is (:361-368)
        dist_to_better[int(np.nonzero(f == np.nanmin(f))[0][0])] = np.inf
    rk_const = math.exp((math.lgamma(1.0 + n / 2.0) + math.log(5.0)) / n) / math.sqrt(math.pi)
    r_k = rk_const * (math.log(n_s) / n_s) ** (1.0 / n)
    to_bounds = np.minimum((1.0 - x).min(axis=1), x.min(axis=1))
    keep = ok & ~np.isinf(f) & (dist_to_better > r_k) & (to_bounds >= mu)
    idx = np.nonzero(keep)[0]
    return idx[np.argsort(f[idx], kind="stable")]


class MultiStart_Solver(NaiveMultiStart_Solver):
    """Sample-then-refine multistart (reference: qsearch/multistart_solvers.py:44-110): 100 uniform samples of the
    error function per circuit (ALL circuits of a frontier expansion scored in ONE launch), APOSMM's start selection on
    every circuit's sample (`aposmm_start_indices`: no better sample within r_k, multistart_solvers.py:85-88), then at
    most `num_threads` local solves per circuit from the selected points, run as one batch, argmin kept.  If fewer
    than `num_threads` points qualify, fewer solves are started, like the reference (:88)."""
    n_samples = 100

    def solve_batch(self, circuits, options, x0s=None, want_matrices=True, starts=None, frontier=False):
        inner = self._inner(options)
        lib = _lib.get_library()
        T = _target_of(options)
        gps = [program_for(c) for c in circuits]
        rng = np.random.default_rng(options.seed) if "seed" in options and options.seed is not None else np.random.default_rng()
        cubes = [rng.uniform(0.0, 1.0, (self.n_samples, gp.n_params)) for gp in gps]
        poi = np.repeat(np.arange(len(gps), dtype="int32"), self.n_samples)
        vals = lib.eval_distance(gps, poi, np.concatenate([(2 * np.pi * s).ravel() for s in cubes]) if gps else np.zeros(0), T)
        vals = vals.reshape(len(gps), self.n_samples)
        pts_per_circuit = []
        for ci in range(len(gps)):
            if gps[ci].n_params == 0:
                pts_per_circuit.append(np.zeros((1, 0)))
                continue
            order = aposmm_start_indices(cubes[ci], vals[ci])[: self.num_threads]
            if len(order) == 0:
                order = np.argsort(vals[ci])[:1]
            pts = 2 * np.pi * cubes[ci][order]
            if x0s is not None and x0s[ci] is not None:
                pts = np.vstack([np.asarray(x0s[ci], dtype="float64")[None, :], pts[:-1]]) if len(pts) > 1 else \
                    np.asarray(x0s[ci], dtype="float64")[None, :]
            pts_per_circuit.append(pts)
        return inner.solve_batch(circuits, options, pts_per_circuit, want_matrices, frontier=frontier)


def default_solver(options=None):
    """The reference picks the least-squares solver whenever the gateset has Jacobians and the objective has a residual
    form, else BFGS when it has a gradient (solvers.py:19-86).  Every gate the lowering accepts has a device Jacobian, so
    the choice only depends on the objective (StateprepObjective has no residuals, objectives.py:81-101)."""
    if options is not None and not _has_residuals(options):
        return BFGS_Jac_Solver()
    return LeastSquares_Jac_Solver()
