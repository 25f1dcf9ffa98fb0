"""Fan-out of one frontier expansion (reference: qsearch/parallelizers.py:34-157).

The reference maps ``evaluate_step`` over the children of the popped nodes with one OS process per child
(parallelizers.py:34-36, 67-84).  ``GpuParallelizer`` instead lowers every child, submits ONE batched solve
(children x start vectors) that the C library splits into contiguous shards over the GPUs of its context (one stream
per GPU, no collective: SURVEY.md section 8e), and yields the reference's tuples ``(step, (U, xopt), depth, weight)``.
It runs in-process: no fork after CUDA initialisation (hard part 4).
"""
from . import _lib
from .solvers import finite_or_inf, rides_on_device


SM_PER_DEVICE = 148      # B200
MAX_BEAMS = 8            # nodes popped per expansion at most; beyond that the search order drifts from best-first
MAX_STARTS = 8           # start vectors per child at most (the frontier's wall time is its slowest instance)


def _num_devices():
    try:
        return _lib.get_library().num_devices
    except Exception:
        return 1


def _branching(opts):
    import math
    try:
        d = getattr(opts.gateset, "d", 2)
        return max(1, opts.gateset.branching_factor(int(round(math.log(opts.target.shape[0]) / math.log(d)))))
    except Exception:
        return 1


def default_num_tasks(opts):
    """The reference sets num_tasks = number of worker processes and derives beams = num_tasks // branching_factor
    (parallelizers.py:74, compiler.py:89-93), i.e. it pops as many nodes per expansion as its workers can solve
    concurrently.  The GPU equivalent of a worker is an SM: the batched LM solver runs one (child, start) instance per
    SM, so one expansion should put beams x branching x starts instances on the SMs of all visible GPUs.  Beams are
    capped at MAX_BEAMS; what is left goes to start vectors (`frontier_starts`)."""
    bf = _branching(opts)
    beams = (_num_devices() * SM_PER_DEVICE) // bf
    return bf * max(1, min(MAX_BEAMS, beams))


def frontier_starts(opts, n_children):
    """Start vectors per child of one expansion: whatever the solver asks for explicitly (LeastSquares_Jac_Solver(starts=k),
    k > 1, or options.frontier_starts), else as many as fill the SMs of all visible GPUs once, at most MAX_STARTS."""
    if "frontier_starts" in opts and opts.frontier_starts:
        return max(1, int(opts.frontier_starts))
    own = getattr(opts.solver, "starts", None)
    if own is None or own > 1:
        return None  # the solver decides (multistart solvers, explicit starts)
    slots = _num_devices() * SM_PER_DEVICE
    return max(1, min(MAX_STARTS, slots // max(1, n_children)))


class Parallelizer:
    def __init__(self, options):
        pass

    def solve_circuits_parallel(self, tuples):
        raise NotImplementedError

    def done(self):
        pass


class GpuParallelizer(Parallelizer):
    def __init__(self, options):
        self.options = options
        options.set_smart_defaults(num_tasks=default_num_tasks)

    def solve_frontier(self, tuples):
        """One batched solve for a whole expansion.  Returns (step, (U, xopt), depth, weight, distance): the final
        objective value rides along, which saves the two host re-evaluations per child the reference performs
        (compiler.py:147,155)."""
        tuples = list(tuples)
        if not tuples:
            return []
        options = self.options
        solver = options.solver
        steps = [t[0] for t in tuples]
        if hasattr(solver, "solve_batch"):
            res = solver.solve_batch(steps, options, starts=frontier_starts(options, len(steps)), frontier=True)
            if rides_on_device(options):
                values = res.distance
            else:  # a user objective with its own evaluation function: evaluate it like the reference (compiler.py:147)
                values = [finite_or_inf(options.objective.gen_eval_func(step, options)(res.x[i])) for i, step in enumerate(steps)]
            return [(step, (res.matrices[i], res.x[i]), depth, weight, values[i])
                    for i, (step, depth, weight) in enumerate(tuples)]
        # a solver without a batched form: evaluate_step one by one (parallelizers.py:34-36)
        backend, out = options.backend, []
        for step, depth, weight in tuples:
            U, x = solver.solve_for_unitary(backend.prepare_circuit(step, options), options)
            out.append((step, (U, x), depth, weight, finite_or_inf(options.objective.gen_eval_func(step, options)(x))))
        return out

    def solve_circuits_parallel(self, tuples):
        """The reference's contract (parallelizers.py:44-52): iterable of (step, (U, xopt), depth, weight)."""
        return [t[:4] for t in self.solve_frontier(tuples)]


class SequentialParallelizer(Parallelizer):
    """One solve at a time through Backend.prepare_circuit + Solver.solve_for_unitary (parallelizers.py:149-157)."""

    def __init__(self, options):
        self.options = options
        options.set_smart_defaults(num_tasks=lambda opts: 1)

    def solve_circuits_parallel(self, tuples):
        o = self.options
        return [(step, o.solver.solve_for_unitary(o.backend.prepare_circuit(step, o), o), depth, weight)
                for step, depth, weight in tuples]
