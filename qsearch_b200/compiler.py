"""SearchCompiler: A* synthesis over circuit structures (reference: qsearch/compiler.py:27-164).

Same search as the reference - root solve, priority queue ordered by the heuristic, pop ``beams`` nodes, expand them
with ``gateset.successors``, keep the best (value, weight) pair, stop below ``threshold`` - but every expansion is ONE
batched device solve through the Parallelizer (children x start vectors), and the objective value of each child comes
back with the solve instead of being re-evaluated twice on the host (compiler.py:147,155).
Checkpointing / logging-to-file of the reference are host I/O outside the hot path and are not reproduced.
"""
import heapq
import math
from timeit import default_timer as timer

import numpy as np

from .defaults import standard_defaults, standard_smart_defaults
from .gates import ProductGate
from .options import Options


class Compiler:
    def __init__(self, options=None):
        self.options = Options()
        self.options.set_defaults(**standard_defaults)
        self.options.set_smart_defaults(**standard_smart_defaults)
        if options is not None:
            self.options = self.options.updated(options)

    def compile(self, options=None):
        raise NotImplementedError


def _require_target(options):
    """``target`` is required (compiler.py:62); a state-preparation Options object supplies it as a smart default
    (defaults.py:80-81,128) - the reference's Project passes ``options.target`` explicitly, same effect."""
    if "target_state" in options and "target" in options:
        options.target = options.target
    options.make_required("target")


def _qudits_of(options):
    U = options.unitary_preprocessor(options.target) if "unitary_preprocessor" in options else options.target
    d = options.gateset.d
    n = int(round(math.log(U.shape[0]) / math.log(d)))
    if d ** n != U.shape[0]:
        raise ValueError("The target matrix of size {} is not compatible with qudits of size {}.".format(U.shape[0], d))
    return n


def _frontier(parallel, options, new_steps):
    """(step, (U, x), depth, weight, value) for every child of one expansion."""
    if hasattr(parallel, "solve_frontier"):
        return parallel.solve_frontier(new_steps)
    ev = options.objective.gen_eval_func
    from .solvers import finite_or_inf
    return [(s, r, d, w, finite_or_inf(ev(s, options)(r[1]))) for s, r, d, w in parallel.solve_circuits_parallel(new_steps)]


def _better(value, weight, best_value, best_weight, threshold):
    """The reference's best-so-far rule (compiler.py:149)."""
    return (value < best_value and (best_value >= threshold or weight <= best_weight)) or \
           (value < threshold and weight < best_weight)


def _solve_root(root, options):
    solver = options.solver
    from .solvers import finite_or_inf, rides_on_device
    if hasattr(solver, "solve_batch"):
        res = solver.solve_batch([root], options)
        if rides_on_device(options):
            return res.x[0], res.distance[0]
        return res.x[0], finite_or_inf(options.objective.gen_eval_func(root, options)(res.x[0]))
    U, x = solver.solve_for_unitary(options.backend.prepare_circuit(root, options), options)
    return x, finite_or_inf(options.objective.gen_eval_func(root, options)(x))


class SearchCompiler(Compiler):
    """A* search (compiler.py:54-164).  Options: target (required), gateset, threshold, heuristic, solver,
    parallelizer, beams, weight_limit, objective, timeout, logger."""

    def compile(self, options=None):
        options = self.options.updated(options) if options is not None else self.options.copy()
        _require_target(options)
        logger, h, weight_limit = options.logger, options.heuristic, options.weight_limit
        start = timer()
        n = _qudits_of(options)
        gateset = options.gateset
        initial_layer = gateset.initial_layer(n)
        branching = gateset.branching_factor(n)
        root = initial_layer if isinstance(initial_layer, ProductGate) else ProductGate(initial_layer)
        if branching <= 0:
            x, value = _solve_root(root, options)
            return {"structure": root, "parameters": np.asarray(x), "distance": value, "weight": 0}
        parallel = options.parallelizer(options)
        beams = int(options.beams)
        if beams < 1:
            beams = max(1, int(options.num_tasks // branching))
        logger.logprint("branching factor {}, beams {}".format(branching, beams))
        stats = {"expansions": 0, "solves": 1}
        try:
            x, best_value = _solve_root(root, options)
            best_pair, best_weight = (root, x), 0
            logger.logprint("New best! {} at weight 0".format(best_value))
            if weight_limit == 0:
                return {"structure": root, "parameters": np.asarray(x), "distance": best_value, "weight": 0}
            queue = [(h(root, x, 0, options, value=best_value), 0, best_value, -1, x, root)]
            tiebreaker = 0
            while queue and timer() - start <= options.timeout and best_value >= options.threshold:
                popped = [heapq.heappop(queue) for _ in range(min(beams, len(queue)))]
                then = timer()
                new_steps = [(succ, node[1], w) for node in popped for succ, w in gateset.successors(node[5])]
                stats["expansions"] += 1
                stats["solves"] += len(new_steps)
                for step, result, cur_weight, weight, value in _frontier(parallel, options, new_steps):
                    new_weight = cur_weight + weight
                    if _better(value, new_weight, best_value, best_weight, options.threshold):
                        best_value, best_pair, best_weight = value, (step, result[1]), new_weight
                        logger.logprint("New best! score: {} at weight: {}".format(best_value, new_weight))
                    if weight_limit is None or new_weight < weight_limit:
                        heapq.heappush(queue, (h(step, result[1], new_weight, options, value=value), new_weight, value,
                                               tiebreaker, result[1], step))
                        tiebreaker += 1
                logger.logprint("Layer completed after {} seconds".format(timer() - then), verbosity=2)
        finally:
            parallel.done()
        logger.logprint("Finished compilation at weight {} with score {} after {} seconds.".format(
            best_weight, best_value, timer() - start))
        return {"structure": best_pair[0], "parameters": np.asarray(best_pair[1]), "distance": best_value,
                "weight": best_weight, "stats": stats, "time": timer() - start}
