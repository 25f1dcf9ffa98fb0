"""LeapCompiler: repeated bounded A* with prefix fixing (reference: qsearch/leap_compiler.py:25-244).

Each round runs the sub-search from the current best circuit (all of its parameters are re-optimised, only the
structure is frozen, leap_compiler.py:105-108); a round ends early when a new best beats the linear-regression
prediction from the earlier bests of that round and improves on the overall best at depth >= min_depth
(leap_compiler.py:220-231).  Every expansion of the sub-search is one batched device solve.
"""
import heapq
from timeit import default_timer as timer

import numpy as np

from .compiler import Compiler, _better, _frontier, _qudits_of, _require_target, _solve_root
from .gates import ProductGate


def _linear_prediction(depths, values, at):
    """Least-squares line through (depths, values) evaluated at `at` (role of scipy.stats.linregress, :223)."""
    d, v = np.asarray(depths, dtype="float64"), np.asarray(values, dtype="float64")
    dm, vm = d.mean(), v.mean()
    den = ((d - dm) ** 2).sum()
    if den == 0:
        return float("nan")
    slope = ((d - dm) * (v - vm)).sum() / den
    return slope * at + (vm - slope * dm)


class SubCompiler(Compiler):
    """One bounded A* round (leap_compiler.py:122-244).  Returns ((circuit, parameters), value, depth)."""

    def compile(self, options=None):
        options = self.options.updated(options) if options is not None else self.options.copy()
        _require_target(options)
        logger, h, depth_limit = options.logger, options.heuristic, options.weight_limit
        n = _qudits_of(options)
        gateset = options.gateset
        initial_layer = options.initial_layer if "initial_layer" in options else gateset.initial_layer(n)
        search_layers = gateset.search_layers(n)
        root = initial_layer if isinstance(initial_layer, ProductGate) else ProductGate(initial_layer)
        stats = options.leap_stats if "leap_stats" in options else {"expansions": 0, "solves": 0}
        if len(search_layers) <= 0:
            x, value = _solve_root(root, options)
            return ((root, x), value, 0)
        parallel = options.parallelizer(options)
        beams = int(options.beams)
        if beams < 1:
            beams = max(1, int(options.num_tasks // len(search_layers)))
        t0 = options.overall_starttime if "overall_starttime" in options else timer()
        overall_best = options.overall_best_value if "overall_best_value" in options else 1.0
        try:
            x, best_value = _solve_root(root, options)
            stats["solves"] += 1
            best_pair, best_depth = (root, x), 0
            logger.logprint("New best! {} at depth 0".format(best_value))
            if depth_limit == 0:
                return (best_pair, best_value, 0)
            queue = [(h(root, x, 0, options, value=best_value), 0, best_value, -1, x, root)]
            tiebreaker = 0
            prev_depths, prev_values = [], []
            while queue and timer() - t0 <= options.timeout and best_value >= options.threshold:
                popped = [heapq.heappop(queue) for _ in range(min(beams, len(queue)))]
                new_steps = [(node[5].appending(layer), node[1], w) for layer, w in search_layers for node in popped]
                stats["expansions"] += 1
                stats["solves"] += len(new_steps)
                for step, result, cur_depth, weight, value in _frontier(parallel, options, new_steps):
                    new_depth = cur_depth + weight
                    if _better(value, new_depth, best_value, best_depth, options.threshold):
                        best_value, best_pair, best_depth = value, (step, result[1]), new_depth
                        logger.logprint("New best! score: {} at depth: {}".format(best_value, new_depth))
                        if len(prev_values) > 1:
                            predicted = _linear_prediction(prev_depths, prev_values, new_depth)
                            if not np.isnan(predicted) and best_value < overall_best and predicted - best_value < 0 and \
                                    ("min_depth" not in options or new_depth >= options.min_depth):
                                return (best_pair, best_value, best_depth)
                        prev_depths.append(best_depth)
                        prev_values.append(best_value)
                    if depth_limit is None or new_depth < depth_limit:
                        heapq.heappush(queue, (h(step, result[1], new_depth, options, value=value), new_depth, value,
                                               tiebreaker, result[1], step))
                        tiebreaker += 1
        finally:
            parallel.done()
        return (best_pair, best_value, best_depth)


class LeapCompiler(Compiler):
    """leap_compiler.py:61-110.  Extra option: min_depth (minimum depth of a round before it may end early)."""

    def compile(self, options=None):
        options = self.options.updated(options) if options is not None else self.options.copy()
        _require_target(options)
        logger = options.logger
        start = timer()
        n = _qudits_of(options)
        sub_cls = options.sub_compiler_class if "sub_compiler_class" in options else SubCompiler
        sub = sub_cls(options)
        total_depth, best_value, depths = 0, 1.0, [0]
        initial_layer = options.gateset.initial_layer(n)
        stats = {"expansions": 0, "solves": 0, "rounds": 0}
        best_pair = None
        while True:
            if timer() - start > options.timeout and best_pair is not None:
                break
            opts = options.updated(initial_layer=initial_layer, overall_starttime=start, overall_best_value=best_value,
                                   leap_stats=stats)
            best_pair, best_value, best_depth = sub.compile(opts)
            stats["rounds"] += 1
            total_depth += best_depth
            depths.append(total_depth)
            logger.logprint("LEAP round {}: depth {} value {}".format(stats["rounds"], total_depth, best_value))
            if best_value < options.threshold or best_depth == 0:
                break
            initial_layer = best_pair[0]
        logger.logprint("Finished all sub-compilations at depth {} with score {} after {} seconds.".format(
            total_depth, best_value, timer() - start))
        return {"structure": best_pair[0], "parameters": np.asarray(best_pair[1]), "cut_depths": depths,
                "distance": best_value, "stats": stats, "time": timer() - start}
