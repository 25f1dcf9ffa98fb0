"""LEAP re-optimisation: another caller of the batched solve (reference: qsearch/post_processing.py:108-295).

LEAP freezes a prefix structure at every round, so the circuit it returns is longer than necessary around the
round boundaries (``cut_depths``).  ``LEAPReoptimizing_PostProcessor`` cuts a window of ``reoptimize_size`` layers
out of the circuit at each boundary midpoint (post_processing.py:165) and runs a bounded A* that re-inserts layers at
that position (``inserting(layer, depth=point)``, :250) until the whole circuit is below the threshold again; if
that took fewer layers than were removed the shorter circuit replaces the old one and the scan restarts (:272-283).
Every expansion of every window search is one batched device solve through the Parallelizer, exactly like the
forward searches.
"""
import heapq
from timeit import default_timer as timer

import numpy as np

from .compiler import Compiler, _better, _frontier, _qudits_of, _require_target, _solve_root
from .gates import ProductGate


class PostProcessor:
    """Interface (reference: qsearch/post_processing.py:28-44)."""

    def __init__(self, options=None):
        pass

    def post_process_circuit(self, result, options=None):
        return result


def _midpoints(cut_depths):
    # post_processing.py:165: first layer after the initial one, then the middle of every LEAP round
    return [1] + [pt + int((pt - prev) / 2) for pt, prev in zip(cut_depths[1:], cut_depths)]


class LEAPReoptimizing_PostProcessor(Compiler, PostProcessor):
    def post_process_circuit(self, result, options=None):
        options = self.options.updated(options) if options is not None else self.options.copy()
        size = options.weight_limit if "weight_limit" in options and options.weight_limit else options.reoptimize_size
        if repr(result["structure"]).count("CNOT") <= size or "cut_depths" not in result:  # post_processing.py:127-130
            return result
        return self.compile(options.updated(best_pair=(result["structure"], result["parameters"]),
                                            cut_depths=result["cut_depths"]))

    def _window_search(self, root, point, window, options, parallel, stats, t0):
        """Bounded A* that grows `root` by inserting layers at `point` (post_processing.py:186-268)."""
        logger, h = options.logger, options.heuristic
        search_layers = options.gateset.search_layers(_qudits_of(options))
        beams = int(options.beams)
        if beams < 1:
            beams = max(1, int(options.num_tasks // max(1, len(search_layers))))
        x, best_value = _solve_root(root, options)
        stats["solves"] += 1
        best_pair, best_depth = (root, x), 0
        queue = [(h(root, x, 0, options, value=best_value), 0, best_value, -1, x, root)]
        tiebreaker = 0
        while queue and timer() - t0 <= options.timeout and best_value >= options.threshold:
            popped = [heapq.heappop(queue) for _ in range(min(beams, len(queue)))]
            new_steps = [(node[5].inserting(layer, depth=point), node[1], w) for layer, w in search_layers for node in popped]
            stats["expansions"] += 1
            stats["solves"] += len(new_steps)
            for step, res, cur_depth, weight, value in _frontier(parallel, options, new_steps):
                new_depth = cur_depth + weight
                if _better(value, new_depth, best_value, best_depth, options.threshold):
                    best_value, best_pair, best_depth = value, (step, res[1]), new_depth
                    logger.logprint("New best! score: {} at depth: {}".format(best_value, new_depth))
                if new_depth < window - 1:  # only circuits that can still end up shorter stay in the queue (:258)
                    heapq.heappush(queue, (h(step, res[1], new_depth, options, value=value), new_depth, value, tiebreaker,
                                           res[1], step))
                    tiebreaker += 1
            if best_depth >= window - 1:  # :262
                break
        return best_pair, best_value

    def compile(self, options=None):
        options = self.options.updated(options) if options is not None else self.options.copy()
        _require_target(options)
        logger = options.logger
        window = options.weight_limit if "weight_limit" in options and options.weight_limit else options.reoptimize_size
        start = timer()
        best_pair = options.best_pair
        best_value = options.objective.gen_eval_func(best_pair[0], options)(best_pair[1])
        midpoints = _midpoints(options.cut_depths) if "cut_depths" in options else None
        stats = {"expansions": 0, "solves": 0, "windows": 0}
        parallel = options.parallelizer(options)
        try:
            improved = True
            while improved and timer() - start <= options.timeout:
                improved = False
                circuit = best_pair[0]
                depth = len(circuit._subgates) - 1
                points = list(midpoints) if midpoints is not None else list(range(1, depth))
                for point in points:
                    if timer() - start > options.timeout:
                        break
                    if point < 1 or point >= depth:
                        continue
                    stats["windows"] += 1
                    root = ProductGate(*circuit._subgates[:point], *circuit._subgates[point + window:])
                    pair, value = self._window_search(root, point, window, options, parallel, stats, start)
                    new_depth = len(pair[0]._subgates) - 1
                    if value < options.threshold and new_depth < depth:
                        logger.logprint("With starting point {} re-optimized from depth {} to depth {}".format(point, depth, new_depth))
                        best_pair, best_value = pair, value
                        if midpoints is not None:  # :277: keep the boundaries behind the window, shifted by the saving
                            midpoints = [i - (depth - new_depth) for i in midpoints if i - (point + window) > 0]
                        improved = True
                        break  # rescan on the shorter circuit (:283)
                    if midpoints is not None:
                        midpoints = [i for i in midpoints if i - (point + window) > 0]
        finally:
            parallel.done()
        return {"structure": best_pair[0], "parameters": np.asarray(best_pair[1]), "distance": best_value, "stats": stats,
                "time": timer() - start}
