"""Search-queue priorities (reference: qsearch/heuristics.py:9-19).  Same signature ``(circ, v, weight, options)``;
an already known objective value can be passed as ``value`` to skip the re-evaluation the reference performs."""


def _value(circ, v, options, value):
    return options.objective.gen_eval_func(circ, options)(v) if value is None else value


def greedy(circ, v, weight, options, value=None):
    return _value(circ, v, options, value)


def astar(circ, v, weight, options, value=None):
    return 10.0 * _value(circ, v, options, value) + weight


def djikstra(circ, v, weight, options, value=None):
    return weight
