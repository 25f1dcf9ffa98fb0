"""Default option values of the search (reference: qsearch/defaults.py:89-122), with the GPU classes in the
slots the reference fills with its CPU ones."""
import numpy as np

from . import backends, gatesets, heuristics, objectives, parallelizers, solvers, utils

standard_defaults = {
    "threshold": 1e-10,
    "gateset": gatesets.Default(),
    "beams": -1,
    "delta": 0,
    "weight_limit": None,
    "objective": objectives.MatrixDistanceObjective(),
    "backend": backends.CudaBackend(),
    "parallelizer": parallelizers.GpuParallelizer,
    "max_quality_optimization": False,
    "unitary_preprocessor": utils.nearest_unitary,
    "timeout": float("inf"),
    "verbosity": 1,
    "stdout_enabled": False,
    "log_file": None,
    "seed": None,
    "reoptimize_size": 7,  # window of LEAPReoptimizing_PostProcessor as used by examples/leap_qft4.py:15
}


def default_heuristic(options):
    kind = options.search_type if "search_type" in options else "astar"
    return {"astar": heuristics.astar, "greedy": heuristics.greedy, "breadth": heuristics.djikstra,
            "djikstra": heuristics.djikstra}[kind]


def default_logger(options):
    return utils.Logger(options.stdout_enabled, options.log_file, options.verbosity)


standard_smart_defaults = {
    "solver": solvers.default_solver,
    "heuristic": default_heuristic,
    "logger": default_logger,
}


# state preparation (reference: qsearch/defaults.py:75-81,124-130): Options(target_state=..., defaults=stateprep_defaults,
# smart_defaults=stateprep_smart_defaults)
def stateprep_initial_state(options):
    v = np.zeros(np.asarray(options.target_state).shape, dtype="complex128")
    v[0] = 1
    return v


def stateprep_target(options):
    return np.eye(np.asarray(options.target_state).shape[0], dtype="complex128")


stateprep_defaults = {"objective": objectives.StateprepObjective()}
stateprep_smart_defaults = {"target": stateprep_target, "initial_state": stateprep_initial_state}
