"""Loader for the reference-generated fixtures in tests/golden/ (see make_golden.py)."""
import os

import numpy as np

from qsearch_b200 import gates as G

GOLDEN_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

_NS = {name: getattr(G, name) for name in dir(G) if name.endswith("Gate")}
_NS["array"] = np.array


def rebuild(circ_repr):
    """Rebuild a circuit tree from the reference's repr() using the mirror classes."""
    return eval(circ_repr, {"__builtins__": {}}, dict(_NS))  # noqa: S307 - fixture strings written by make_golden.py


class Case:
    def __init__(self, blob, i):
        pre = "c{}_".format(i)
        self.repr = str(blob[pre + "repr"])
        if pre + "seed" in blob.files:  # a circuit with a CPIPhaseGate: its phases come from np.random (make_golden_qutrit.py)
            state = np.random.get_state()
            np.random.seed(int(blob[pre + "seed"]))
            self.circuit = rebuild(self.repr)
            np.random.set_state(state)
        else:
            self.circuit = rebuild(self.repr)
        self.xs = blob[pre + "x"]
        self.targets = {k[len(pre) + 2:]: blob[k] for k in blob.files if k.startswith(pre + "T_")}
        self.samples = []
        for s in range(len(self.xs)):
            sp = "{}s{}_".format(pre, s)
            self.samples.append({k[len(sp):]: blob[k] for k in blob.files if k.startswith(sp)})

    def __repr__(self):
        return self.repr[:60]


def load(name):
    blob = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    return [Case(blob, i) for i in range(int(blob["n_cases"]))]


def proj_weights(n):
    return np.random.default_rng(7).standard_normal(n)


ALL_FILES = ["gates_small", "gateset_layers", "microbench", "leap_long", "qutrit"]
