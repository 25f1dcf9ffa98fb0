"""Synthetic inputs for the batched objective+Jacobian microbenchmark (BASELINE.json configs[3]).

Definition: for n in {4, 5}: QubitCNOTRing; rng = default_rng(20261104 + n);
C circuits, each = initial layer + `depth` search layers drawn uniformly from search_layers(n);
V parameter vectors per circuit, uniform in [0, 2*pi).  Targets: qft(2^n) and a seeded Haar-random unitary.

`gatesets_module` may be this package's gatesets (default) or the reference's qsearch.gatesets: the same
draws produce trees with identical repr (checked by tests/golden/make_golden.py).
"""
import numpy as np

SEED_BASE = 20261104


def microbench_circuits(n, n_circuits=100, depth=10, gatesets_module=None, gates_module=None, rng=None):
    if gatesets_module is None:
        from . import gatesets as gatesets_module
    if gates_module is None:
        from . import gates as gates_module
    if rng is None:
        rng = np.random.default_rng(SEED_BASE + n)
    g = gatesets_module.QubitCNOTRing()
    layers = g.search_layers(n)
    circuits = []
    for _ in range(n_circuits):
        c = gates_module.ProductGate(g.initial_layer(n))
        for _ in range(depth):
            c = c.appending(layers[int(rng.integers(len(layers)))][0])
        circuits.append(c)
    return circuits, rng


def microbench_params(rng, n_circuits, n_vectors, n_params):
    return rng.random((n_circuits, n_vectors, n_params)) * (2 * np.pi)


def haar_unitary(dim, seed):
    """Seeded Haar-random unitary: Q of the QR of a complex Gaussian, phases fixed by diag(R)."""
    rng = np.random.default_rng(seed)
    z = (rng.standard_normal((dim, dim)) + 1j * rng.standard_normal((dim, dim))) / np.sqrt(2)
    q, r = np.linalg.qr(z)
    d = np.diag(r)
    return q * (d / np.abs(d))


def microbench_workload(n, n_circuits=100, n_vectors=1000, depth=10):
    """(circuits, params[C,V,P], targets dict) for one n."""
    from . import unitaries
    circuits, rng = microbench_circuits(n, n_circuits, depth)
    P = circuits[0].num_inputs
    params = microbench_params(rng, n_circuits, n_vectors, P)
    targets = {"qft": unitaries.qft(1 << n), "haar": haar_unitary(1 << n, SEED_BASE + 100 + n)}
    return circuits, params, targets


YORKTOWN = [(0, 1), (0, 2), (1, 2), (2, 3), (2, 4), (3, 4)]  # examples/qft5_ibm_yorktown.py:21


def frontier_workload(n, n_parents=None, depth=None, starts=4, seed=0):
    """One frontier expansion as the search produces it (compiler.py:136-146, leap_compiler.py:202-212): `n_parents`
    (default: 48 children) seeded parent circuits of `depth` layers, each extended by every search layer of the gateset
    (n = 4: QubitCNOTLinear, 3 layers; n = 5: the Yorktown adjacency list, 6 layers), `starts` uniform start vectors per
    child.  Returns (children, x0s [per child: (starts, P)], targets): targets["unreachable"] = qft(2^n) (no child reaches
    it at this depth: the regime the search spends its time in), targets["reachable"] = (child 0, seeded parameters)."""
    from . import gates, gatesets, unitaries
    rng = np.random.default_rng(SEED_BASE + 7000 + 10 * n + seed)
    g = gatesets.QubitCNOTLinear() if n <= 4 else gatesets.QubitCNOTAdjacencyList(YORKTOWN)
    layers = g.search_layers(n)
    if depth is None:
        depth = 10 if n <= 4 else 12
    if n_parents is None:
        n_parents = max(1, 48 // len(layers))
    children = []
    for _ in range(n_parents):
        c = gates.ProductGate(g.initial_layer(n))
        for _ in range(depth - 1):
            c = c.appending(layers[int(rng.integers(len(layers)))][0])
        children += [c.appending(layer) for layer, _ in layers]
    x0s = [rng.random((starts, c.num_inputs)) * (2 * np.pi) for c in children]
    x_star = rng.random(children[0].num_inputs) * (2 * np.pi)
    return children, x0s, {"unreachable": unitaries.qft(1 << n), "reachable": (children[0], x_star)}
