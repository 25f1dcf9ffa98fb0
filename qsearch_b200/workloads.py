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
