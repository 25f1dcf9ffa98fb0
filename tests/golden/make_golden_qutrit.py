"""Qutrit (d = 3) fixtures FROM THE REFERENCE ITSELF (run in the build container only):

    python tests/golden/make_golden_qutrit.py        ->  tests/golden/qutrit.npz

Same format as make_golden.py (circuits stored as repr(), rebuilt class for class by tests/_golden.py).  Covers the
gates the reference's own tests list for d = 3 (tests/test_circuit.py:36,41, tests/test_gradients.py:25:
SingleQutritGate, CPIGate) plus CSUMGate, CPIPhaseGate, UpgradedConstantGate, every layer of the two qutrit gatesets
(gatesets.py:283-305) and deeper products on 1, 2 and 3 qutrits.

CPIPhaseGate draws its phases from np.random when it is constructed (gates.py:590) and its repr does not show them:
a fixture circuit holds at most one CPIPhaseGate and records the np.random seed set before it was built
(``c<i>_seed``); the loader sets the same seed before rebuilding.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)
sys.path.insert(0, "/root/reference")

from qsearch import gates as rg, gatesets as rgs, unitaries as ru  # noqa: E402  (the reference)
from qsearch_b200 import workloads  # noqa: E402
import make_golden  # noqa: E402


def main():
    rng = np.random.default_rng(2468)
    sq, csum, cpi = rg.SingleQutritGate(), rg.CSUMGate(), rg.CPIGate()
    ucnot = rg.UpgradedConstantGate(rg.CNOTGate())
    i3 = rg.IdentityGate(d=3)
    seeds = {}
    circuits = []

    def add(c, seed=None):
        if seed is not None:
            seeds[len(circuits)] = seed
        circuits.append(c)

    def seeded(seed, build):
        np.random.seed(seed)
        add(build(), seed)

    # leaves and constants
    for c in (sq, csum, cpi, ucnot, i3, rg.IdentityGate(qudits=2, d=3)):
        add(c)
    seeded(41, lambda: rg.CPIPhaseGate())
    # small trees
    add(rg.ProductGate(sq, sq))
    add(rg.KroneckerGate(sq, sq))
    add(rg.KroneckerGate(sq, i3, sq))
    add(rg.ProductGate(rg.KroneckerGate(sq, sq), csum, rg.KroneckerGate(sq, sq)))
    add(rg.KroneckerGate(i3, rg.ProductGate(cpi, rg.KroneckerGate(sq, sq))))
    add(rg.KroneckerGate(csum, sq))
    # every layer of the qutrit gatesets
    for n in (2, 3):
        gs = rgs.QutritCNOTLinear()
        add(gs.initial_layer(n))
        for layer, _ in gs.search_layers(n):
            add(layer)
    for n, seed in ((2, 51), (3, 52)):
        np.random.seed(seed)
        layers = rgs.QutritCPIPhaseLinear().search_layers(n)  # one CPIPhaseGate, shared by the layers of this call
        for layer, _ in layers:
            add(layer, seed)
    # deeper circuits, the shapes the search produces
    gs = rgs.QutritCNOTLinear()
    c = rg.ProductGate(gs.initial_layer(2))
    for _ in range(3):
        c = c.appending(gs.search_layers(2)[0][0])
    add(c)
    c = rg.ProductGate(gs.initial_layer(3))
    for k in range(4):
        c = c.appending(gs.search_layers(3)[k % 2][0])
    add(c)
    c = rg.ProductGate(rgs.fill_row(sq, 3), rg.KroneckerGate(csum, i3), rg.KroneckerGate(sq, i3, sq), rg.KroneckerGate(i3, cpi),
                       rgs.fill_row(sq, 3), rg.KroneckerGate(i3, ucnot))
    add(c)

    cases = []
    for c in circuits:
        N = 3 ** c.qudits
        xs = rng.random((3 if N <= 9 else 2, c.num_inputs)) * 2 * np.pi
        if c.num_inputs:
            xs[-1, :] = 0.0 if N == 3 else xs[-1, :]
        targets = {"haar": workloads.haar_unitary(N, 700 + N)}
        if N == 9:
            targets["qft"] = ru.qft(9)
        cases.append((c, xs, targets))
    path = os.path.join(HERE, "qutrit.npz")
    make_golden.pack(cases, path)
    blob = dict(np.load(path))
    for i, seed in seeds.items():
        blob["c{}_seed".format(i)] = np.array(seed)
        # the constant itself, so that the loader can verify it rebuilt the same gate
        blob["c{}_U0".format(i)] = np.array(circuits[i].matrix(np.zeros(circuits[i].num_inputs)))
    np.savez_compressed(path, **blob)
    print(path, os.path.getsize(path) // 1024, "KiB,", len(cases), "cases,", len(seeds), "seeded")


if __name__ == "__main__":
    main()
