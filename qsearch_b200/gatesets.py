"""Gatesets / topologies: which circuit shapes the search hands to the hot path.

Host-side structure only; mirrors the interface of qsearch/gatesets.py (`initial_layer`, `search_layers`,
`branching_factor`, `successors`) so the resulting trees - and therefore parameter order and the lowered
gate programs - are identical to the reference's.
"""
from .gates import (CNOTGate, CZGate, ISwapGate, XXGate, IdentityGate, KroneckerGate, NonadjacentCNOTGate,
                    ProductGate, U3Gate, XZXZGate, ZXZXZGate, SingleQutritGate, CPIPhaseGate, UpgradedConstantGate)


def fill_row(gate, n):
    """One copy of `gate` on every qubit (reference: gatesets.py:317-318)."""
    return KroneckerGate(*[gate] * n)


def linear_topology(double_gate, single_gate, n, d=2, identity_gate=None, single_alt=None,
                    double_weight=1, single_weight=0, skip_index=None):
    """Two-qubit block on every adjacent pair (i, i+1) (reference: gatesets.py:308-315).

    Each layer is Kron(I*i, Product(double_gate, Kron(first, single_gate)), I*(n-i-2)) where `first` is
    single_alt (control side) when given, else single_gate.
    """
    ident = identity_gate if identity_gate else IdentityGate(d=d)
    first = single_gate if single_alt is None else single_alt
    weight = double_weight + 2 * single_weight
    block = ProductGate(double_gate, KroneckerGate(first, single_gate))
    return [(KroneckerGate(*[ident] * i, block, *[ident] * (n - i - 2)), weight)
            for i in range(n - 1) if skip_index is None or i != skip_index]


def find_last_3_cnots_linear(circuit):
    """Index i if the last three layers all sit on pair (i, i+1), else None (reference: gatesets.py:321-334)."""
    layers = circuit._subgates
    if len(layers) < 4:
        return None
    pos = [[type(s).__name__ for s in layer._subgates].index("ProductGate") for layer in layers[-1:-4:-1]]
    return pos[0] if pos[0] == pos[1] == pos[2] else None


class Gateset:
    """Interface (reference: gatesets.py:26-97)."""
    d = 2

    def initial_layer(self, qudits):
        raise NotImplementedError

    def search_layers(self, qudits):
        return []

    def branching_factor(self, qudits):
        return len(self.search_layers(qudits))

    def successors(self, circ, qudits=None):
        if qudits is None:
            qudits = circ.qudits
        return [(circ.appending(layer), w) for layer, w in self.search_layers(qudits)]

    def __eq__(self, other):
        return self is other or (type(self) == type(other) and self.__dict__ == other.__dict__)

    def __hash__(self):
        return hash(type(self).__name__)


class _LinearTwoQubit(Gateset):
    """Shared shape of the *Linear gatesets (reference: gatesets.py:125-151, 176-243)."""
    two_gate_cls = CNOTGate

    def __init__(self, single_gate=None, single_alt=None):
        self.single_gate = single_gate if single_gate is not None else U3Gate()
        self.single_alt = single_alt if single_alt is not None else XZXZGate()
        self.two_gate = self.two_gate_cls()
        self.cnot = self.two_gate

    def initial_layer(self, n):
        return fill_row(self.single_gate, n)

    def search_layers(self, n):
        return linear_topology(self.two_gate, self.single_gate, n, self.d, single_alt=self.single_alt)

    def branching_factor(self, qudits):
        return qudits - 1

    def successors(self, circ, qudits=None):
        if qudits is None:
            qudits = circ.qudits
        skip = find_last_3_cnots_linear(circ)
        return [(circ.appending(layer), w) for layer, w in
                linear_topology(self.two_gate, self.single_gate, qudits, self.d, single_alt=self.single_alt,
                                skip_index=skip)]


class QubitCNOTLinear(_LinearTwoQubit):
    two_gate_cls = CNOTGate


class QubitCZLinear(_LinearTwoQubit):
    two_gate_cls = CZGate


class QubitISwapLinear(_LinearTwoQubit):
    two_gate_cls = ISwapGate


class QubitXXLinear(_LinearTwoQubit):
    two_gate_cls = XXGate


class _PlainLinear(Gateset):
    single_cls = U3Gate

    def __init__(self):
        self.single_gate = self.single_cls()
        self.cnot = CNOTGate()

    def initial_layer(self, n):
        return fill_row(self.single_gate, n)

    def search_layers(self, n):
        return linear_topology(self.cnot, self.single_gate, n, self.d)


class U3CNOTLinear(_PlainLinear):
    """reference: gatesets.py:112-123"""
    single_cls = U3Gate


class ZXZXZCNOTLinear(_PlainLinear):
    """reference: gatesets.py:99-110"""
    single_cls = ZXZXZGate


class QubitCNOTRing(Gateset):
    """Linear pairs plus the wrap-around CNOT(n-1 -> 0) (reference: gatesets.py:153-174)."""

    def __init__(self, single_gate=None, single_alt=None):
        self.single_gate = single_gate if single_gate is not None else U3Gate()
        self.single_alt = single_alt if single_alt is not None else XZXZGate()
        self.cnot = CNOTGate()

    def initial_layer(self, n):
        return fill_row(self.single_gate, n)

    def search_layers(self, n):
        ident = IdentityGate()
        layers = linear_topology(self.cnot, self.single_gate, n, self.d, identity_gate=ident,
                                 single_alt=self.single_alt)
        if n == 2:
            return layers
        wrap = ProductGate(NonadjacentCNOTGate(n, n - 1, 0),
                           KroneckerGate(self.single_gate, *[ident] * (n - 2), self.single_alt))
        return layers + [(wrap, 1)]


class QubitCNOTAdjacencyList(Gateset):
    """CNOTs on an explicit list of (control, target) pairs (reference: gatesets.py:245-281)."""

    def __init__(self, adjacency, single_gate=None, single_alt=None):
        self.adjacency = adjacency
        self.single_gate = single_gate if single_gate is not None else U3Gate()
        self.single_alt = single_alt if single_alt is not None else XZXZGate()
        self.I = IdentityGate()

    def initial_layer(self, n):
        return fill_row(self.single_gate, n)

    def search_layers(self, n):
        if n == 2:
            return [(ProductGate(CNOTGate(), KroneckerGate(self.single_gate, self.single_gate)), 1)]
        layers = []
        for c, t in self.adjacency:
            if c >= n or t >= n:
                continue
            # single_gate sits on the control, single_alt on the target (reference: gatesets.py:279)
            singles = [self.single_gate if i == c else self.single_alt if i == t else self.I for i in range(n)]
            layers.append((ProductGate(NonadjacentCNOTGate(n, c, t), KroneckerGate(*singles)), 1))
        return layers


class QutritCPIPhaseLinear(Gateset):
    """Single-qutrit gates and CPIPhase on adjacent pairs (reference: gatesets.py:283-293).  As in the reference, every call
    of ``search_layers`` makes a new CPIPhaseGate, i.e. draws new phases."""
    d = 3

    def __init__(self):
        self.single_gate = SingleQutritGate()

    def initial_layer(self, n):
        return fill_row(self.single_gate, n)

    def search_layers(self, n):
        return linear_topology(CPIPhaseGate(), self.single_gate, n, self.d)


class QutritCNOTLinear(Gateset):
    """Single-qutrit gates and the qubit CNOT embedded in two qutrits (reference: gatesets.py:295-305)."""
    d = 3

    def __init__(self):
        self.single_gate = SingleQutritGate()

    def initial_layer(self, n):
        return fill_row(self.single_gate, n)

    def search_layers(self, n):
        return linear_topology(UpgradedConstantGate(CNOTGate()), self.single_gate, n, self.d)


DefaultQubit = QubitCNOTLinear
DefaultQutrit = QutritCPIPhaseLinear
Default = DefaultQubit
