"""Shared cut-circuit cases of the cutqc tests (host helpers, no device work)."""
import networkx as nx
import numpy as np

from qdc_b200.circuit import QuantumCircuit
from qdc_b200.qsplit import qsplit_circuit_cutter as qcc
from test_oracle_golden import _cut_dqva, _ghz3


def _rot(circ, rng, qubits):
    for q in qubits:
        circ.u3(*rng.uniform(0, 2 * np.pi, 3), q)


def there_and_back():
    """Wire 2 leaves the first fragment, crosses a second one and comes back: the first
    fragment has an O end and a rho end of the same wire (path length 3, K = 2)."""
    rng = np.random.default_rng(7)
    circ = QuantumCircuit(5, name="q")
    _rot(circ, rng, range(5))
    circ.cx(0, 1); circ.cx(1, 2)                       # block A on wires 0, 1, 2
    _rot(circ, rng, (1, 2))
    circ.cx(2, 3); circ.cx(3, 4); circ.crx(0.7, 4, 2)  # block B on wires 2, 3, 4
    _rot(circ, rng, (2, 3))
    circ.cx(2, 1); circ.cx(1, 0)                       # block C on wires 0, 1, 2
    _rot(circ, rng, (0, 2))
    ops2 = [op for op in circ.data if circ.qubits[2] in op.qargs]
    first_b = next(i for i, op in enumerate(ops2) if op.name == "cx" and op.qargs[1] is circ.qubits[3])
    first_c = next(i for i, op in enumerate(ops2) if op.name == "cx" and op.qargs[1] is circ.qubits[1])
    cuts = [(circ.qubits[2], first_b), (circ.qubits[2], first_c)]
    frags, wpm = qcc.cut_circuit(circ, cuts)
    return circ, frags, wpm


def chain3():
    """Three fragments in a chain, one cut between neighbours (K = 2, F = 3)."""
    rng = np.random.default_rng(11)
    circ = QuantumCircuit(6, name="q")
    _rot(circ, rng, range(6))
    circ.cx(0, 1); circ.cx(1, 2)
    _rot(circ, rng, (1, 2))
    circ.cx(2, 3); circ.cx(3, 4)
    _rot(circ, rng, (3, 4))
    circ.cx(4, 5); circ.rz(0.3, 5); circ.h(5)
    ops2 = [op for op in circ.data if circ.qubits[2] in op.qargs]
    ops4 = [op for op in circ.data if circ.qubits[4] in op.qargs]
    c2 = next(i for i, op in enumerate(ops2) if op.name == "cx" and op.qargs[1] is circ.qubits[3])
    c4 = next(i for i, op in enumerate(ops4) if op.name == "cx" and op.qargs[1] is circ.qubits[5])
    frags, wpm = qcc.cut_circuit(circ, [(circ.qubits[2], c2), (circ.qubits[4], c4)])
    return circ, frags, wpm


def ckt5():
    """The circuit of circuit-knitting-toolbox's own wire-cutting test
    (circuit-knitting-toolbox/test/circuit_cutting/test_circuit_cutting.py:28-47: H, CX, T, RX, RY on 5
    qubits) with its MANUAL cut `subcircuit_vertices=[[0, 1], [2, 3]]` (:69-72): the two-qubit gates in
    circuit order are cx(0,1), cx(0,2) | cx(2,4), cx(2,3), so wire 2 is cut between cx(0,2) and
    cx(2,4).  Known answer held by the reference (:77-82): the reconstruction equals the full
    statevector distribution (mean squared error 0)."""
    circ = QuantumCircuit(5, name="q")
    for i in range(5):
        circ.h(i)
    circ.cx(0, 1)
    for i in range(2, 5):
        circ.t(i)
    circ.cx(0, 2)
    circ.rx(np.pi / 2, 4)
    circ.rx(np.pi / 2, 0)
    circ.rx(np.pi / 2, 1)
    circ.cx(2, 4)
    circ.t(0)
    circ.t(1)
    circ.cx(2, 3)
    circ.ry(np.pi / 2, 4)
    for i in range(5):
        circ.h(i)
    ops2 = [op for op in circ.data if circ.qubits[2] in op.qargs]
    cut = next(i for i, op in enumerate(ops2) if op.name == "cx" and op.qargs[1] is circ.qubits[4])
    frags, wpm = qcc.cut_circuit(circ, [(circ.qubits[2], cut)])
    return circ, frags, wpm


def ghz3():
    circ, cuts = _ghz3()
    frags, wpm = qcc.cut_circuit(circ, cuts)
    return circ, frags, wpm


def dqva6(max_cuts=2, seed=1, fused=False):
    G = nx.Graph()
    G.add_edges_from([(0, 1), (1, 2), (2, 0), (2, 3), (3, 4), (4, 5), (5, 3), (1, 4)])
    circ, _, frags, wpm = _cut_dqva(G, [[0, 1, 2], [3, 4, 5]], max_cuts, seed, fused)
    return circ, frags, wpm, G


CASES = {"ghz3": ghz3, "ckt5": ckt5, "there_and_back": there_and_back, "chain3": chain3,
         "dqva6": lambda: dqva6()[:3], "dqva6_fused": lambda: dqva6(2, 5, True)[:3]}
