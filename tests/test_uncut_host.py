"""Host logic of the uncut ansatz modules (ansatz/dqv_ansatz.py, ansatz/qaoa_plus.py of the
reference) checked with the numpy oracle: fused == gate-faithful, parameter parsing, QAOA+
objective."""
import itertools

import networkx as nx
import numpy as np

from oracle import statevector as sv
from qdc_b200.ansatz import dqv_ansatz, qaoa_plus


def test_dqva_fused_equals_faithful_and_parameter_layout():
    G = nx.random_regular_graph(3, 8, seed=3)
    rng = np.random.default_rng(0)
    init = "00100001"
    params = list(rng.uniform(0, 2 * np.pi, 9 * 2))
    order = [int(v) for v in rng.permutation(8)]
    a = dqv_ansatz.gen_dqva(G, P=2, params=params, init_state=init, mixer_order=order)
    b = dqv_ansatz.gen_dqva(G, P=2, params=params, init_state=init, mixer_order=order, fused=True)
    assert a.num_qubits == b.num_qubits == 9 and len(a.ancillas) == 1
    pa = sv.probabilities(sv.simulate(9, a.to_instructions()))
    pb = sv.probabilities(sv.simulate(9, b.to_instructions()))
    assert np.abs(pa - pb).max() < 1e-12
    assert pa[256:].sum() < 1e-20                      # ancilla returned to |0>
    # 6 zero-init qubits: P=2 chunks of 6 alphas + gamma, 4 leftover alphas -> 6 + 6 + 4 mixers
    assert sum(1 for g in b.data if g.name == "zcrx") == 16
    assert sum(1 for g in b.data if g.name == "rz") == 16
    assert sum(1 for g in a.data if g.name == "crx") == 16


def test_qaoa_plus_expectation_value():
    G = nx.cycle_graph(4)
    circ = qaoa_plus.construct_qaoa_plus(2, G, [0.3, 0.7, 1.1, 0.2])
    assert sum(1 for g in circ.data if g.name == "cx") == 2 * 2 * 4
    p = sv.probabilities(sv.simulate(4, circ.to_instructions()))
    counts = {format(x, "04b"): p[x] for x in range(16)}
    want = sum(p[x] * (bin(x).count("1") - 1.5 * sum(1 for i, j in G.edges() if (x >> i) & (x >> j) & 1))
               for x in range(16))
    assert abs(qaoa_plus.expectation_value(counts, G, 1.5) - want) < 1e-12
