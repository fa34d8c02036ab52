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


def _indset(x, G):
    return all(not ((x >> i) & (x >> j) & 1) for i, j in G.edges())


def test_qao_and_qls_ansatz_stay_in_the_independent_set_subspace():
    """Domain invariant of the constrained mixers (ansatz/qaoa.py:21-69, qls_ansatz.py:15-108):
    from an independent-set initial state every amplitude outside the independent sets, and
    every amplitude with the ancilla set, is zero; fused == gate-faithful."""
    from qdc_b200.ansatz import qaoa, qls_ansatz
    G = nx.random_regular_graph(3, 8, seed=5)
    rng = np.random.default_rng(1)
    init = ["0"] * 8
    init[7 - 2] = "1"                                  # node 2 alone is an independent set
    init = "".join(init)
    order = [int(v) for v in rng.permutation(8)]
    cases = []
    prm = list(rng.uniform(0, 2 * np.pi, 4))
    cases.append([qaoa.gen_qaoa(G, 2, prm, init_state=init, mixer_order=order, fused=f) for f in (False, True)])
    nz = 7
    prm = list(rng.uniform(0, 2 * np.pi, min(2 * 9, 3 * (nz + 1))))
    cases.append([qls_ansatz.gen_qlsa(G, 2, prm, init_state=init, mixer_order=order, fused=f) for f in (False, True)])
    prm = list(rng.uniform(0, 2 * np.pi, 5))
    cases.append([qls_ansatz.gen_qlsa(G, 1, prm, init_state=init, mixer_order=order, param_lim=5, fused=f)
                  for f in (False, True)])
    for faithful, fused in cases:
        pa = sv.probabilities(sv.simulate(9, faithful.to_instructions()))
        pb = sv.probabilities(sv.simulate(9, fused.to_instructions()))
        assert np.abs(pa - pb).max() < 1e-12 and abs(pa.sum() - 1) < 1e-12
        assert pa[256:].sum() < 1e-20
        assert sum(pa[x] for x in range(256) if not _indset(x, G)) < 1e-20
        assert pa[1 << 2] < 1 - 1e-3                   # the mixers did move amplitude
    # qls parameter parsing (:178-206): gamma first, then alpha blocks each followed by a gamma
    alphas, gammas = qls_ansatz._parse(list(range(12)), 4)
    assert gammas == [0, 9] and alphas == [[1, 2, 3, 4], [5, 6, 7, 8], [10, 11]]
    # every mixer of the QAO-Ansatz shares beta: 8 nodes x 2 layers
    assert sum(1 for g in cases[0][1].data if g.name == "zcrx") == 16
