"""QAOA+ ansatz (ansatz/qaoa_plus.py:8-53 of the reference): H on every qubit, per layer an
edge-wise phase separator RZ(g/2) (x) RZ(g/2) . CX . RZ(-g/2) . CX and an RX(-2 beta) mixer.
`expectation_value` is the reference's penalised objective on a {bitstring: count} mapping."""
from __future__ import annotations

from ..circuit import QuantumCircuit
from ..utils.helper_funcs import hamming_weight

__all__ = ["construct_qaoa_plus", "expectation_value"]


def construct_qaoa_plus(P, G, params, barriers=False, measure=False):
    assert len(params) == 2 * P, "Number of parameters should be 2P"
    nq = len(G.nodes())
    circ = QuantumCircuit(nq, name='q')
    circ.h(range(nq))
    gammas = [p for i, p in enumerate(params) if i % 2 == 0]
    betas = [p for i, p in enumerate(params) if i % 2 == 1]
    for i in range(P):
        for q_i, q_j in G.edges():
            circ.rz(gammas[i] / 2, [q_i, q_j])
            circ.cx(q_i, q_j)
            circ.rz(-1 * gammas[i] / 2, q_j)
            circ.cx(q_i, q_j)
            if barriers:
                circ.barrier()
        for q_i in range(nq):
            circ.rx(-2 * betas[i], q_i)
    return circ


def expectation_value(counts, G, Lambda):
    """sum_x p(x) (|x| - Lambda * #edges with both ends 1)  (:40-53)."""
    total = sum(counts.values())
    energy = 0
    for bitstr, count in counts.items():
        temp = hamming_weight(bitstr)
        rev = list(reversed(bitstr))
        for q_i, q_j in G.edges():
            if rev[q_i] == '1' and rev[q_j] == '1':
                temp += -1 * Lambda
        energy += count * temp / total
    return energy
