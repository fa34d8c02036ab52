"""QAO-Ansatz for MIS (reference: ansatz/qaoa.py:21-134): every node gets the same mixer angle
beta per layer, `params = [beta_0, gamma_0, beta_1, gamma_1, ...]`, one ancilla.  Same gate
order as the reference; `fused=True` (extension) emits each partial mixer as one
open-controlled RX instead of X^N . MCX . X^N . CRX . X^N . MCX . X^N."""
from __future__ import annotations

import numpy as np

from ..circuit import AncillaRegister, QuantumCircuit
from ..utils.helper_funcs import hamming_weight

__all__ = ["gen_qaoa", "apply_mixer", "apply_phase_separator", "partial_mixer"]


def partial_mixer(circ, ctrl_qubits, anc, target, theta, fused=False):
    """RX(theta) on `target` iff all `ctrl_qubits` are |0> (ansatz/qaoa.py:42-66, the same
    block in dqv_ansatz.py:77-104 and qls_ansatz.py:77-104)."""
    if fused:
        circ.zcrx(theta, ctrl_qubits, target)
        return
    for _ in range(2):
        for ctrl in ctrl_qubits:
            circ.x(ctrl)
        circ.mcx(ctrl_qubits, anc)
        for ctrl in ctrl_qubits:
            circ.x(ctrl)
        if _ == 0:
            circ.crx(theta, anc, target)


def apply_mixer(circ, G, beta, anc_idx, barriers, decompose_toffoli, mixer_order, verbose=0,
                fused=False):
    """U_M(beta): one partial mixer per node, all with angle 2 beta (:21-69)."""
    if decompose_toffoli == 0 and not fused:
        raise NotImplementedError("decompose_toffoli=0 (custom open-controlled gate) is what "
                                  "fused=True emits; the reference cannot simulate it either")
    if mixer_order is None:
        mixer_order = list(G.nodes)
    for qubit in mixer_order:
        ctrl_qubits = [circ.qubits[i] for i in G.neighbors(qubit)]
        partial_mixer(circ, ctrl_qubits, circ.ancillas[anc_idx], circ.qubits[qubit], 2 * beta, fused)
        if barriers > 1:
            circ.barrier()


def apply_phase_separator(circ, gamma, G):
    """RZ(2 gamma) on every node (:71-73)."""
    for qb in G.nodes:
        circ.rz(2 * gamma, qb)


def gen_qaoa(G, P, params=[], init_state=None, barriers=1, decompose_toffoli=1, mixer_order=None,
             verbose=0, fused=False):
    """(:75-134).  init_state 'W' (a dense `initialize`, :93-100) is not supported by the gate
    set of this package; decompose_toffoli > 1 (unrolling to basis gates) changes nothing the
    simulator sees and is ignored."""
    nq = len(G.nodes)
    if init_state is None:
        init_state = '0' * nq
    if init_state == 'W':
        raise NotImplementedError("init_state='W' needs a state initialiser; not on the hot path")
    circ = QuantumCircuit(nq, name='q')
    circ.add_register(AncillaRegister(1, 'anc'))
    for qb, bit in enumerate(reversed(init_state)):
        if bit == '1':
            circ.x(qb)
    if barriers > 0:
        circ.barrier()
    assert len(params) == 2 * P, "Incorrect number of parameters!"
    for beta, gamma in zip(params[0::2], params[1::2]):
        apply_mixer(circ, G, beta, 0, barriers, decompose_toffoli, mixer_order, verbose, fused)
        if barriers > 0:
            circ.barrier()
        apply_phase_separator(circ, gamma, G)
        if barriers > 0:
            circ.barrier()
    return circ
