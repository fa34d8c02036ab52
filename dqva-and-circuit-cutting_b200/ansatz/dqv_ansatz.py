"""Uncut DQVA ansatz: `gen_dqva(G, P, params, init_state, ..., mixer_order)`.

Host-side circuit construction with the reference's names, argument meaning and gate order
(ansatz/dqv_ansatz.py:15-195), emitting this package's qiskit-free `QuantumCircuit`; one ancilla.
`fused=True` (extension) emits each partial mixer as one open-controlled RX (`zcrx`), the same
unitary with the ancilla left in |0>; the engine's compiler performs the same fusion on the
gate-faithful form, so both run identically on the device."""
from __future__ import annotations

from ..circuit import AncillaRegister, QuantumCircuit
from ..utils.helper_funcs import hamming_weight

__all__ = ["gen_dqva", "apply_mixer", "apply_phase_separator"]


def apply_mixer(circ, alpha, init_state, G, barriers, decompose_toffoli, mixer_order,
                verbose=0, fused=False):
    """U_M(alpha): one partial mixer per qubit that is 0 in `init_state` (:15-108)."""
    if mixer_order is None:
        mixer_order = list(G.nodes)
    pad_alpha = [None] * len(init_state)
    next_alpha = 0
    rev = list(reversed(init_state))
    for qubit in mixer_order:
        if rev[qubit] == '1' or next_alpha >= len(alpha):
            continue
        pad_alpha[qubit] = alpha[next_alpha]
        next_alpha += 1
    for qubit in mixer_order:
        if pad_alpha[qubit] is None or not G.has_node(qubit):
            continue
        ctrl_qubits = [circ.qubits[i] for i in G.neighbors(qubit)]
        anc = circ.ancillas[0]
        if fused:
            circ.zcrx(2 * pad_alpha[qubit], ctrl_qubits, circ.qubits[qubit])
            continue
        def _toffoli():                                   # open controls = X . MCX . X (:77-84)
            for ctrl in ctrl_qubits:
                circ.x(ctrl)
            circ.mcx(ctrl_qubits, anc)
            for ctrl in ctrl_qubits:
                circ.x(ctrl)
        _toffoli()
        circ.crx(2 * pad_alpha[qubit], anc, circ.qubits[qubit])
        _toffoli()
        if barriers > 1:
            circ.barrier()


def apply_phase_separator(circ, gamma, G):
    """RZ(2 gamma) on every node (:111-116)."""
    for qb in G.nodes:
        circ.rz(2 * gamma, qb)


def gen_dqva(G, P=1, params=[], init_state=None, barriers=1, decompose_toffoli=1,
             mixer_order=None, verbose=0, fused=False):
    """(:122-195) params = P x [alphas of the non-zero-init qubits..., gamma], leftover
    parameters become one more mixer layer."""
    nq = len(G.nodes)
    if init_state is None:
        init_state = '0' * nq
    circ = QuantumCircuit(nq, name='q')
    circ.add_register(AncillaRegister(1, 'anc'))
    for qb, bit in enumerate(reversed(init_state)):
        if bit == '1':
            circ.x(qb)
    if barriers > 0:
        circ.barrier()
    num_nonzero = nq - hamming_weight(init_state)
    assert len(params) == (nq + 1) * P, "Incorrect number of parameters!"
    alpha_list, gamma_list, last_idx = [], [], 0
    for p in range(P):
        chunk = num_nonzero + 1
        cur = params[p * chunk:(p + 1) * chunk]
        alpha_list.append(cur[:-1])
        gamma_list.append(cur[-1])
        last_idx = (p + 1) * chunk
    alpha_list.append(params[last_idx:])
    for i, alphas in enumerate(alpha_list):
        apply_mixer(circ, alphas, init_state, G, barriers, decompose_toffoli, mixer_order,
                    verbose=verbose, fused=fused)
        if barriers > 0:
            circ.barrier()
        if i < len(gamma_list):
            apply_phase_separator(circ, gamma_list[i], G)
            if barriers > 0:
                circ.barrier()
    return circ
