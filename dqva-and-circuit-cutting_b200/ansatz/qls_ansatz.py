"""Quantum Local Search ansatz (reference: ansatz/qls_ansatz.py:15-234): DQVA-style partial
mixers with individual angles, a parameter budget (`param_lim`) and the reference's parameter
parsing: params[0] is the first gamma, then blocks of `num_nonzero` alphas each followed by a
gamma while parameters remain, a short last alpha block taking what is left (:178-206)."""
from __future__ import annotations

from ..circuit import AncillaRegister, QuantumCircuit
from ..utils.helper_funcs import hamming_weight
from .qaoa import partial_mixer

__all__ = ["gen_qlsa", "apply_mixer", "apply_phase_separator"]


def apply_mixer(circ, alpha, init_state, G, barriers, decompose_toffoli, mixer_order, verbose=0,
                fused=False):
    """U_M(alpha) (:15-108): angles are handed out in mixer order to the qubits that are 0 in
    `init_state` (little-endian string) until `alpha` runs out; the others get no mixer."""
    if mixer_order is None:
        mixer_order = list(G.nodes)
    start = list(reversed(init_state))
    angle_of, used = {}, 0
    for qubit in mixer_order:
        if start[qubit] != '1' and used < len(alpha):
            angle_of[qubit] = alpha[used]
            used += 1
    for qubit in mixer_order:
        if qubit not in angle_of or not G.has_node(qubit):
            continue
        ctrl_qubits = [circ.qubits[i] for i in G.neighbors(qubit)]
        partial_mixer(circ, ctrl_qubits, circ.ancillas[0], circ.qubits[qubit], 2 * angle_of[qubit], fused)
        if barriers > 1:
            circ.barrier()


def apply_phase_separator(circ, gamma, G):
    """RZ(2 gamma) on every node (:110-115)."""
    for qb in G.nodes:
        circ.rz(2 * gamma, qb)


def _parse(params, num_nonzero):
    """(:178-206) -> (alpha blocks, gammas)."""
    alphas, gammas, i, need_driver = [], [], 0, False
    total = len(params)
    while i < total:
        if i == 0:
            gammas.append(params[0])
            i = 1
            need_driver = False
        elif total - i >= num_nonzero:
            alphas.append(params[i:i + num_nonzero])
            i += num_nonzero
            if i < total and need_driver:
                gammas.append(params[i])
                i += 1
            need_driver = True
        else:
            alphas.append(params[i:])
            i = total
    return alphas, gammas


def gen_qlsa(G, P=1, params=[], init_state=None, barriers=1, decompose_toffoli=1,
             mixer_order=None, verbose=0, param_lim=None, fused=False):
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
    num_params = min(P * (nq + 1), (P + 1) * (num_nonzero + 1)) if param_lim is None else param_lim
    assert len(params) == num_params, "Incorrect number of parameters!"
    if num_nonzero == 0:
        raise ValueError("init_state has no free qubit: the reference's parser does not terminate")
    alpha_list, gamma_list = _parse(list(params), num_nonzero)
    for i, alphas in enumerate(alpha_list):
        apply_mixer(circ, alphas, init_state, G, barriers, decompose_toffoli, mixer_order, verbose, fused)
        if barriers == 1:
            circ.barrier()
        if i < len(gamma_list):
            apply_phase_separator(circ, gamma_list[i], G)
            if barriers == 1:
                circ.barrier()
    return circ
