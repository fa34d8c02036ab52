"""Cut-friendly DQVA ansatz: `gen_dqva(...) -> (circuit, cuts)`.

Host-side circuit construction with the reference's names, argument meaning and gate order
(ansatz/dqv_cut_ansatz.py:21-356), emitting this package's qiskit-free `QuantumCircuit`.
It defines the gate set the CUDA statevector engine runs:

  X on every '1' of `init_state` (:283-285); per active node q a partial mixer
  X^N(q) . MCX(N(q) -> anc) . X^N(q) . CRX(2 alpha_q; anc -> q) . X^N(q) . MCX . X^N(q)
  (:127-151); multi-qubit identity `Id` blocks tying each subgraph together (:234-249);
  RZ(2 gamma) on every problem qubit (:166-171).  Only the first layer is built (:337-343).

`fused=True` is an extension: each partial mixer is emitted as one open-controlled RX
(`zcrx`), which is the same unitary with the ancilla left in |0>.  Cut positions are
computed on the gate-faithful operation count either way, so both modes cut at the same
place in time.
"""
from __future__ import annotations

from typing import Dict, List

import networkx as nx

from ..circuit import AncillaRegister, QuantumCircuit, Qubit
from ..utils.graph_funcs import get_subgraphs

__all__ = ["gen_dqva", "apply_mixer", "apply_phase_separator", "find_cuts", "num_gates",
           "tie_subgraph_qubits_together"]


def num_gates(circuit: QuantumCircuit, qubit: Qubit) -> int:
    """Operations (barriers excluded) acting on `qubit` so far (:174-180)."""
    return sum(1 for inst in circuit.data if inst.name != "barrier" and qubit in inst.qargs)


def tie_subgraph_qubits_together(circ: QuantumCircuit, subgraph: nx.Graph, anc_idx: int) -> None:
    """Append one `Id` instruction over the subgraph's qubits and its ancilla (:234-249)."""
    targets = [circ.qubits[node] for node in subgraph.nodes] + [circ.ancillas[anc_idx]]
    circ.identity_block(targets, name="Id")


def find_cuts(G, circ, subgraphs, subgraph_dict: Dict[int, int], mixer_order: List[int],
              qubit_index: int, qubit: int, hot_nodes, cut_nodes, active_qubits) -> List[Qubit]:
    """Qubits to sever when the mixer sequence moves from one subgraph to the next
    (:183-231): cut nodes the previous subgraph has touched that a later subgraph still
    needs (as a mixer target, a mixer control, or simply because they live there)."""
    prev_subgraph = subgraph_dict[mixer_order[qubit_index - 1]]
    tie_subgraph_qubits_together(circ, subgraphs[prev_subgraph], prev_subgraph)

    touched = set()
    for prev_node in subgraphs[prev_subgraph].nodes:
        if prev_node in cut_nodes:
            touched.add(prev_node)
        if active_qubits[prev_node]:
            for node in list(G.neighbors(prev_node)) + [prev_node]:
                if node in cut_nodes:
                    touched.add(node)

    later_subgraphs = {subgraph_dict[mixer_order[i]] for i in range(qubit_index, len(mixer_order))}
    to_cut = set()
    for node in touched:
        for later in later_subgraphs:
            for later_node in subgraphs[later].nodes:
                if active_qubits[later_node] and \
                        (node == later_node or node in G.neighbors(later_node)):
                    to_cut.add(node)
                if node == later_node:
                    to_cut.add(node)
    # the reference returns list(set(Qubit...)) (hash order); ascending node order here
    return [circ.qubits[node] for node in sorted(to_cut)]


def apply_mixer(circ, alpha, init_state, G, barriers, decompose_toffoli, mixer_order,
                subgraph_dict, cut_nodes, hot_nodes, subgraphs, verbose=0, fused=False):
    """Partial mixers in `mixer_order`; returns the cut list [(Qubit, op_number)] (:21-163)."""
    if mixer_order is None:
        mixer_order = list(G.nodes)

    # alpha assignment: a cut node that is not hot gets no mixer; every other node takes the
    # next alpha in sequence, including nodes already in |1> (:66-83)
    pad_alpha = [None] * len(init_state)
    next_alpha = 0
    for qubit in mixer_order:
        if qubit in cut_nodes and qubit not in hot_nodes:
            continue
        pad_alpha[qubit] = alpha[next_alpha]
        next_alpha += 1
    active_qubits = [v is not None for v in pad_alpha]

    # first mixer of every subgraph after the first one (:89-94)
    swap_qubits = []
    cur = mixer_order[0]
    for qubit in mixer_order[1:]:
        if subgraph_dict[qubit] != subgraph_dict[cur]:
            cur = qubit
            swap_qubits.append(qubit)

    # operation counts for cut positions are always taken on the gate-faithful sequence so
    # fused and unfused circuits are cut at the same logical place
    shadow = circ if not fused else circ.copy()

    cuts = []
    qubit = mixer_order[-1]
    for qubit_index, qubit in enumerate(mixer_order):
        if qubit in swap_qubits and len(hot_nodes) > 0:
            if fused:
                shadow_cut = find_cuts(G, shadow, subgraphs, subgraph_dict, mixer_order,
                                       qubit_index, qubit, hot_nodes, cut_nodes, active_qubits)
                prev_subgraph = subgraph_dict[mixer_order[qubit_index - 1]]
                tie_subgraph_qubits_together(circ, subgraphs[prev_subgraph], prev_subgraph)
                for qb in shadow_cut:
                    cuts.append((qb, num_gates(circ, qb)))
            else:
                for qb in find_cuts(G, circ, subgraphs, subgraph_dict, mixer_order,
                                    qubit_index, qubit, hot_nodes, cut_nodes, active_qubits):
                    cuts.append((qb, num_gates(circ, qb)))

        if pad_alpha[qubit] is None or not G.has_node(qubit):
            continue

        neighbors = list(G.neighbors(qubit))
        anc = circ.ancillas[subgraph_dict[qubit]]
        ctrl_qubits = [circ.qubits[i] for i in neighbors]

        def _faithful(c):
            for ctrl in ctrl_qubits:
                c.x(ctrl)
            c.mcx(ctrl_qubits, anc)
            for ctrl in ctrl_qubits:
                c.x(ctrl)
            c.crx(2 * pad_alpha[qubit], anc, c.qubits[qubit])
            for ctrl in ctrl_qubits:
                c.x(ctrl)
            c.mcx(ctrl_qubits, anc)
            for ctrl in ctrl_qubits:
                c.x(ctrl)

        if fused:
            circ.zcrx(2 * pad_alpha[qubit], ctrl_qubits, circ.qubits[qubit])
            _faithful(shadow)
        else:
            _faithful(circ)
        if barriers > 1:
            circ.barrier()

    if len(hot_nodes) > 0:                      # for-else of the reference (:157-161)
        subgraph_idx = subgraph_dict[qubit]
        tie_subgraph_qubits_together(circ, subgraphs[subgraph_idx], subgraph_idx)
    return cuts


def apply_phase_separator(circ, gamma, G):
    for qb in G.nodes:
        circ.rz(2 * gamma, qb)


def gen_dqva(G, partition, cut_nodes, hot_nodes, mixer_order, P=1, params=[], init_state=None,
             barriers=1, decompose_toffoli=1, verbose=0, fused=False):
    nq = len(G.nodes)
    if P != 1:
        raise Exception("P != 1 currently unsupported")

    subgraphs, _ = get_subgraphs(G, partition)
    subgraph_dict = {q: i for i, sub in enumerate(subgraphs) for q in sub}

    if init_state is None:
        init_state = '0' * nq

    dqva_circ = QuantumCircuit(nq, name='q')
    dqva_circ.add_register(AncillaRegister(len(partition), 'anc'))

    for qb, bit in enumerate(reversed(init_state)):
        if bit == '1':
            dqva_circ.x(qb)
    if barriers > 0:
        dqva_circ.barrier()

    # parameter layout: P chunks of (nq alphas, 1 gamma), leftovers = extra mixer layer
    chunk = nq + 1
    alpha_list, gamma_list = [], []
    for p in range(P):
        section = params[p * chunk:(p + 1) * chunk]
        alpha_list.append(section[:-1])
        gamma_list.append(section[-1])
    if len(params[P * chunk:]) > 0:
        alpha_list.append(params[P * chunk:])

    cuts = []
    for i, alphas in enumerate(alpha_list):
        cuts = apply_mixer(dqva_circ, alphas, init_state, G, barriers, decompose_toffoli,
                           mixer_order, subgraph_dict, cut_nodes, hot_nodes, subgraphs,
                           verbose=verbose, fused=fused)
        if barriers > 0:
            dqva_circ.barrier()
        if i < len(gamma_list):
            apply_phase_separator(dqva_circ, gamma_list[i], G)
            if barriers > 0:
                dqva_circ.barrier()
        break       # only the first layer is cut-compatible (:337-343)

    return dqva_circ, cuts
