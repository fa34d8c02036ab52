"""Kronecker / summation term tables of the CutQC reconstruction (reference:
cutqc/post_process.py), as integer arrays instead of label tuples.

With K cut pairs (O end, rho end) every combination kappa in {I,X,Y,Z}^K (itertools.product
order, first pair most significant: combination index t = sum_k a_k 4^(K-1-k)) contributes
    0.5^K * kron_s  v_s[kappa restricted to the pairs that touch subcircuit s]
and v_s[.] = sum_j coefficient_j * measured instance_j, the expansion of each rho label into
initialisations (post_process.py:20-30):
    I -> +zero +one    X -> 2 plus -zero -one    Y -> 2 plusI -zero -one    Z -> +zero -one
The reference dedupes identical kron terms per subcircuit in order of first appearance
(post_process.py:119-124); a term is determined by kappa restricted to the subcircuit's own
pairs, and first appearance in product order is lexicographic order of that restriction, so
    kron index = base-4 number of the restricted labels (first pair most significant).
`tests/test_cutqc_host.py` checks this numbering against the label-by-label oracle.
"""
from __future__ import annotations

import itertools
from typing import Dict, List, Sequence

import numpy as np

from .evaluator import InstanceTable, PAULI_LABELS, _table_for, find_subcircuit_O_rho_qubits

__all__ = ["get_combinations", "find_init_meas", "build", "TermTables", "term_tables"]

# rho label -> [(coefficient, initialisation digit)] in the reference's listing order
_RHO_EXPANSION = ([(1, 0), (1, 1)], [(2, 2), (-1, 0), (-1, 1)], [(2, 3), (-1, 0), (-1, 1)],
                  [(1, 0), (-1, 1)])


def get_pairs(complete_path_map):
    """(O end, rho end) per cut in path order (post_process.py:52-60, cutter.py:439-448)."""
    return [(path[j], path[j + 1]) for path in complete_path_map.values()
            for j in range(len(path) - 1)]


def get_combinations(complete_path_map):
    pairs = get_pairs(complete_path_map)
    return pairs, list(itertools.product(PAULI_LABELS, repeat=len(pairs)))


class TermTables:
    """Everything the device stages need, per subcircuit s:
    tables[s]     InstanceTable (instance numbering)
    pairs_of[s]   indices of the cut pairs touching s, ascending
    kron_terms[s] list over kron index of [(coefficient, instance index), ...]
    term_index[s] int64[4^K]: kron index used by summation term t
    """

    def __init__(self, subcircuits, complete_path_map):
        self.pairs = get_pairs(complete_path_map)
        K = self.num_cuts = len(self.pairs)
        self.tables: List[InstanceTable] = []
        self.pairs_of: List[List[int]] = []
        self.kron_terms: List[List[list]] = []
        self.term_index: List[np.ndarray] = []
        t = np.arange(4 ** K, dtype=np.int64)
        digit = [(t >> (2 * (K - 1 - k))) & 3 for k in range(K)]        # label of pair k in term t
        for s, sub in enumerate(subcircuits):
            O_q, rho_q = find_subcircuit_O_rho_qubits(complete_path_map, s)
            tab = _table_for(sub, O_q, rho_q)
            qubits = list(sub.qubits)
            mine = [k for k, (o, r) in enumerate(self.pairs)
                    if o["subcircuit_idx"] == s or r["subcircuit_idx"] == s]
            # position of every pair end in the instance table's rho / O lists
            rho_pair = [k for k in mine if self.pairs[k][1]["subcircuit_idx"] == s]
            O_pair = [k for k in mine if self.pairs[k][0]["subcircuit_idx"] == s]
            assert [qubits.index(self.pairs[k][1]["subcircuit_qubit"]) for k in rho_pair] == tab.rho
            assert [qubits.index(self.pairs[k][0]["subcircuit_qubit"]) for k in O_pair] == tab.O
            # the reference expands initialisations in subcircuit-qubit order (post_process.py:18-33)
            rho_by_qubit = sorted(range(len(rho_pair)), key=lambda j: tab.rho[j])
            terms = []
            for labels in itertools.product(range(4), repeat=len(mine)):
                lab = dict(zip(mine, labels))
                meas_digits = [lab[k] for k in O_pair]
                term = []
                for choice in itertools.product(*[_RHO_EXPANSION[lab[rho_pair[j]]] for j in rho_by_qubit]):
                    coeff, init_digits = 1, [0] * len(rho_pair)
                    for j, (c, a) in zip(rho_by_qubit, choice):
                        coeff *= c
                        init_digits[j] = a
                    term.append((coeff, tab.index(init_digits, meas_digits)))
                terms.append(term)
            idx = np.zeros(4 ** K, dtype=np.int64)
            for k in mine:
                idx = idx * 4 + digit[k]
            self.tables.append(tab)
            self.pairs_of.append(mine)
            self.kron_terms.append(terms)
            self.term_index.append(idx)


def term_tables(subcircuits, complete_path_map) -> TermTables:
    return TermTables(subcircuits, complete_path_map)


def find_init_meas(combination, O_rho_pairs, subcircuits):
    """Reference-layout view of one combination (post_process.py:3-50):
    {subcircuit: [(init labels with '+zero' / '2plus' / ... , meas labels with '+X' ...), ...]}."""
    names = ("zero", "one", "plus", "plusI")
    out = {}
    for s, sub in enumerate(subcircuits):
        qubits = list(sub.qubits)
        init = [[("", "zero")] for _ in qubits]            # (sign prefix, label) choices per qubit
        meas = ["comp"] * len(qubits)
        for lab, (o, r) in zip(combination, O_rho_pairs):
            a = PAULI_LABELS.index(lab)
            if r["subcircuit_idx"] == s:
                init[qubits.index(r["subcircuit_qubit"])] = [
                    ({1: "+", -1: "-", 2: "2"}[c], names[d]) for c, d in _RHO_EXPANSION[a]]
            if o["subcircuit_idx"] == s:
                meas[qubits.index(o["subcircuit_qubit"])] = "+" + lab
        out[s] = [(tuple(p + l if p else l for p, l in choice), tuple(meas))
                  for choice in itertools.product(*init)]
    return out


def build(full_circuit, combinations, O_rho_pairs, subcircuits, all_indexed_combinations):
    """Reference-layout result (post_process.py:65-129): (kronecker_terms, summation_terms) with
    kronecker_terms[s] = {((coefficient, instance), ...): kron index} and summation_terms[t] =
    {s: kron index}.  Computed from the array tables; `all_indexed_combinations` is only used
    to look up instance numbers, as in the reference."""
    coeff = {"+": 1, "-": -1, "2": 2}
    kron: Dict[int, dict] = {s: {} for s in range(len(subcircuits))}
    summation = []
    for combination in combinations:
        per_sub = find_init_meas(combination, O_rho_pairs, subcircuits)
        row = {}
        for s in range(len(subcircuits)):
            term = []
            for init, meas in per_sub[s]:
                c, labels = 1, []
                for x in init:
                    if x == "zero":
                        labels.append(x)
                    else:
                        c *= coeff[x[0]]
                        labels.append(x[1:])
                meas = tuple(x if x == "comp" else x[1:] for x in meas)
                term.append((c, all_indexed_combinations[s][(tuple(labels), meas)]))
            term = tuple(term)
            row[s] = kron[s].setdefault(term, len(kron[s]))
        summation.append(row)
    return kron, summation
