"""Circuit-knitting-toolbox flavoured entry points over the same kernels.

The CKT copy of CutQC bundled with the reference (circuit-knitting-toolbox/.../wire_cutting*.py)
does in float64 numpy what cutqc/*.c does in float32: `_attribute_shots` = vertical collapse,
`build` = sum over summation terms of kron products / 2^num_cuts, `generate_reconstructed_output`
= the bit permutation to full-circuit order.  These wrappers take the same arguments (numpy
arrays or device tensors) and return device tensors.  `evaluate_subcircuits` /
`reconstruct_full_distribution` are the two top-level calls of wire_cutting.py (:88-187) over them:
`cuts` is the dict CKT's cutter returns ("subcircuits", "complete_path_map", "num_cuts").
"""
from __future__ import annotations

import itertools
from typing import Dict, List, Sequence

import numpy as np
import torch

from .. import _lib
from . import kernels

__all__ = ["_attribute_shots", "build", "generate_reconstructed_output", "generate_summation_terms",
           "_generate_metadata", "evaluate_subcircuits", "reconstruct_full_distribution"]


def _device():
    if not torch.cuda.is_available():
        raise _lib.QdcError("the CKT wrappers run on a CUDA device; there is no CPU fallback")
    return torch.device("cuda")


def _stack(vectors: Dict[int, object], dev) -> torch.Tensor:
    rows = [torch.as_tensor(np.asarray(vectors[i]) if not isinstance(vectors[i], torch.Tensor)
                            else vectors[i], dtype=torch.float64, device=dev)
            for i in range(len(vectors))]
    return torch.stack(rows).contiguous()


def _attribute_shots(subcircuit_entries, subcircuit_instance_probs):
    """subcircuit_entry_probs[subcircuit][entry] = sum coefficient * instance probabilities
    (wire_cutting.py:252-310).  Returns {subcircuit: tensor [n_entries, 2^effective]}."""
    dev = _device()
    out = {}
    for s, entries in subcircuit_entries.items():
        terms = [None] * len(entries)
        for _, (entry_idx, kronecker_term) in entries.items():
            for term in kronecker_term:
                if len(term) != 2:
                    raise ValueError(f"Ill-formed Kronecker term: {term}")
            terms[entry_idx] = [(int(c), int(i)) for c, i in kronecker_term]
        if not subcircuit_instance_probs.get(s):
            raise ValueError("Something unexpected happened during shot attribution.")
        measured = _stack(subcircuit_instance_probs[s], dev)
        out[s], _ = kernels.vertical_collapse(measured, terms)
    return out


def build(summation_terms: Sequence[Dict[int, int]], subcircuit_entry_probs, num_cuts: int,
          num_threads: int = 1):
    """(reconstructed_prob, smart_order, overhead) as wire_cutting_post_processing.py:386-443;
    `num_threads` is accepted and ignored (one device pass).  `overhead` counts what the numpy
    version would have done (naive_compute, :340-383)."""
    dev = _device()
    subs = list(subcircuit_entry_probs.keys())
    vecs = {s: (subcircuit_entry_probs[s] if isinstance(subcircuit_entry_probs[s], torch.Tensor)
                else _stack(subcircuit_entry_probs[s], dev)) for s in subs}
    smart_order = sorted(subs, key=lambda s: vecs[s].shape[1])
    index = [np.array([term[s] for term in summation_terms], dtype=np.int64) for s in smart_order]
    prob = kernels.build([vecs[s] for s in smart_order], index, num_cuts)
    overhead = {"additions": 0, "multiplications": 0}
    length = 1
    per_term_mults = 0
    for pos, s in enumerate(smart_order):
        length *= vecs[s].shape[1]
        if pos > 0:
            per_term_mults += length
    overhead["multiplications"] = per_term_mults * len(summation_terms)
    overhead["additions"] = length * max(len(summation_terms) - 1, 0)
    return prob, smart_order, overhead


def generate_reconstructed_output(full_circuit, subcircuits, unordered, smart_order,
                                  complete_path_map) -> torch.Tensor:
    """Bit permutation of the kron-ordered vector to full-circuit order
    (wire_cutting_verification.py:80-136), one scatter on the device."""
    dev = _device()
    unordered = torch.as_tensor(unordered, dtype=torch.float64, device=dev)
    full_qubits = list(full_circuit.qubits)
    out_qubits = {s: [] for s in smart_order}
    for wire, path in complete_path_map.items():
        last = path[-1]
        out_qubits[last["subcircuit_idx"]].append(
            (list(subcircuits[last["subcircuit_idx"]].qubits).index(last["subcircuit_qubit"]),
             full_qubits.index(wire)))
    wires = []
    for s in smart_order:
        wires += [w for _, w in sorted(out_qubits[s], reverse=True)]
    n = len(full_qubits)
    assert unordered.numel() == 1 << n and sorted(wires) == list(range(n))
    u = torch.arange(1 << n, dtype=torch.int64, device=dev)
    dest = torch.zeros_like(u)
    for j, w in enumerate(wires):                         # character j of the zero-filled binary string
        dest |= ((u >> (n - 1 - j)) & 1) << w
    out = torch.zeros_like(unordered)
    out[dest] = unordered
    return out


# ---- the top-level entries (wire_cutting.py:88-187) ------------------------------------------------
_INIT_OF_LABEL = {"I": ((1, "zero"), (1, "one")), "X": ((2, "plus"), (-1, "zero"), (-1, "one")),
                  "Y": ((2, "plusI"), (-1, "zero"), (-1, "one")), "Z": ((1, "zero"), (-1, "one"))}


def generate_summation_terms(subcircuits, complete_path_map, num_cuts: int):
    """CKT's bookkeeping of the 4^num_cuts terms (wire_cutting_post_processing.py:225-337), with its
    numbering: cut c is base-4 digit c of the term index (least significant first, :48-68); entries
    and instances of a subcircuit are numbered in order of first appearance.  Returns
    (summation_terms, subcircuit_entries, subcircuit_instances) in the reference's dict layout.

    Host side: a subcircuit only sees the digits of its own cut ends, so its entry is looked up by
    that digit tuple instead of by re-deriving label strings per term."""
    from .post_process import get_pairs
    pairs = get_pairs(complete_path_map)
    if len(pairs) != num_cuts:
        raise ValueError(f"num_cuts = {num_cuts} but the path map holds {len(pairs)} cut wires")
    n_sub = len(subcircuits)
    ends = []                                    # per subcircuit: [(cut, "meas" | "init", qubit index)]
    for s in range(n_sub):
        qubits = list(subcircuits[s].qubits)
        mine = []
        for c, (O_q, rho_q) in enumerate(pairs):
            if O_q["subcircuit_idx"] == s:
                mine.append((c, "meas", qubits.index(O_q["subcircuit_qubit"])))
            if rho_q["subcircuit_idx"] == s:
                mine.append((c, "init", qubits.index(rho_q["subcircuit_qubit"])))
        ends.append(mine)
    entries = {s: {} for s in range(n_sub)}
    instances = {s: {} for s in range(n_sub)}
    by_digits = [dict() for _ in range(n_sub)]
    terms = []
    for idx in range(4 ** num_cuts):
        digit = [(idx >> (2 * c)) & 3 for c in range(num_cuts)]
        term = {}
        for s in range(n_sub):
            mine = ends[s]
            key = tuple(digit[c] for c, _, _ in mine)
            if key not in by_digits[s]:
                d = subcircuits[s].num_qubits
                init, meas = ["zero"] * d, ["comp"] * d
                for (c, kind, q), a in zip(mine, key):
                    (meas if kind == "meas" else init)[q] = "IXYZ"[a]
                choices = [_INIT_OF_LABEL[x] if x != "zero" else ((1, "zero"),) for x in init]
                kron_term = []
                for combo in itertools.product(*choices):
                    coeff = int(np.prod([c for c, _ in combo])) if combo else 1
                    inst_key = (tuple(lab for _, lab in combo), tuple(meas))
                    kron_term.append((coeff, instances[s].setdefault(inst_key, len(instances[s]))))
                label = ("".join("IXYZ"[a] for (_, kind, _), a in zip(mine, key) if kind == "init"),
                         "".join("IXYZ"[a] for (_, kind, _), a in zip(mine, key) if kind == "meas"))
                entries[s][label] = (len(entries[s]), kron_term)
                by_digits[s][key] = entries[s][label][0]
            term[s] = by_digits[s][key]
        terms.append(term)
    return terms, entries, instances


def _generate_metadata(cuts):
    """wire_cutting.py:190-217."""
    return generate_summation_terms(cuts["subcircuits"], cuts["complete_path_map"], cuts["num_cuts"])


def evaluate_subcircuits(cuts, backend=None, params=None, param_order=None):
    """`subcircuit_instance_probabilities[subcircuit][instance]` (wire_cutting.py:88-146) from the
    statevector engine: every subcircuit is compiled once and its physical initialisations x basis
    changes run as one batch (evaluator.simulate_subcircuits); the signed marginals of
    `measure_prob` (wire_cutting_evaluation.py:249-296) are one `qdc_cutqc_measure` launch per
    subcircuit over all instances.  `backend` replaces the reference's `service` / `backend_names` /
    `options` (Qiskit Runtime is out of scope): a B200Backend, default one on the current device.
    Returns {subcircuit: {instance index: device tensor [2^effective]}}."""
    from ..engine import B200Backend
    from .evaluator import INIT_LABELS, simulate_subcircuits
    dev = _device()
    backend = backend if backend is not None else B200Backend(device=dev)
    subcircuits = cuts["subcircuits"]
    _, _, instances = _generate_metadata(cuts)
    sims = simulate_subcircuits(backend, subcircuits, cuts["complete_path_map"], params, param_order)
    out = {}
    for s, (tab, raw) in enumerate(sims):
        if raw.shape[2] != 1:
            raise _lib.QdcError("evaluate_subcircuits takes one parameter vector at a time")
        raw = raw[:, :, 0]                                           # [n_init, 3^O, 2^d]
        rows, signs, which = [], [], []
        for (init, meas), idx in instances[s].items():
            for q, lab in enumerate(init):
                if lab not in INIT_LABELS or (lab != "zero" and q not in tab.rho):
                    raise Exception("Illegal initialization :", lab)
            i = 0
            for q in tab.rho:
                i = i * 4 + INIT_LABELS.index(init[q])
            m3, sign = 0, 0
            for q in tab.O:
                if meas[q] not in ("I", "X", "Y", "Z"):
                    raise Exception("Illegal measurement basis:", meas[q])
                m3 = m3 * 3 + {"I": 0, "Z": 0, "X": 1, "Y": 2}[meas[q]]
                sign |= (1 << q) if meas[q] != "I" else 0
            rows.append(i * tab.n_meas3 + m3)
            signs.append(sign)
            which.append(idx)
        flat = raw.reshape(tab.n_init * tab.n_meas3, 1 << tab.d)
        picked = flat.index_select(0, torch.as_tensor(rows, device=flat.device)).contiguous()
        measured = kernels.measure(picked, tab.keep_mask, signs)
        out[s] = {idx: measured[j] for j, idx in enumerate(which)}
    return out


def reconstruct_full_distribution(circuit, subcircuit_instance_probabilities, cuts, num_threads: int = 1):
    """The reconstructed probability vector of the uncut circuit, full-circuit qubit order, as a
    device tensor (wire_cutting.py:149-187: _generate_metadata -> _attribute_shots -> build ->
    generate_reconstructed_output).  `num_threads` is accepted and ignored."""
    summation_terms, subcircuit_entries, _ = _generate_metadata(cuts)
    entry_probs = _attribute_shots(subcircuit_entries, subcircuit_instance_probabilities)
    unordered, smart_order, _ = build(summation_terms=summation_terms, subcircuit_entry_probs=entry_probs,
                                      num_cuts=cuts["num_cuts"], num_threads=num_threads)
    return generate_reconstructed_output(circuit, cuts["subcircuits"], unordered, smart_order,
                                         cuts["complete_path_map"])
