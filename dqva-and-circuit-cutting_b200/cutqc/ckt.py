"""Circuit-knitting-toolbox flavoured entry points over the same kernels.

The CKT copy of CutQC bundled with the reference (circuit-knitting-toolbox/.../wire_cutting*.py)
does in float64 numpy what cutqc/*.c does in float32: `_attribute_shots` = vertical collapse,
`build` = sum over summation terms of kron products / 2^num_cuts, `generate_reconstructed_output`
= the bit permutation to full-circuit order.  These wrappers take the same arguments (numpy
arrays or device tensors) and return device tensors.
"""
from __future__ import annotations

from typing import Dict, List, Sequence

import numpy as np
import torch

from .. import _lib
from . import kernels

__all__ = ["_attribute_shots", "build", "generate_reconstructed_output"]


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
