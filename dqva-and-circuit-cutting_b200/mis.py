"""Cost functions of the reference's variational solvers (mis.py), evaluated on the B200.

Only the per-evaluation hot path lives here; the outer optimisation loops of
`solve_mis_cut_dqva` / `solve_mis_dqva` (scipy COBYLA, partition rounds) are unchanged host
Python in the reference and call these where they call their closures.

* `cut_cost`   -- `avg_cost` of solve_mis_cut_dqva (mis.py:117-134): sim_with_cutting, strip the
                  ancillas, sum p(x) |x| [x independent set]; batched over parameter vectors.
* `uncut_cost` -- `f(params)` of solve_mis_dqva / solve_mis_qaoa (mis.py:639-665): full
                  statevector, strip the ancilla, sum p(x) |x| (no independent-set mask there).
                  The score is taken straight from the amplitudes (no probability vector is
                  written).  The reference rounds probabilities to 5 decimals in its statevector
                  branch (:652); this evaluates the exact sum.
* `top_strings`-- top_probs / better_strs (mis.py:232-242).
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import numpy as np
import torch

from . import recombine as _rc
from .engine import B200Backend
from .utils.cutting_funcs import sim_with_cutting

__all__ = ["cut_cost", "uncut_cost", "qaoa_plus_cost", "top_strings"]


def _adjacency(graph, n_wires):
    n = graph.number_of_nodes()
    return [sum(1 << int(v) for v in graph.neighbors(i)) for i in range(n)] + [0] * (n_wires - n)


def cut_cost(fragments, wire_path_map, frag_shots, backend: B200Backend, graph=None,
             params=None, param_order=None, mode="likely", shard: bool = False) -> np.ndarray:
    """-avg_cost for every parameter vector (a length-1 array for bound fragments).
    shard=True under torchrun: the parameter sets are dealt over the ranks and the costs
    all-gathered (`dist.sharded_over_parameter_sets`); every rank returns the full vector."""
    if graph is not None:
        backend.graph = graph

    def local(rows):
        dists = sim_with_cutting(fragments, wire_path_map, frag_shots, backend, mode=mode,
                                 params=rows, param_order=param_order, store=False)
        if not isinstance(dists, list):
            dists = [dists]
        return np.array([-d.score.get()["sum_pw_indset"] for d in dists])

    if shard and params is not None:
        from . import dist as _dist
        return _dist.sharded_over_parameter_sets(local, params)
    return local(params)


def uncut_cost(circuit, backend: B200Backend, graph, params=None, param_order=None,
               masked: bool = False, max_states: Optional[int] = None,
               shard: bool = False) -> np.ndarray:
    """-sum_x p(x) |x| of the uncut circuit for every parameter vector (masked=True adds the
    independent-set indicator).  Wires beyond graph.number_of_nodes() are ancillas.
    shard=True under torchrun: parameter sets dealt over the ranks ("replicas only": a state
    fits one GPU), costs all-gathered."""
    if shard and params is not None:
        from . import dist as _dist
        return _dist.sharded_over_parameter_sets(
            lambda rows: uncut_cost(circuit, backend, graph, rows, param_order, masked, max_states),
            params)
    if param_order is None:
        param_order = list(circuit.parameters)
    prog, auto = backend.compile(circuit, param_order=param_order)
    n = prog.n_qubits
    table = np.zeros((1, 0)) if params is None else np.asarray(params, dtype=np.float64)
    if table.ndim == 1:
        table = table.reshape(1, -1)
    n_sets = table.shape[0]
    # the MIS reduction is fused into the last gate group (engine.run_scored / qdc_sv_run_score): the final
    # state is scored tile by tile from shared memory, never re-read and never written back
    adj = _adjacency(graph, n)
    node_mask = (1 << graph.number_of_nodes()) - 1
    state_bytes = 16 << n
    if max_states is None:
        max_states = max(1, backend.max_state_bytes // state_bytes)
    chunk = min(n_sets, max_states)
    scratch = torch.empty(chunk, 1 << n, dtype=torch.complex128, device=backend.device)
    out = np.empty(n_sets)
    pending = []
    for lo in range(0, n_sets, chunk):
        hi = min(lo + chunk, n_sets)
        pending.append((lo, hi, backend.run_scored(prog, auto, table[lo:hi], adj, node_mask, scratch=scratch)))
    for lo, hi, sums in pending:                          # one read-back at the end: the chunks queue behind each other
        h = sums.cpu().numpy()
        out[lo:hi] = -(h[:, 1] if masked else h[:, 2])
    return out


def qaoa_plus_cost(circuit, backend: B200Backend, graph, Lambda: float, params=None,
                   param_order=None, max_states: Optional[int] = None) -> np.ndarray:
    """The QAOA+ cost closure f(params) = -expectation_value(counts, G, Lambda)
    (ansatz/qaoa_plus.py:40-53, 60-70) in the exact limit, for every parameter vector:
    -(sum_x p(x) |x| - Lambda sum_x p(x) #{edges inside x}), reduced on the device from the
    amplitudes."""
    if param_order is None:
        param_order = list(circuit.parameters)
    prog, auto = backend.compile(circuit, param_order=param_order)
    n = prog.n_qubits
    table = np.zeros((1, 0)) if params is None else np.asarray(params, dtype=np.float64)
    if table.ndim == 1:
        table = table.reshape(1, -1)
    n_sets = table.shape[0]
    adj = _adjacency(graph, n)
    node_mask = (1 << graph.number_of_nodes()) - 1
    if max_states is None:
        max_states = max(1, backend.max_state_bytes // (16 << n))
    chunk = min(n_sets, max_states)
    states = torch.empty(chunk, 1 << n, dtype=torch.complex128, device=backend.device)
    out = np.empty(n_sets)
    for lo in range(0, n_sets, chunk):
        hi = min(lo + chunk, n_sets)
        backend.run(prog, auto, table[lo:hi], out=states[:hi - lo])
        sums = [_rc.penalty_reduce(torch.view_as_real(states[s]).reshape(-1), n, adj, node_mask,
                                   index_flip=prog.flip_mask, amplitudes=True) for s in range(hi - lo)]
        for s, v in enumerate(torch.stack(sums).cpu().numpy()):
            out[lo + s] = -(v[2] - Lambda * v[1])
    return out


def top_strings(dist, graph, cutoff=1, threshold=1e-5, best_weight=0, num_anc=None):
    """Bitstrings among the `cutoff` most probable (p > threshold) that are independent sets
    heavier than `best_weight` (mis.py:232-242), from the fused on-device top-k when present.

    The reference strips the ancillas FIRST (helper_funcs.py:3-16 sums entries that differ only in
    ancilla bits) and applies `threshold` and `cutoff` afterwards; the device ranks full-width
    bitstrings.  The fused candidates are therefore merged by node string before the threshold and
    the cut-off are applied, and the fused list is only trusted when it cannot have missed a
    partner: every candidate above the threshold was returned (`n_candidates <= k`), or the merged
    list holds `cutoff` entries and the last of them beats the runner-up by more than the
    2^num_anc - 1 partners that may not have been returned could add (each is below the smallest
    probability returned).  Otherwise, when the distribution was stored, the reference's dict path
    runs; when it was not, the merged list is the best available answer."""
    from .utils.graph_funcs import is_indset
    from .utils.helper_funcs import hamming_weight
    n = graph.number_of_nodes()
    top = None
    score = getattr(dist, "score", None)
    if score is not None and len(score.top_x) >= cutoff:
        got = score.get()
        merged = {}
        for x, p in got["top"]:
            key = format(x & ((1 << n) - 1), f"0{n}b")
            merged[key] = merged.get(key, 0.0) + p               # insertion order = device rank
        ranked = sorted(((k, v) for k, v in merged.items() if v > threshold), key=lambda t: -t[1])
        complete = got.get("n_candidates", len(got["top"]) + 1) <= len(got["top"])
        anc = num_anc if num_anc is not None else max(getattr(dist, "n_wires", n) - n, 0)
        floor = min((p for _, p in got["top"]), default=0.0)       # nothing that was not returned is above this
        slack = ((1 << anc) - 1) * floor                           # what unseen partners could add to a string
        runner_up = ranked[cutoff][1] if len(ranked) > cutoff else floor
        safe = len(ranked) >= cutoff and (anc == 0 or ranked[cutoff - 1][1] > runner_up + slack)
        if complete or safe or getattr(dist, "P", None) is None:
            top = ranked
    if top is None:
        from .utils.helper_funcs import strip_ancillas
        anc = num_anc if num_anc is not None else dist.n_wires - n
        probs = strip_ancillas(dict(dist), num_anc=anc) if anc else dict(dist)
        top = sorted([(k, v) for k, v in probs.items() if v > threshold], key=lambda t: -t[1])
    top = top[:cutoff]
    return [(s, hamming_weight(s)) for s, _ in top if is_indset(s, graph)
            and hamming_weight(s) > best_weight]
