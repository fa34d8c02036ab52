"""`solve_mis_cut_dqva` on the B200 path (reference: mis.py:32-263), SURVEY.md 8(f) items 1-2.

Same outer structure as the reference -- partition the graph, m rounds of mixer permutations,
inner rounds that re-cut the circuit around the current independent set, optimise the fragment
parameters, read the most probable bitstrings, keep improvements -- and the same return tuple
and `history` record (mis.py:245-253).  What changes is the evaluation: `avg_cost` is
`qdc_b200.mis.cut_cost` (variants -> tensors -> contraction + fused MIS score on the device), and
the optimiser can be *batched*: the reference's COBYLA calls the cost one vector at a time
(mis.py:221), `optimizer="SPSA"` here evaluates a whole population of perturbed vectors per
iteration in one device pass (the "64 parameter sets" batch of BASELINE configs[1]).

Differences that are deliberate:
  * parameters are bound by name over the whole circuit; the reference binds each fragment
    positionally from index 0 over an unordered qiskit parameter set (A.6-3), which is not
    reproducible;
  * the evaluation is the infinite-shot limit, `shots` only sets `frag_shots` in the history;
  * partition_alg 'metis' needs the metis package (absent here): 'klb' (networkx
    Kernighan-Lin, the reference's other option, mis.py:145-147) or an explicit `partition`.
"""
from __future__ import annotations

import copy
import time
from typing import List, Optional

import networkx as nx
import numpy as np

from . import _lib, mis as _mis
from .ansatz import dqv_cut_ansatz
from .circuit import Parameter
from .engine import B200Backend
from .qsplit import qsplit_circuit_cutter as qcc
from .qsplit import qsplit_mlrecon_methods as qmm
from .utils.cutting_funcs import simple_choose_nodes, sim_with_cutting
from .utils.graph_funcs import get_subgraphs, is_indset
from .utils.helper_funcs import hamming_weight, strip_ancillas

__all__ = ["solve_mis_cut_dqva", "minimize_spsa"]


def minimize_spsa(batch_cost, x0, iterations=30, pairs=16, a=0.6, c=0.25, seed=None):
    """Batched SPSA: every iteration draws `pairs` Rademacher directions, evaluates the 2 * pairs
    perturbed vectors (+ the current one) in ONE call of `batch_cost(table) -> costs`, averages the
    gradient estimates and steps.  Returns a dict shaped like scipy's OptimizeResult
    ('x', 'fun', 'nfev')."""
    rng = np.random.default_rng(seed)
    x = np.asarray(x0, dtype=np.float64).copy()
    best_x, best_f, nfev = x.copy(), np.inf, 0
    for it in range(iterations):
        ak = a / (it + 1 + 0.1 * iterations) ** 0.602
        ck = c / (it + 1) ** 0.101
        delta = rng.choice([-1.0, 1.0], size=(pairs, x.size))
        table = np.concatenate([x[None], x + ck * delta, x - ck * delta])
        costs = np.asarray(batch_cost(table), dtype=np.float64)
        nfev += table.shape[0]
        j = int(np.argmin(costs))
        if costs[j] < best_f:
            best_f, best_x = float(costs[j]), table[j].copy()
        grad = ((costs[1:1 + pairs] - costs[1 + pairs:])[:, None] / (2 * ck) * delta).mean(axis=0)
        x = x - ak * grad
    final = float(np.asarray(batch_cost(x[None]))[0])
    nfev += 1
    if final < best_f:
        best_f, best_x = final, x
    return {"x": best_x, "fun": best_f, "nfev": nfev}


def _partition(graph, partition_alg, num_frags, partition):
    if partition is not None:
        return [list(p) for p in partition]
    if partition_alg == "klb":
        if num_frags != 2:
            raise ValueError("Kernighan-Lin bisection gives 2 subgraphs; pass `partition` for more")
        return [sorted(p) for p in nx.algorithms.community.kernighan_lin_bisection(graph)]
    if partition_alg == "metis":
        try:
            import metis                                          # della_requirements.txt:50
        except ImportError as exc:
            raise _lib.QdcError("partition_alg='metis' needs the metis package (not installed); "
                                "use 'klb' or pass partition=[[...], ...]") from exc
        assignment = metis.part_graph(graph, nparts=num_frags)[1]
        parts = [[] for _ in set(assignment)]
        for node, a in zip(list(graph), assignment):
            parts[a].append(node)
        return parts
    raise ValueError(f"Unknown graph partitioning algorithm: {partition_alg}")


def solve_mis_cut_dqva(init_state, graph, P=1, m=4, threshold=1e-5, cutoff=1, sim="b200",
                       shots=8192, verbose=0, max_cuts=1, num_frags=2, optimizer="COBYLA",
                       partition_alg="klb", backend: Optional[B200Backend] = None, partition=None,
                       optimizer_options: Optional[dict] = None, fused=False):
    """Find a large independent set of `graph` with the cut DQVA (mis.py:32-263).

    Returns (best_indset, best_params, best_init_state, best_perm, partition, cut_nodes, history).
    optimizer: any scipy.optimize.minimize method name (sequential, like the reference) or "SPSA"
    (batched, `optimizer_options` -> minimize_spsa)."""
    if max_cuts < num_frags - 1:
        raise ValueError(f"Number of cuts {max_cuts} is too few for {num_frags} fragments")
    if backend is None:
        backend = B200Backend(graph=graph)
    backend.graph = graph
    opts = dict(optimizer_options or {})

    partition = _partition(graph, partition_alg, num_frags, partition)
    subgraphs, cut_edges = get_subgraphs(graph, partition)
    subgraph_dict = {q: i for i, sub in enumerate(subgraphs) for q in sub}
    if verbose:
        print(f"Partitioned graph into {len(subgraphs)} subgraphs = {partition}\nwith cut edges = {cut_edges}")

    def sort_mixers(order):                                       # mis.py:42-56
        groups = [[] for _ in range(len(set(subgraph_dict.values())))]
        for node in order:
            groups[subgraph_dict[node]].append(node)
        out = []
        for g in np.random.permutation(np.array(groups, dtype=object)):
            out.extend(g)
        return [int(v) for v in out]

    n = graph.number_of_nodes()
    cur_permutation = sort_mixers(list(np.random.permutation(list(graph.nodes))))
    best_indset = best_init_state = cur_init_state = init_state
    best_params, best_perm = None, copy.copy(cur_permutation)
    cut_nodes, history = [], []
    new_hamming_weight = hamming_weight(cur_init_state)

    for mixer_round in range(1, m + 1):
        mixer_history, inner_round = [], 1
        while True:
            if verbose:
                print(f"\nStart round {mixer_round}.{inner_round}, Initial state = {cur_init_state}")
            num_params = P * (n + 1)
            params = [Parameter(f"var_{i}") for i in range(num_params)]
            cut_nodes, hot_nodes = simple_choose_nodes(graph, partition, cut_edges, max_cuts)
            if len(hot_nodes) == 0:
                raise Exception("No hot nodes selected!")          # mis.py:72-76
            circuit, cuts = dqv_cut_ansatz.gen_dqva(graph, partition, cut_nodes, hot_nodes,
                                                    cur_permutation, params=params,
                                                    init_state=cur_init_state, verbose=0, P=P,
                                                    fused=fused)
            fragments, wire_path_map = qcc.cut_circuit(circuit, cuts)
            if len(cuts) > max_cuts:
                raise Exception("TOO MANY CUTS!")
            if len(cuts) == 0:
                raise Exception("DIDNT FIND ANY CUTS!")
            if len(fragments) != len(partition):
                raise Exception("WRONG NUMBER OF FRAGMENTS!")
            used = circuit.parameters
            num_used_params = len(used)
            frag_shots = shots // qmm.fragment_variants(wire_path_map)

            def batch_cost(table):
                return _mis.cut_cost(fragments, wire_path_map, frag_shots, backend, graph,
                                     params=np.atleast_2d(table), param_order=used)

            init_params = np.random.uniform(low=0.0, high=2 * np.pi, size=num_used_params)
            t0 = time.time()
            if optimizer == "SPSA":
                out = minimize_spsa(batch_cost, init_params, **opts)
            else:
                from scipy.optimize import minimize
                out = minimize(lambda x: float(batch_cost(x)[0]), init_params, method=optimizer,
                               options=opts or None)
            opt_params, opt_cost = np.asarray(out["x"]), float(out["fun"])
            if verbose:
                print(f"\tOptimal cost: {opt_cost}\n\t{out['nfev']} function evaluations in "
                      f"{time.time() - t0:.2f} s")

            # distribution at the optimum -> top `cutoff` strings above threshold (mis.py:229-242)
            dist = sim_with_cutting(fragments, wire_path_map, frag_shots, backend,
                                    params=opt_params[None], param_order=used,
                                    threshold=threshold, topk=max(cutoff, 1), store=False)
            best_hamming_weight = hamming_weight(best_indset)
            better_strs = _mis.top_strings(dist, graph, cutoff=cutoff, threshold=threshold,
                                           best_weight=best_hamming_weight,
                                           num_anc=len(fragments))
            better_strs = sorted(better_strs, key=lambda t: t[1], reverse=True)
            mixer_history.append({
                "mixer_round": mixer_round, "inner_round": inner_round, "cost": opt_cost,
                "function_evals": out["nfev"], "init_state": cur_init_state,
                "mixer_order": copy.copy(cur_permutation), "num_params": num_used_params,
                "frag_shots": frag_shots, "frag_qubits": [f.num_qubits for f in fragments],
                "cuts": cuts, "hot_nodes": hot_nodes, "better_strs": better_strs,
                "cur_params": opt_params, "cur_args": (fragments, wire_path_map, frag_shots)})
            if len(better_strs) == 0:
                if verbose:
                    print("\tNone of the measured bitstrings had higher Hamming weight than:", best_indset)
                break
            best_indset, new_hamming_weight = better_strs[0]
            best_init_state, best_params = cur_init_state, opt_params
            best_perm = copy.copy(cur_permutation)
            cur_init_state = best_indset
            if verbose:
                print(f"\tFound new independent set: {best_indset}, Hamming weight = {new_hamming_weight}")
            inner_round += 1
        history.append(mixer_history)
        cur_permutation = sort_mixers(list(np.random.permutation(list(graph.nodes))))
    if verbose:
        print("\tRETURNING, best hamming weight:", new_hamming_weight)
    return best_indset, best_params, best_init_state, best_perm, partition, cut_nodes, history
