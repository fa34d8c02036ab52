"""`sim_with_cutting` -- the seam `mis.solve_mis_cut_dqva` calls -- plus hot-node selection.

Mirrors the reference's utils/cutting_funcs.py: `sim_with_cutting` :167-205 (same signature,
same `mode` values, same exception), `simple_choose_nodes` :125-164, `_cut_cost` :105-122,
`_is_connected` :89-102.  Node selection is host-side graph logic; `sim_with_cutting` runs the
fragment simulation, frame inversion and 4^K contraction on the GPU through the backend object
passed where the reference passes its Aer backend.
"""
from __future__ import annotations

import itertools
import random
import time
from typing import List, Tuple

import networkx as nx

__all__ = ["sim_with_cutting", "simple_choose_nodes"]


def _is_connected(graph, partition, hot_nodes, subgraph_dict):
    """Do the chosen hot nodes link all parts into one meta-graph? (:89-102)"""
    meta = nx.Graph()
    meta.add_nodes_from(range(len(partition)))
    for hot in hot_nodes:
        for u, v in graph.edges:
            if hot in (u, v) and subgraph_dict[u] != subgraph_dict[v]:
                meta.add_edge(subgraph_dict[u], subgraph_dict[v])
    return nx.is_connected(meta)


def _cut_cost(graph, partition, hot_nodes, subgraph_dict, cut_nodes):
    """Number of wire cuts a hot-node choice needs: for every cut node, the number of
    foreign subgraphs whose active mixers touch it (:105-122)."""
    seen = {c: set() for c in cut_nodes}
    for sub_idx, nodes in enumerate(partition):
        for node in nodes:
            if node in cut_nodes and node not in hot_nodes:
                continue                                  # mixer switched off
            for touched in list(graph.neighbors(node)) + [node]:
                if touched in seen:
                    seen[touched].add(sub_idx)
    return sum(1 for c, subs in seen.items() for s in subs if s != subgraph_dict[c])


def simple_choose_nodes(graph: nx.Graph, partition: List[List[int]],
                        cut_edges: List[Tuple[int, int]], max_cuts: int,
                        ) -> Tuple[List[int], List[int]]:
    """Uniformly random feasible hot-node set (:125-164).  Consumes one `random.choice`
    from the global `random` stream, like the reference."""
    subgraph_dict = {node: i for i, nodes in enumerate(partition) for node in nodes}
    cut_nodes = list(set(itertools.chain.from_iterable(cut_edges)))

    feasible = []
    for r in range(1, min(len(cut_nodes), max_cuts) + 1):
        for cand in itertools.combinations(cut_nodes, r):
            if not _is_connected(graph, partition, cand, subgraph_dict):
                continue
            if _cut_cost(graph, partition, cand, subgraph_dict, cut_nodes) <= max_cuts:
                feasible.append(cand)
    hot_nodes = random.choice(feasible)
    return cut_nodes, list(hot_nodes)


def sim_with_cutting(fragments, wire_path_map, frag_shots, backend, mode="likely",
                     verbose=0, params=None, param_order=None, threshold=1e-5, topk=1,
                     store=None, compact="auto", sample=False, seed=None, contraction=None):
    """Simulate a fragmented circuit and recombine (utils/cutting_funcs.py:167-205).

    Returns the recombined distribution as a Mapping `{bitstring: probability}` that keeps the
    device array and, when `backend.graph` is set, the fused MIS objective / top-k (`.score`).
    `backend` must be a `qdc_b200.engine.B200Backend`; `frag_shots` only scales the dict views
    (the engine evaluates the infinite-shot limit).  Extensions: `params`
    [n_sets, n_params] + `param_order` evaluate a batch of parameter vectors of unbound fragments
    in one pass and return a list of distributions; `compact` as in
    `recombine_fragment_models` (contract only over fragment outcomes that occur);
    `contraction` as in `recombine_fragment_models` ("f64" or "i8[:S]");
    `sample=True` draws `frag_shots` outcomes per variant circuit on the device (the reference's
    finite-shot behaviour) instead of the exact limit."""
    from ..qsplit import qsplit_mlrecon_methods as qmm

    t0 = time.time()
    frag_data = qmm.collect_fragment_data(fragments, wire_path_map, shots=frag_shots,
                                          tomography_backend=backend, params=params,
                                          param_order=param_order, sample=sample, seed=seed)
    direct_models = qmm.direct_fragment_model(frag_data)
    if mode == "direct":
        models = direct_models
    elif mode == "likely":
        models = qmm.maximum_likelihood_model(direct_models)
    else:
        raise Exception('Unknown recombination mode:', mode)
    model_time = time.time() - t0

    t1 = time.time()
    recombined_dist = qmm.recombine_fragment_models(models, wire_path_map, backend=backend,
                                                    threshold=threshold, topk=topk, store=store,
                                                    compact=compact, contraction=contraction)
    recombine_time = time.time() - t1
    if verbose:
        print(f"\tModel time: {model_time:.3f}, Recombine time: {recombine_time:.3f}")
    return recombined_dist
