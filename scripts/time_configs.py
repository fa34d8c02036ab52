"""Ad-hoc timing of whole cost evaluations (sim_with_cutting: variants -> tensors -> contraction +
MIS scoring) at the BASELINE config shapes C1 / C2 / C3, with and without support compaction."""
import importlib, os, random, sys, time
import networkx as nx, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
importlib.import_module("dqva-and-circuit-cutting_b200")
from qdc_b200.engine import B200Backend
from qdc_b200.utils import graph_funcs
from qdc_b200.utils.cutting_funcs import sim_with_cutting, simple_choose_nodes
from qdc_b200.ansatz import dqv_cut_ansatz
from qdc_b200.circuit import Parameter
from qdc_b200.qsplit import qsplit_circuit_cutter as qcc


def build(G, partition, max_cuts, seed, fused):
    random.seed(seed); np.random.seed(seed)
    n = G.number_of_nodes()
    _, cut_edges = graph_funcs.get_subgraphs(G, partition)
    cut_nodes, hot_nodes = simple_choose_nodes(G, partition, cut_edges, max_cuts)
    order = [int(v) for part in partition for v in np.random.permutation(part)]
    params = [Parameter(f"var_{i}") for i in range(n + 1)]
    circ, cuts = dqv_cut_ansatz.gen_dqva(G, partition, cut_nodes, hot_nodes, order, params=params,
                                         init_state="0" * n, fused=fused)
    frags, wpm = qcc.cut_circuit(circ, cuts)
    return circ, cuts, frags, wpm


def kl(G, max_cuts):
    for seed in range(20):
        parts = [sorted(p) for p in nx.algorithms.community.kernighan_lin_bisection(G, seed=seed)]
        _, cut_edges = graph_funcs.get_subgraphs(G, parts)
        try:
            random.seed(0); simple_choose_nodes(G, parts, cut_edges, max_cuts); return parts
        except IndexError:
            continue
    raise SystemExit("no feasible partition")


def run(name, G, partition, max_cuts, n_sets, seed=1):
    circ, cuts, frags, wpm = build(G, partition, max_cuts, seed, fused=False)
    used = circ.parameters
    table = np.random.default_rng(1).uniform(0, 2 * np.pi, (n_sets, len(used)))
    be = B200Backend(device=torch.device("cuda:0"), graph=G)
    N = circ.num_qubits
    widths = [f.num_qubits for f in frags]
    for compact in (False, True):
        def go():
            d = sim_with_cutting(frags, wpm, 1, be, mode="direct", params=table, param_order=used,
                                 store=False, compact=compact)
            d = d if isinstance(d, list) else [d]
            return [x.score.get() for x in d]
        go(); torch.cuda.synchronize()
        t0 = time.time(); out = go(); torch.cuda.synchronize(); t = time.time() - t0
        print(f"{name}: N={N} frags={widths} cuts={len(cuts)} sets={n_sets} compact={compact}: "
              f"{t*1e3/n_sets:.2f} ms per evaluation (wall, incl. host), {n_sets*2.0**N/t:.3e} probs/s, "
              f"cost[0]={-out[0]['sum_pw_indset']:.9f} sum_p={out[0]['sum_p']:.12f}", flush=True)


if __name__ == "__main__":
    gd = os.path.join(ROOT, "tests", "golden", "graphs")
    G = graph_funcs.graph_from_file(os.path.join(gd, "N18_com2_pin80_pout2_G1.txt"))
    run("C1", G, kl(G, 1), 1, 64)
    G = graph_funcs.graph_from_file(os.path.join(gd, "N26_com2_pin20_pout10_G1.txt"))
    for mc in (2, 3):
        run(f"C2 max_cuts={mc}", G, kl(G, mc), mc, 64)
    G = graph_funcs.graph_from_file(os.path.join(gd, "N33_com3_pin80_pout2_G1.txt"))
    part = [list(range(0, 11)), list(range(11, 22)), list(range(22, 33))]
    for mc in (2, 3, 4):
        try:
            run(f"C3 max_cuts={mc}", G, part, mc, 4, seed=2); break
        except IndexError:
            continue
