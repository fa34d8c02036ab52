"""GPU parity at the shapes BASELINE.json names: C1 = N18_com2_pin80_pout2/G1, 2
fragments, 1 cut; C2 = N26_com2_pin20_pout10/G1, 2 fragments, 2-3 cuts, batch of parameter
vectors; C3 = N33_com3_pin80_pout2/G1, 3 fragments.  The cut path (gate-faithful fragments ->
engine -> Pauli tensors -> contraction + MIS scoring) is compared with the UNCUT circuit: numpy
oracle statevector where it fits (C1), the engine's own uncut simulation beyond that (C2; the
engine is pinned to the oracle at <= 16 qubits in test_gpu_statevec.py), and size-independent
properties where nothing else fits (C3: total probability, non-negativity of the top entry)."""
import os
import random

import networkx as nx
import numpy as np
import pytest
import torch

from oracle import mis_oracle as mo
from oracle import statevector as sv

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _graph(golden_dir, name):
    from qdc_b200.utils import graph_funcs
    return graph_funcs.graph_from_file(os.path.join(golden_dir, "graphs", name + ".txt"))


def _cut_circuit(G, partition, max_cuts, seed, fused, symbolic=False):
    from qdc_b200.ansatz import dqv_cut_ansatz
    from qdc_b200.circuit import Parameter
    from qdc_b200.qsplit import qsplit_circuit_cutter as qcc
    from qdc_b200.utils import graph_funcs
    from qdc_b200.utils.cutting_funcs import simple_choose_nodes
    random.seed(seed)
    np.random.seed(seed)
    n = G.number_of_nodes()
    _, cut_edges = graph_funcs.get_subgraphs(G, partition)
    cut_nodes, hot_nodes = simple_choose_nodes(G, partition, cut_edges, max_cuts)
    order = [int(v) for part in partition for v in np.random.permutation(part)]
    params = ([Parameter(f"var_{i}") for i in range(n + 1)] if symbolic
              else list(np.random.uniform(0, 2 * np.pi, n + 1)))
    circ, cuts = dqv_cut_ansatz.gen_dqva(G, partition, cut_nodes, hot_nodes, order, params=params,
                                         init_state="0" * n, fused=fused)
    frags, wpm = qcc.cut_circuit(circ, cuts)
    return circ, cuts, frags, wpm, (cut_nodes, hot_nodes, order, params)


def _kl_partition(G, max_cuts):
    """Kernighan-Lin bisection (the reference's default partition_alg, mis.py:145-149); the first
    seed whose cut admits a hot-node set within max_cuts is used and recorded."""
    from qdc_b200.utils import graph_funcs
    from qdc_b200.utils.cutting_funcs import simple_choose_nodes
    for seed in range(20):
        parts = [sorted(p) for p in nx.algorithms.community.kernighan_lin_bisection(G, seed=seed)]
        _, cut_edges = graph_funcs.get_subgraphs(G, parts)
        try:
            random.seed(0)
            simple_choose_nodes(G, parts, cut_edges, max_cuts)
            return parts, seed
        except IndexError:
            continue
    pytest.skip("no feasible hot-node set for this graph / max_cuts")


def test_c1_n18_one_cut_matches_uncut_oracle(cuda, golden_dir):
    from qdc_b200.engine import B200Backend
    from qdc_b200.utils.cutting_funcs import sim_with_cutting
    G = _graph(golden_dir, "N18_com2_pin80_pout2_G1")
    partition, kl_seed = _kl_partition(G, 1)
    circ, cuts, frags, wpm, meta = _cut_circuit(G, partition, 1, 0, fused=False)
    assert len(frags) == 2 and len(cuts) == 1
    be = B200Backend(device=cuda, graph=G)
    dist = sim_with_cutting(frags, wpm, 1000, be, mode="likely", topk=3)
    # uncut reference: numpy oracle on the fused form of the same circuit (same unitary)
    from qdc_b200.ansatz import dqv_cut_ansatz
    cut_nodes, hot_nodes, order, params = meta
    fused_circ, _ = dqv_cut_ansatz.gen_dqva(G, partition, cut_nodes, hot_nodes, order, params=params,
                                            init_state="0" * 18, fused=True)
    want = sv.probabilities(sv.simulate(fused_circ.num_qubits, fused_circ.to_instructions()))
    got = dist.P.cpu().numpy()
    assert got.shape == want.shape == (1 << 20,)
    assert np.abs(got - want).max() < TOL
    s_p, s_pwok = mo.score_array(want, 18, 2, mo.adjacency_masks(G, 18))
    out = dist.score.get()
    assert abs(out["sum_p"] - 1) < TOL and abs(out["sum_pw_indset"] - s_pwok) < TOL
    idx = np.nonzero(want > 1e-5)[0]
    order_ = idx[np.lexsort((idx, -want[idx]))][:3]
    assert [x for x, _ in out["top"]] == [int(i) for i in order_]


@pytest.mark.parametrize("max_cuts", [2, 3])
def test_c2_n26_parameter_batch_matches_uncut_engine(cuda, golden_dir, max_cuts):
    from qdc_b200 import recombine as rc
    from qdc_b200.engine import B200Backend
    from qdc_b200.utils.cutting_funcs import sim_with_cutting
    G = _graph(golden_dir, "N26_com2_pin20_pout10_G1")
    partition, kl_seed = _kl_partition(G, max_cuts)
    circ, cuts, frags, wpm, meta = _cut_circuit(G, partition, max_cuts, 1, fused=False, symbolic=True)
    assert len(frags) == 2 and 1 <= len(cuts) <= max_cuts
    used = circ.parameters
    n_sets = 3
    table = np.random.default_rng(1).uniform(0, 2 * np.pi, (n_sets, len(used)))
    be = B200Backend(device=cuda, graph=G)
    dists = sim_with_cutting(frags, wpm, 1, be, mode="direct", params=table, param_order=used,
                             store=True)
    if n_sets == 1:
        dists = [dists]
    N = circ.num_qubits                                   # 26 + 2 ancillas
    prog, auto = be.compile(circ, param_order=used)
    adj = [sum(1 << int(v) for v in G.neighbors(i)) for i in range(26)] + [0, 0]
    for s in range(n_sets):
        state = be.run(prog, auto, table[s:s + 1])
        want = be.probabilities(state, prog)[0]
        got = dists[s].P
        assert float((got - want).abs().max()) < TOL
        tabs = rc.build_tables(N - 14, 14, list(range(N)), adj, (1 << 26) - 1, device=cuda,
                               index_flip=prog.flip_mask)
        ref = rc.reduce_and_score(torch.view_as_real(state[0]).reshape(-1), tabs, amplitudes=True).get()
        out = dists[s].score.get()
        assert abs(out["sum_p"] - 1) < 1e-9
        assert abs(out["sum_pw_indset"] - ref["sum_pw_indset"]) < 1e-9
        assert abs(out["sum_pw"] - ref["sum_pw"]) < 1e-9
        del state, want


def test_c3_n33_three_fragments_properties(cuda, golden_dir):
    """2^36 outputs reduced on the fly (nothing stored): total probability and sanity of top-1."""
    from qdc_b200.engine import B200Backend
    from qdc_b200.utils.cutting_funcs import sim_with_cutting
    G = _graph(golden_dir, "N33_com3_pin80_pout2_G1")
    partition = [list(range(0, 11)), list(range(11, 22)), list(range(22, 33))]   # planted split
    for max_cuts in (2, 3, 4):
        try:
            circ, cuts, frags, wpm, meta = _cut_circuit(G, partition, max_cuts, 2, fused=False)
            break
        except IndexError:
            continue
    else:
        pytest.skip("no feasible hot-node set")
    assert len(frags) == 3
    be = B200Backend(device=cuda, graph=G)
    dist = sim_with_cutting(frags, wpm, 1, be, mode="direct", store=False, threshold=1e-3, topk=2)
    out = dist.score.get()
    assert abs(out["sum_p"] - 1) < 1e-9
    assert 0 <= out["sum_pw_indset"] <= out["sum_pw"] + 1e-12 <= 33
    assert all(0 < p <= 1 for _, p in out["top"])


def test_c3_n33_int8_contraction_matches_fp64(cuda, golden_dir):
    """The same 3-fragment, 2^36-output evaluation through the int8-sliced contraction
    (contraction="i8:6" and "i8:5") against the FP64 DMMA kernel: costs within 1e-10, same top-k."""
    from qdc_b200.engine import B200Backend
    from qdc_b200.utils.cutting_funcs import sim_with_cutting
    G = _graph(golden_dir, "N33_com3_pin80_pout2_G1")
    partition = [list(range(0, 11)), list(range(11, 22)), list(range(22, 33))]
    for max_cuts in (2, 3, 4):
        try:
            circ, cuts, frags, wpm, meta = _cut_circuit(G, partition, max_cuts, 2, fused=False)
            break
        except IndexError:
            continue
    else:
        pytest.skip("no feasible hot-node set")
    be = B200Backend(device=cuda, graph=G)
    kw = dict(mode="direct", store=False, threshold=1e-3, topk=2)
    ref = sim_with_cutting(frags, wpm, 1, be, contraction="f64", **kw).score.get()
    for spec in ("i8:6", "i8:5"):
        out = sim_with_cutting(frags, wpm, 1, be, contraction=spec, **kw).score.get()
        for key in ("sum_p", "sum_pw", "sum_pw_indset"):
            assert abs(out[key] - ref[key]) < 1e-10, (spec, key, out[key], ref[key])
        assert [x for x, _ in out["top"]] == [x for x, _ in ref["top"]]
        for (_, p), (_, q) in zip(out["top"], ref["top"]):
            assert abs(p - q) < 1e-10


# ---- config-level pins against the ORACLE (not against this engine) ------------------------------

@pytest.mark.parametrize("max_cuts", [2, 3])
def test_c2_sub_instance_matches_numpy_oracle(cuda, golden_dir, max_cuts):
    """C2's graph (N26_com2_pin20_pout10/G1) restricted to 10 + 10 of its nodes, so that the UNCUT
    circuit (22 qubits) fits the numpy oracle: a batch of parameter vectors through the cut path,
    every one of the 2^22 stored probabilities and the three sums against oracle.statevector +
    oracle.mis_oracle (the full 26-node instance above can only be compared with this engine)."""
    from qdc_b200.ansatz import dqv_cut_ansatz
    from qdc_b200.engine import B200Backend
    from qdc_b200.utils.cutting_funcs import sim_with_cutting
    G26 = _graph(golden_dir, "N26_com2_pin20_pout10_G1")
    keep = list(range(0, 10)) + list(range(13, 23))
    G = nx.relabel_nodes(G26.subgraph(keep), {v: i for i, v in enumerate(keep)})
    n = 20
    partition = [list(range(10)), list(range(10, 20))]
    try:
        circ, cuts, frags, wpm, meta = _cut_circuit(G, partition, max_cuts, 3, fused=False, symbolic=True)
    except IndexError:
        pytest.skip("no feasible hot-node set for this sub-instance / max_cuts")
    assert len(frags) == 2 and 1 <= len(cuts) <= max_cuts
    cut_nodes, hot_nodes, order, syms = meta
    used = circ.parameters
    n_sets = 3
    table = np.random.default_rng(7).uniform(0, 2 * np.pi, (n_sets, len(used)))
    be = B200Backend(device=cuda, graph=G)
    dists = sim_with_cutting(frags, wpm, 1, be, mode="likely", params=table, param_order=used, store=True, topk=3)
    adj = mo.adjacency_masks(G, n)
    for s in range(n_sets):
        value = dict(zip([p.name for p in used], table[s]))
        numeric = [value.get(p.name, 0.0) for p in syms]            # unused symbols never enter the circuit
        fused_circ, _ = dqv_cut_ansatz.gen_dqva(G, partition, cut_nodes, hot_nodes, order, params=numeric,
                                                init_state="0" * n, fused=True)
        want = sv.probabilities(sv.simulate(fused_circ.num_qubits, fused_circ.to_instructions()))
        got = dists[s].P.cpu().numpy()
        assert got.shape == want.shape == (1 << (n + 2),)
        assert np.abs(got - want).max() < TOL
        s_p, s_pwok = mo.score_array(want, n, 2, adj, masked=True)
        _, s_pw = mo.score_array(want, n, 2, adj, masked=False)
        out = dists[s].score.get()
        assert abs(out["sum_p"] - s_p) < TOL and abs(out["sum_pw_indset"] - s_pwok) < TOL
        assert abs(out["sum_pw"] - s_pw) < TOL
        idx = np.nonzero(want > 1e-5)[0]
        best = idx[np.lexsort((idx, -want[idx]))][:3]
        assert [x for x, _ in out["top"]] == [int(i) for i in best]


def _oracle_wire_marginals(frags, wpm):
    """P(circuit wire i = 1) of the recombined distribution from the ORACLE's fragment tensors: by
    linearity the other output bits are summed inside each fragment before the contraction over the
    cut-basis index (oracle.qsplit_oracle: fragment_pauli_tensor, _network_subscripts,
    united_axis_permutation), so 2^36 outputs never exist on the CPU either."""
    from oracle import qsplit_oracle as qo
    targets = qo.identify_frag_targets(wpm)
    Ts = [qo.fragment_pauli_tensor(frag, targets[f]["prep"], targets[f]["meas"]) for f, frag in enumerate(frags)]
    labels = qo._network_subscripts(wpm)
    F = len(Ts)

    def contract(tensors):
        operands, out = [], []
        for f in reversed(range(F)):
            operands += [tensors[f], [40 + f] + labels[f]]
            out.append(40 + f)
        return np.einsum(*operands, out)

    sums = [T.sum(axis=0, keepdims=True) for T in Ts]
    norm = float(contract(sums).reshape(-1)[0])
    widths = [int(np.log2(T.shape[0])) for T in Ts]
    N = sum(widths)
    perm = qo.united_axis_permutation(wpm)
    united_of_circuit = [N - 1 - perm[N - 1 - i] for i in range(N)]    # as in qo.recombine_arrays
    marg = np.zeros(N)
    for i in range(N):
        u, f = united_of_circuit[i], 0
        while u >= widths[f]:
            u -= widths[f]
            f += 1
        b = np.arange(Ts[f].shape[0])
        one = Ts[f][((b >> u) & 1) == 1].sum(axis=0, keepdims=True)
        tensors = list(sums)
        tensors[f] = one
        marg[i] = float(contract(tensors).reshape(-1)[0]) / norm
    return marg


def test_c3_n33_wire_marginals_match_oracle(cuda, golden_dir):
    """C3 (N33, 3 fragments, 2^36 outputs, nothing stored) against the oracle, not only against
    sum p = 1: with an edgeless graph on the first k wires the fused score sum p w equals
    sum_x p(x) popcount(x & (2^k - 1)), so consecutive k give P(wire = 1) for every wire -- each
    one a full pass of the contraction + scoring over all 2^36 outputs -- and the oracle gets the
    same marginals from its own fragment tensors."""
    from qdc_b200.engine import B200Backend
    from qdc_b200.qsplit import qsplit_mlrecon_methods as qmm
    G = _graph(golden_dir, "N33_com3_pin80_pout2_G1")
    partition = [list(range(0, 11)), list(range(11, 22)), list(range(22, 33))]
    for max_cuts in (2, 3, 4):
        try:
            circ, cuts, frags, wpm, meta = _cut_circuit(G, partition, max_cuts, 2, fused=False)
            break
        except IndexError:
            continue
    else:
        pytest.skip("no feasible hot-node set")
    assert len(frags) == 3
    N = circ.num_qubits
    want = _oracle_wire_marginals(frags, wpm)
    be = B200Backend(device=cuda, graph=None)
    models = qmm.direct_fragment_model(qmm.collect_fragment_data(frags, wpm, shots=1, tomography_backend=be))
    prefix = [0.0]
    for k in range(1, N + 1):
        out = qmm.recombine_fragment_models(models, wpm, graph=nx.empty_graph(k), store=False,
                                            threshold=2.0).score.get()
        assert abs(out["sum_p"] - 1.0) < 1e-9
        assert abs(out["sum_pw"] - out["sum_pw_indset"]) < 1e-12        # no edges: every string is independent
        prefix.append(out["sum_pw"])
    got = np.diff(np.array(prefix))
    assert np.abs(got - want).max() < 1e-9, float(np.abs(got - want).max())
    # the DQVA ancillas return to |0>
    n_nodes = G.number_of_nodes()
    assert np.abs(want[n_nodes:]).max() < 1e-12 and np.abs(got[n_nodes:]).max() < 1e-9


def test_c5_uncut_24_nodes_matches_oracle(cuda):
    """C5's path (uncut DQVA statevector + MIS cost straight from the amplitudes, mis.uncut_cost) at
    24 nodes + 1 ancilla = 25 qubits against oracle.statevector + oracle.mis_oracle, and the QAOA+
    objective (ansatz/qaoa_plus.py:40-53) at the same size; the engine compiles the gate-faithful
    circuit (ancilla + MCX mixers), the oracle simulates the fused form of the same unitary."""
    from qdc_b200 import mis
    from qdc_b200.ansatz import dqv_ansatz, qaoa_plus
    from qdc_b200.circuit import Parameter
    from qdc_b200.engine import B200Backend
    n = 24
    G = nx.random_regular_graph(3, n, seed=11)
    syms = [Parameter(f"var_{i}") for i in range(n + 1)]
    order = [int(v) for v in np.random.default_rng(3).permutation(n)]
    circ = dqv_ansatz.gen_dqva(G, P=1, params=syms, init_state="0" * n, mixer_order=order)
    used = circ.parameters
    n_sets = 1                                            # (the 25-qubit numpy oracle takes ~25 s per vector)
    table = np.random.default_rng(5).uniform(0, 2 * np.pi, (n_sets, len(used)))
    be = B200Backend(device=cuda)
    got = mis.uncut_cost(circ, be, G, params=table, param_order=used)
    got_masked = mis.uncut_cost(circ, be, G, params=table, param_order=used, masked=True)
    adj = mo.adjacency_masks(G, n)
    for s in range(n_sets):
        value = dict(zip([p.name for p in used], table[s]))
        numeric = [value.get(p.name, 0.0) for p in syms]
        fused = dqv_ansatz.gen_dqva(G, P=1, params=numeric, init_state="0" * n, mixer_order=order, fused=True)
        p = sv.probabilities(sv.simulate(n + 1, fused.to_instructions()))
        _, s_pw = mo.score_array(p, n, 1, adj, masked=False)
        _, s_pwok = mo.score_array(p, n, 1, adj, masked=True)
        assert abs(got[s] + s_pw) < TOL and abs(got_masked[s] + s_pwok) < TOL
        del p
    # QAOA+ at the same width: -(sum p |x| - Lambda sum p #edges(x))
    qc = qaoa_plus.construct_qaoa_plus(1, G, [0.37, 1.1])
    cost = mis.qaoa_plus_cost(qc, be, G, Lambda=1.5)
    p = sv.probabilities(sv.simulate(n, qc.to_instructions()))
    xs = np.arange(1 << n, dtype=np.uint64)
    weight = mo.popcount64(xs)
    edges = np.zeros(1 << n)
    for a, b in G.edges():
        edges += ((xs >> np.uint64(a)) & np.uint64(1)) * ((xs >> np.uint64(b)) & np.uint64(1))
    want = -(float(p @ weight) - 1.5 * float(p @ edges))
    assert abs(cost[0] - want) < 1e-9, (cost[0], want)


def test_c5_uncut_30_qubits_properties(cuda):
    """C5 at its own width (29 nodes + 1 ancilla = 30 qubits, 17 GB state): nothing on the CPU fits,
    so size-independent properties: the all-zero parameter vector leaves |0...0> (cost 0 exactly);
    the masked cost never exceeds the unmasked one; DQVA from an independent-set start only ever
    populates independent sets, so both are equal; and the cost is bounded by the graph's
    independence bound n."""
    from qdc_b200 import mis
    from qdc_b200.ansatz import dqv_ansatz
    from qdc_b200.circuit import Parameter
    from qdc_b200.engine import B200Backend
    if torch.cuda.get_device_properties(cuda).total_memory < 60e9:
        pytest.skip("needs a 17 GB state plus workspace")
    n = 29
    G = nx.random_regular_graph(4, n + 1, seed=2)
    G.remove_node(n)                                      # 29 nodes, degrees 3-4
    syms = [Parameter(f"var_{i}") for i in range(n + 1)]
    circ = dqv_ansatz.gen_dqva(G, P=1, params=syms, init_state="0" * n, mixer_order=list(range(n)))
    used = circ.parameters
    table = np.zeros((2, len(used)))
    table[1] = np.random.default_rng(9).uniform(0, 2 * np.pi, len(used))
    be = B200Backend(device=cuda)
    plain = mis.uncut_cost(circ, be, G, params=table, param_order=used, max_states=1)
    masked = mis.uncut_cost(circ, be, G, params=table, param_order=used, masked=True, max_states=1)
    assert plain[0] == 0.0 and masked[0] == 0.0
    assert -n <= plain[1] < -0.5                          # a random angle vector flips some nodes
    assert abs(plain[1] - masked[1]) < 1e-9               # DQVA stays inside the independent sets
