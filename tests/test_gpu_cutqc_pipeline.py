"""GPU parity of the CutQC path end to end (rows a13-a16): engine-simulated instances ->
measure -> vertical_collapse -> merge -> build -> reorder, against the cutqc oracle stage by
stage and against the uncut statevector.  float64 on the device: 1e-10."""
import networkx as nx
import numpy as np
import pytest
import torch

from oracle import cutqc_oracle as co
from oracle import mis_oracle as mo
from oracle import statevector as sv
import cutqc_cases as cc

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _be(cuda, **kw):
    from qdc_b200.engine import B200Backend
    return B200Backend(device=cuda, **kw)


@pytest.mark.parametrize("name", list(cc.CASES))
def test_stages_match_oracle_and_uncut(cuda, name):
    from qdc_b200.cutqc.main import CutQC
    circ, frags, wpm = cc.CASES[name]()
    cpm = co.complete_path_map_from_wire_path_map(wpm)
    counter, kron_o, terms_o, K, inter = co.evaluate(frags, cpm, early_termination=1)
    job = CutQC.from_fragments(circ, frags, wpm, _be(cuda, tile_bits=3)).evaluate(early_termination=(0, 1))
    assert job.counter == counter and job.num_cuts == K
    for s in range(len(frags)):
        measured = job.measured[s].cpu().numpy()
        assert measured.shape[0] == len(inter["measured"][s])
        for idx, want in inter["measured"][s].items():
            assert np.abs(measured[idx] - want).max() < TOL
        kron = job.kron[1][s].cpu().numpy()
        present = job.kron_present[1][s].cpu().numpy()
        assert bool(job.kron_present[0][s].all())
        # "all zero" (vertical_collapse.c:64-72, |x| <= 1e-16) is decided on rounding noise for
        # terms that vanish analytically: a term the oracle keeps must be kept; one it drops
        # may survive here only as noise
        for k in range(kron.shape[0]):
            if k in kron_o[s]:
                assert np.abs(kron[k] - kron_o[s][k]).max() < TOL
                assert present[k] or np.abs(kron_o[s][k]).max() < 1e-12
            else:
                assert np.abs(kron[k]).max() < 1e-12
    n = circ.num_qubits
    full = sv.probabilities(sv.simulate(n, circ.to_instructions()))
    for et in (0, 1):
        layer, = job.post_process(early_termination=et)
        assert np.abs(layer.dense().cpu().numpy() - full).max() < TOL
        got = layer.to_dict()
        assert set(got) == {format(x, f"0{n}b") for x in range(1 << n) if full[x] > 1e-10}


@pytest.mark.parametrize("name,limit", [("chain3", 4), ("dqva6_fused", 3), ("there_and_back", 2)])
def test_dynamic_definition_layers_match_oracle(cuda, name, limit):
    from qdc_b200.cutqc.main import CutQC
    circ, frags, wpm = cc.CASES[name]()
    cpm = co.complete_path_map_from_wire_path_map(wpm)
    counter, kron_o, terms_o, K, _ = co.evaluate(frags, cpm)
    want = co.post_process(list(circ.qubits), frags, cpm, counter, kron_o, terms_o, K,
                           qubit_limit=limit, recursion_depth=4)
    job = CutQC.from_fragments(circ, frags, wpm, _be(cuda)).evaluate()
    got = job.post_process(qubit_limit=limit, recursion_depth=4)
    assert len(got) == len(want)
    for layer, (schedule, vec, dist) in zip(got, want):
        assert layer.schedule == schedule
        assert np.abs(layer.prob.cpu().numpy() - vec).max() < TOL
        d = layer.to_dict()
        assert d.keys() == dist.keys() and all(abs(d[k] - dist[k]) < TOL for k in dist)


def test_float32_output_and_fused_mis_score(cuda):
    from qdc_b200.cutqc.main import CutQC
    circ, frags, wpm, G = cc.dqva6(2, 5, True)
    job = CutQC.from_fragments(circ, frags, wpm, _be(cuda)).evaluate()
    layer, = job.post_process()
    n = circ.num_qubits
    full = sv.probabilities(sv.simulate(n, circ.to_instructions()))
    num_anc = n - G.number_of_nodes()
    res = layer.score(G, k=3, num_anc=num_anc).get()
    probs = {format(x, f"0{n}b"): full[x] for x in range(1 << n)}
    assert abs(res["sum_p"] - 1) < TOL
    assert abs(-res["sum_pw_indset"] - mo.avg_cost(mo.strip_ancillas(probs, num_anc), G)) < TOL
    layer32, = job.post_process(out_dtype=torch.float32)
    assert layer32.prob.dtype == torch.float32
    assert np.abs(layer32.dense().cpu().numpy() - full).max() < 2e-6


@pytest.mark.parametrize("name", list(cc.CASES))
def test_fused_evaluate_matches_oracle_kron_vectors(cuda, name):
    """fused=True: kron vectors straight from the basis-input amplitudes == vertical_collapse of
    the measured instances (oracle), and the reconstruction == uncut."""
    from qdc_b200.cutqc.main import CutQC
    circ, frags, wpm = cc.CASES[name]()
    cpm = co.complete_path_map_from_wire_path_map(wpm)
    _, kron_o, _, _, _ = co.evaluate(frags, cpm, early_termination=0)
    job = CutQC.from_fragments(circ, frags, wpm, _be(cuda)).evaluate(early_termination=(0, 1), fused=True)
    for s in range(len(frags)):
        kron = job.kron[0][s].cpu().numpy()
        assert kron.shape[0] == len(kron_o[s])
        for k, want in kron_o[s].items():
            assert np.abs(kron[k] - want).max() < TOL
    n = circ.num_qubits
    full = sv.probabilities(sv.simulate(n, circ.to_instructions()))
    for et in (0, 1):
        layer, = job.post_process(early_termination=et)
        assert np.abs(layer.dense().cpu().numpy() - full).max() < TOL


@pytest.mark.parametrize("name", ["chain3", "dqva6_fused"])
def test_ckt_wrappers_match_numpy_twin(cuda, name):
    """S5: CKT-named `_attribute_shots`, `build`, `generate_reconstructed_output` on the device ==
    the numpy twin (oracle) and the uncut distribution, from the same instance probabilities."""
    from qdc_b200.cutqc import ckt
    circ, frags, wpm = cc.CASES[name]()
    cpm = co.complete_path_map_from_wire_path_map(wpm)
    counter, _, terms, K, inter = co.evaluate(frags, cpm)
    entries = {s: {("label", k): (k, list(term)) for term, k in inter["kronecker_terms"][s].items()}
               for s in range(len(frags))}
    inst = {s: {i: v for i, v in inter["measured"][s].items()} for s in range(len(frags))}
    want_entries = co.ckt_attribute_shots(entries, inst)
    got_entries = ckt._attribute_shots(entries, inst)
    for s in want_entries:
        for k, v in want_entries[s].items():
            assert np.abs(got_entries[s][k].cpu().numpy() - v).max() < TOL
    want, order_w, over_w = co.ckt_build(terms, want_entries, K)
    got, order_g, over_g = ckt.build(terms, got_entries, K, num_threads=4)
    assert order_g == order_w and over_g == over_w
    assert np.abs(got.cpu().numpy() - want).max() < TOL
    n = circ.num_qubits
    full = sv.probabilities(sv.simulate(n, circ.to_instructions()))
    ordered = ckt.generate_reconstructed_output(circ, frags, got, order_g, cpm).cpu().numpy()
    assert np.abs(ordered - co.ckt_generate_reconstructed_output(list(circ.qubits), frags, want, order_w, cpm)).max() < TOL
    assert np.abs(ordered - full).max() < TOL


@pytest.mark.parametrize("name", ["ckt5", "there_and_back", "chain3", "dqva6_fused"])
def test_ckt_reconstruct_full_distribution(cuda, name):
    """CKT's two top-level calls (wire_cutting.py:88-187): `evaluate_subcircuits` instance by
    instance against the oracle's restatement of `_run_subcircuit_batch` + `measure_prob`, and
    `reconstruct_full_distribution` against the oracle's and against the uncut statevector - the
    assertion of the reference's own test (test_circuit_cutting.py:69-82, MSE ~ 0) on its own
    5-qubit circuit and on cases with 2 cuts / 3 subcircuits."""
    from qdc_b200.cutqc import ckt
    circ, frags, wpm = cc.CASES[name]()
    cpm = co.complete_path_map_from_wire_path_map(wpm)
    K = sum(len(p) - 1 for p in cpm.values())
    cuts = {"subcircuits": frags, "complete_path_map": cpm, "num_cuts": K}
    assert ckt._generate_metadata(cuts) == co.ckt_generate_summation_terms(frags, cpm, K)
    want_inst = co.ckt_run_subcircuit_instances(frags, co.ckt_generate_summation_terms(frags, cpm, K)[2])
    got_inst = ckt.evaluate_subcircuits(cuts, backend=_be(cuda))
    for s in want_inst:
        assert sorted(got_inst[s]) == sorted(want_inst[s])
        for i, v in want_inst[s].items():
            assert np.abs(got_inst[s][i].cpu().numpy() - v).max() < TOL
    got = ckt.reconstruct_full_distribution(circ, got_inst, cuts, num_threads=2).cpu().numpy()
    want = co.ckt_reconstruct_full_distribution(list(circ.qubits), want_inst, cuts)
    full = sv.probabilities(sv.simulate(circ.num_qubits, circ.to_instructions()))
    assert np.abs(got - want).max() < TOL
    assert np.mean((got - full) ** 2) < 1e-24 and np.abs(got - full).max() < TOL
    # numpy instance probabilities (what a CKT user holds) are accepted as well
    got_np = ckt.reconstruct_full_distribution(circ, want_inst, cuts).cpu().numpy()
    assert np.abs(got_np - full).max() < TOL


def test_dynamic_definition_state_on_disk_and_resume(cuda, tmp_path):
    """f4: `post_process(dest_folder=)` writes the reference's recursion state (meta_data.pckl,
    dynamic_definition_<layer>/build_output.pckl with its three keys, distributor.py:121-141,
    main.py:452-457); stopping after one layer and resuming from disk gives the same layers as one
    in-memory run, and the zoomed_ctr of a zoomed layer is advanced on disk as the reference does."""
    from qdc_b200.cutqc.main import CutQC
    from qdc_b200.cutqc.helper_fun import get_dirname, read_dict
    circ, frags, wpm = cc.CASES["chain3"]()
    job = CutQC.from_fragments(circ, frags, wpm, _be(cuda)).evaluate()
    want = job.post_process(qubit_limit=2, recursion_depth=4)
    assert len(want) >= 3
    dest = str(tmp_path / get_dirname("chain3", 3, 0, "sv", 1, 2, "build"))
    assert dest.endswith("cutqc_data/chain3/cc_3/sv_0_2_1")
    first = job.post_process(qubit_limit=2, recursion_depth=1, dest_folder=dest)
    assert len(first) == 1
    meta = read_dict(f"{dest}/meta_data.pckl")
    assert set(meta) == {"counter", "dynamic_definition_schedule"} and list(meta["dynamic_definition_schedule"]) == [0]
    out0 = read_dict(f"{dest}/dynamic_definition_0/build_output.pckl")
    assert set(out0) == {"zoomed_ctr", "max_states", "reconstructed_prob"} and out0["zoomed_ctr"] == 0
    prob0 = want[0].prob.cpu().numpy()
    assert np.array_equal(out0["reconstructed_prob"], prob0)
    assert out0["max_states"] == sorted(range(len(prob0)), key=lambda x: prob0[x], reverse=True)   # main.py:455
    got = job.post_process(qubit_limit=2, recursion_depth=4, dest_folder=dest, resume=True)
    assert len(got) == len(want)
    for a, b in zip(got, want):
        assert a.schedule == b.schedule and torch.equal(a.prob, b.prob)
    meta = read_dict(f"{dest}/meta_data.pckl")
    assert list(meta["dynamic_definition_schedule"]) == list(range(len(want)))
    assert sum(read_dict(f"{dest}/dynamic_definition_{l}/build_output.pckl")["zoomed_ctr"]
               for l in range(len(want))) == len(want) - 1
    with pytest.raises(Exception):
        job.post_process(resume=True)
