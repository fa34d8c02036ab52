"""CPU tests: the product's array formulation of the CutQC host logic (cutqc/evaluator.py,
post_process.py, distributor.py, main.py bookkeeping) against the label-by-label oracle, and the
oracle pipeline against the uncut statevector (the invariant that pins it)."""
import itertools
import random

import numpy as np
import pytest
import torch

from oracle import cutqc_oracle as co
from oracle import statevector as sv
import cutqc_cases as cc
from qdc_b200.cutqc import distributor as dist
from qdc_b200.cutqc import evaluator as ev
from qdc_b200.cutqc import post_process as pp
from qdc_b200.cutqc import main as cm


@pytest.fixture(scope="module", params=list(cc.CASES))
def case(request):
    circ, frags, wpm = cc.CASES[request.param]()
    return circ, frags, wpm, co.complete_path_map_from_wire_path_map(wpm)


def test_oracle_pipeline_reconstructs_uncut_distribution(case):
    circ, frags, _, cpm = case
    for et in (0, 1):
        counter, kron, terms, K, _ = co.evaluate(frags, cpm, early_termination=et)
        (_, vec, got), = co.post_process(list(circ.qubits), frags, cpm, counter, kron, terms, K)
        n = circ.num_qubits
        full = sv.probabilities(sv.simulate(n, circ.to_instructions()))
        assert abs(vec.sum() - 1) < 1e-12
        for x in range(1 << n):
            assert abs(got.get(format(x, f"0{n}b"), 0.0) - full[x]) < 1e-10 + (full[x] <= 1e-10) * 1e-10


def test_instance_numbering_matches_oracle(case):
    _, frags, _, cpm = case
    assert cm.complete_path_map_from_wire_path_map(case[2]) == cpm
    for s, sub in enumerate(frags):
        O_q, rho_q = co.find_subcircuit_O_rho_qubits(cpm, s)
        assert ev.find_subcircuit_O_rho_qubits(cpm, s) == (O_q, rho_q)
        combos_o, indexed_o = co.find_all_combinations(O_q, rho_q, list(sub.qubits))
        combos_p, indexed_p = ev.find_all_combinations(O_q, rho_q, list(sub.qubits))
        assert [(list(i), list(m)) for i, m in combos_p] == [(list(i), list(m)) for i, m in combos_o]
        assert indexed_p == indexed_o
        tab = ev._table_for(sub, O_q, rho_q)
        # parent / sign mask of every mutated instance
        for (inits, meas), k in indexed_o.items():
            assert tab.key(k) == (inits, meas)
            mut = k % tab.n_mut
            parent_meas = tuple("I" if x == "Z" else x for x in meas)
            want_parent = [tuple(m) for _, m in combos_o[:tab.n_meas3]].index(parent_meas)
            assert tab.parent[mut] == want_parent
            assert tab.sign_mask[mut] == sum(1 << q for q, x in enumerate(meas) if x in "XYZ")
    assert ev.mutate_measurement_basis(("X", "comp")) == [("X", "comp")]
    assert ev.mutate_measurement_basis(("I", "comp", "I")) == co.mutate_measurement_basis(("I", "comp", "I"))


def test_explicit_instances_match_oracle_probabilities(case):
    """get_subcircuit_instance circuits (evaluator.py:82-124) give the oracle's raw vectors."""
    _, frags, _, cpm = case
    circ_dict, indexed = ev.generate_subcircuit_instances(frags, cpm)
    keys = list(circ_dict)
    random.Random(0).shuffle(keys)
    for s, inits, meas in keys[:12]:
        inst = circ_dict[(s, inits, meas)]
        got = sv.probabilities(sv.simulate(inst["circuit"].num_qubits, inst["circuit"].to_instructions()))
        assert np.abs(got - co.instance_probabilities(frags[s], inits, meas)).max() < 1e-12
        assert inst["shots"] == 8192


def test_term_tables_match_oracle(case):
    circ, frags, _, cpm = case
    pairs, combos = co.get_combinations(cpm)
    indexed = {s: co.find_all_combinations(*co.find_subcircuit_O_rho_qubits(cpm, s), list(f.qubits))[1]
               for s, f in enumerate(frags)}
    kron_o, sum_o = co.build_terms(combos, pairs, frags, indexed)
    tt = pp.term_tables(frags, cpm)
    assert tt.num_cuts == len(pairs)
    for s in range(len(frags)):
        by_index = {k: term for term, k in kron_o[s].items()}
        assert len(by_index) == len(tt.kron_terms[s])
        for k, term in enumerate(tt.kron_terms[s]):
            assert tuple(term) == by_index[k]                  # same pairs, same order
        assert [row[s] for row in sum_o] == tt.term_index[s].tolist()
    # reference-layout views
    assert pp.get_combinations(cpm) == (pairs, combos)
    for combo in combos[:: max(1, len(combos) // 5)]:
        assert pp.find_init_meas(combo, pairs, frags) == co.find_init_meas(combo, pairs, frags)
    kron_p, sum_p = pp.build(circ, combos, pairs, frags, indexed)
    assert kron_p == kron_o and sum_p == sum_o
    assert cm.get_counter(frags, pairs) == co.get_counter(frags, pairs)


def test_schedule_matches_oracle():
    rng = random.Random(3)
    for _ in range(200):
        counter = {s: {"effective": rng.randint(0, 6)} for s in range(rng.randint(1, 4))}
        counter[0]["effective"] += 1                      # the reference divides by the total
        limit = rng.randint(1, 10)
        want = co.initialize_dynamic_definition_schedule(counter, limit)
        got = dist.initialize_dynamic_definition_schedule(counter, limit)
        assert got[0] == want
        layer, cur_o, cur_p = 0, want, got[0]
        for _ in range(3):
            active = sum(st.count("active") for st in cur_o["subcircuit_state"].values())
            if sum(st.count("merged") for st in cur_o["subcircuit_state"].values()) == 0:
                break                                     # nothing left to zoom into (distributor.py:107)
            state = rng.randrange(1 << active)
            cur_o = co.next_dynamic_definition_schedule(layer, cur_o, state, limit)
            cur_p = dist.next_dynamic_definition_schedule(layer, cur_p, state, limit)
            assert cur_p == cur_o
            layer += 1
    for total, caps in [(5, [3, 3]), (0, [2, 2]), (7, [1, 2, 4]), (4, [0, 4, 0])]:
        assert dist.distribute_load(total, caps) == co.distribute_load(total, caps)
    assert dist.find_process_jobs(list(range(10)), 1, 4) == [3, 4, 5]
    assert dist.find_process_jobs(list(range(10)), 3, 4) == [8, 9]


@pytest.mark.parametrize("limit,depth", [(None, 1), (3, 3)])
def test_layer_bookkeeping_matches_oracle_reorder(case, limit, depth):
    """ReconstructedLayer.to_dict / dense on the oracle's bin vectors == oracle reorder."""
    circ, frags, _, cpm = case
    counter, kron, terms, K, _ = co.evaluate(frags, cpm)
    layers = co.post_process(list(circ.qubits), frags, cpm, counter, kron, terms, K,
                             qubit_limit=limit, recursion_depth=depth)
    obj = cm.CutQC.__new__(cm.CutQC)
    obj.circuit, obj.subcircuits, obj.complete_path_map = circ, list(frags), cpm
    n = circ.num_qubits
    for schedule, vec, want in layers:
        layer = obj._wrap(torch.as_tensor(vec), schedule)
        got = layer.to_dict()
        assert got.keys() == want.keys()
        assert all(abs(got[k] - want[k]) < 1e-15 for k in want)
        assert obj.reorder(circ, vec, cpm, frags, schedule) == got
        dense = layer.dense().numpy()
        if not layer.fixed:
            assert abs(dense.sum() - vec.sum()) < 1e-12
        for k, v in want.items():
            assert abs(dense[int(k, 2)] - v) < 1e-15
