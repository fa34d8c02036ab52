"""ORACLE (test infrastructure, never on the product path): CPU restatement of the reference's
CutQC host pipeline in float64 numpy, label by label as the reference does it, so the array /
bit-mask formulation of the product (`qdc_b200.cutqc.*`) has an independent checker.

Reference lines followed (paths under /root/reference/cutqc):
  evaluator.py:20-32   find_subcircuit_O_rho_qubits      evaluator.py:34-67  find_all_combinations
  evaluator.py:69-80   mutate_measurement_basis          evaluator.py:82-124 get_subcircuit_instance
  measure.c:123-161    measure_instance                  cutter.py:439-461   get_pairs / get_counter
  post_process.py:3-50 find_init_meas                    post_process.py:52-63 get_combinations
  post_process.py:65-129 build (kronecker / summation terms)
  vertical_collapse.c:27-104                             distributor.py:15-95 DD schedule
  merge.c:28-122       merge                             build.c:69-196      build
  main.py:141-196      reorder

Pinning: the stage arithmetic (measure / vertical_collapse / merge / build) is checked against
the reference's own C programs compiled into oracle/_ref (tests/test_oracle_cutqc.py); the
Python host logic has no golden vectors in the reference, so it is pinned through the
invariant every reference scenario reduces to: reconstructed distribution == distribution of
the uncut circuit.
"""
from __future__ import annotations

import copy
import itertools

import numpy as np

from . import statevector as sv

_INIT_GATES = {           # evaluator.py:89-107 -- gates in the order they ACT (front insertions
    "zero": [],           # are listed back to front in the reference)
    "one": ["x"],
    "plus": ["h"],
    "minus": ["x", "h"],
    "plusI": ["h", "s"],
    "minusI": ["x", "h", "s"],
}
_MEAS_GATES = {"I": [], "comp": [], "Z": [], "X": ["h"], "Y": ["sdg", "h"]}   # :109-119


def complete_path_map_from_wire_path_map(wire_path_map):
    """qsplit's `{wire: ((frag, frag_wire), ...)}` in cutqc's form
    `{wire: [{'subcircuit_idx', 'subcircuit_qubit'}, ...]}` (cutter.py:299-386 output layout)."""
    return {wire: [{"subcircuit_idx": f, "subcircuit_qubit": q} for f, q in path]
            for wire, path in wire_path_map.items()}


def find_subcircuit_O_rho_qubits(complete_path_map, subcircuit_idx):           # evaluator.py:20-32
    O_qubits, rho_qubits = [], []
    for input_qubit in complete_path_map:
        path = complete_path_map[input_qubit]
        if len(path) > 1:
            O_qubits += [q for q in path[:-1] if q["subcircuit_idx"] == subcircuit_idx]
            rho_qubits += [q for q in path[1:] if q["subcircuit_idx"] == subcircuit_idx]
    return O_qubits, rho_qubits


def mutate_measurement_basis(meas):                                            # evaluator.py:69-80
    if all(x != "I" for x in meas):
        return [tuple(meas)]
    return list(itertools.product(*[["I", "Z"] if x == "I" else [x] for x in meas]))


def find_all_combinations(O_qubits, rho_qubits, qubits):                       # evaluator.py:34-67
    complete_inits = []
    for init in itertools.product(["zero", "one", "plus", "plusI"], repeat=len(rho_qubits)):
        full = ["zero"] * len(qubits)
        for lab, q in zip(init, rho_qubits):
            full[qubits.index(q["subcircuit_qubit"])] = lab
        complete_inits.append(full)
    complete_meas = []
    for meas in itertools.product(["I", "X", "Y"], repeat=len(O_qubits)):
        full = ["comp"] * len(qubits)
        for lab, q in zip(meas, O_qubits):
            full[qubits.index(q["subcircuit_qubit"])] = lab
        complete_meas.append(full)
    combinations = list(itertools.product(complete_inits, complete_meas))
    indexed, ctr = {}, 0
    for inits, meas in combinations:
        for mutated in mutate_measurement_basis(meas):
            indexed[(tuple(inits), tuple(mutated))] = ctr
            ctr += 1
    return combinations, indexed


def instance_probabilities(subcircuit, inits, meas):            # evaluator.py:82-124, 126-129
    """Statevector probabilities of one subcircuit instance, index = qiskit little-endian."""
    n = len(subcircuit.qubits)
    front, back = [], []
    for q, lab in enumerate(inits):
        if lab not in _INIT_GATES:
            raise Exception("Illegal initialization : ", lab)
        front += [(g, (q,), (), None) for g in _INIT_GATES[lab]]
    for q, lab in enumerate(meas):
        if lab not in _MEAS_GATES:
            raise Exception("Illegal measurement basis:", lab)
        back += [(g, (q,), (), None) for g in _MEAS_GATES[lab]]
    state = sv.simulate(n, front + list(subcircuit.to_instructions()) + back)
    return sv.probabilities(state)


def measure_instance(raw, meas):                                               # measure.c:123-161
    """Signed marginal over the qubits whose label is not 'comp'; sigma flips for every set bit
    of a qubit measured in X / Y / Z ('I' sums unsigned).  Effective index keeps the 'comp'
    qubits in ascending qubit order (measure.c:163-214)."""
    d = len(meas)
    comp = [q for q in range(d) if meas[q] == "comp"]
    out = np.zeros(1 << len(comp))
    for full in range(1 << d):
        sigma = 1
        for q in range(d):
            if (full >> q) & 1 and meas[q] not in ("I", "comp"):
                sigma = -sigma
        eff = sum(((full >> q) & 1) << j for j, q in enumerate(comp))
        out[eff] += sigma * raw[full]
    return out


def get_pairs(complete_path_map):                                              # cutter.py:439-448
    pairs = []
    for input_qubit in complete_path_map:
        path = complete_path_map[input_qubit]
        for a, b in zip(path[:-1], path[1:]):
            pairs.append((a, b))
    return pairs


def get_counter(subcircuits, O_rho_pairs):                                     # cutter.py:450-461
    counter = {s: {"effective": len(c.qubits), "rho": 0, "O": 0, "d": len(c.qubits)}
               for s, c in enumerate(subcircuits)}
    for O_qubit, rho_qubit in O_rho_pairs:
        counter[O_qubit["subcircuit_idx"]]["effective"] -= 1
        counter[O_qubit["subcircuit_idx"]]["O"] += 1
        counter[rho_qubit["subcircuit_idx"]]["rho"] += 1
    return counter


def get_combinations(complete_path_map):                                       # post_process.py:52-63
    pairs = get_pairs(complete_path_map)
    return pairs, list(itertools.product(["I", "X", "Y", "Z"], repeat=len(pairs)))


_INIT_EXPANSION = {"zero": ["zero"], "I": ["+zero", "+one"], "X": ["2plus", "-zero", "-one"],
                   "Y": ["2plusI", "-zero", "-one"], "Z": ["+zero", "-one"]}   # post_process.py:20-32
_INIT_TERM = {"zero": (1, "zero"), "+zero": (1, "zero"), "+one": (1, "one"), "2plus": (2, "plus"),
              "-zero": (-1, "zero"), "-one": (-1, "one"), "2plusI": (2, "plusI")}   # :79-99


def find_init_meas(combination, O_rho_pairs, subcircuits):                     # post_process.py:3-50
    table = {s: [["zero"] * len(c.qubits), ["comp"] * len(c.qubits)]
             for s, c in enumerate(subcircuits)}
    for s, (O_qubit, rho_qubit) in zip(combination, O_rho_pairs):
        rq = subcircuits[rho_qubit["subcircuit_idx"]].qubits
        oq = subcircuits[O_qubit["subcircuit_idx"]].qubits
        table[rho_qubit["subcircuit_idx"]][0][rq.index(rho_qubit["subcircuit_qubit"])] = s
        table[O_qubit["subcircuit_idx"]][1][oq.index(O_qubit["subcircuit_qubit"])] = s
    out = {}
    for s, (init, meas) in table.items():
        inits = itertools.product(*[_INIT_EXPANSION[x] for x in init])
        meas = tuple(x if x == "comp" else "+" + x for x in meas)
        out[s] = [(tuple(i), meas) for i in inits]
    return out


def build_terms(combinations, O_rho_pairs, subcircuits, all_indexed_combinations):   # post_process.py:65-129
    kronecker_terms = {s: {} for s in range(len(subcircuits))}
    summation_terms = []
    for combination in combinations:
        summation_term = {}
        all_init_meas = find_init_meas(combination, O_rho_pairs, subcircuits)
        for s in range(len(subcircuits)):
            term = []
            for init, meas in all_init_meas[s]:
                coefficient = 1
                labels = []
                for x in init:
                    c, lab = _INIT_TERM[x]
                    coefficient *= c
                    labels.append(lab)
                meas = tuple(x if x == "comp" else x[1:] for x in meas)
                term.append((coefficient, all_indexed_combinations[s][(tuple(labels), meas)]))
            term = tuple(term)
            if term not in kronecker_terms[s]:
                kronecker_terms[s][term] = len(kronecker_terms[s])
            summation_term[s] = kronecker_terms[s][term]
        summation_terms.append(summation_term)
    return kronecker_terms, summation_terms


def vertical_collapse(kronecker_terms, measured, early_termination=0):   # vertical_collapse.c:27-104
    """measured[instance index] -> {kron index: vector}; with early_termination=1 all-zero
    vectors are not written (:76-78)."""
    out = {}
    for term, kron_index in kronecker_terms.items():
        vec = None
        for coefficient, inst in term:
            vec = coefficient * measured[inst] if vec is None else vec + coefficient * measured[inst]
        if early_termination == 1 and not np.any(np.abs(vec) > 1e-16):      # :64-72
            continue
        out[kron_index] = vec
    return out


# ---------------------------------------------------------------------------------------------
# dynamic-definition schedule (distributor.py:15-95)
# ---------------------------------------------------------------------------------------------
def distribute_load(total_load, capacities):                                   # distributor.py:15-27
    assert total_load <= sum(capacities)
    loads = [int(c / sum(capacities) * total_load) for c in capacities]
    total_load -= sum(loads)
    for slot in reversed(range(len(loads))):
        while total_load > 0 and loads[slot] < capacities[slot]:
            loads[slot] += 1
            total_load -= 1
    assert total_load == 0
    return loads


def initialize_dynamic_definition_schedule(counter, recursion_qubit):          # distributor.py:29-60
    smart_order = sorted(counter.keys(), key=lambda s: counter[s]["effective"])
    capacities = [counter[s]["effective"] for s in smart_order]
    active = distribute_load(min(sum(capacities), recursion_qubit), capacities)
    state = {s: ["active"] * a + ["merged"] * (counter[s]["effective"] - a)
             for s, a in zip(smart_order, active)}
    return {"smart_order": smart_order, "subcircuit_state": state, "upper_bin": None}


def next_dynamic_definition_schedule(recursion_layer, schedule, state_idx, recursion_qubit):   # :62-95
    schedule = copy.deepcopy(schedule)
    num_active = sum(st.count("active") for st in schedule["subcircuit_state"].values())
    bits = bin(state_idx)[2:].zfill(num_active)
    ptr = 0
    for s in schedule["smart_order"]:
        st = schedule["subcircuit_state"][s]
        for i, x in enumerate(st):
            if x == "active":
                st[i] = bits[ptr]
                ptr += 1
    schedule["smart_order"] = sorted(schedule["smart_order"],
                                     key=lambda s: schedule["subcircuit_state"][s].count("merged"))
    capacities = [schedule["subcircuit_state"][s].count("merged") for s in schedule["smart_order"]]
    active = distribute_load(min(sum(capacities), recursion_qubit), capacities)
    for s, a in zip(schedule["smart_order"], active):
        st = schedule["subcircuit_state"][s]
        for i, x in enumerate(st):
            if x == "merged" and a > 0:
                st[i] = "active"
                a -= 1
    schedule["upper_bin"] = (recursion_layer, state_idx)
    return schedule


def merge(vec, qubit_states):                                                  # merge.c:28-122
    """qubit_states[i] describes bit num_effective-1-i (decToBinary is MSB first)."""
    n = len(qubit_states)
    num_active = qubit_states.count("active")
    out = np.zeros(1 << num_active)
    for state in range(1 << n):
        bits = [(state >> (n - 1 - i)) & 1 for i in range(n)]
        dest, seen = 0, 0
        for i, x in enumerate(qubit_states):
            if x == "merged":
                continue
            if x == "active":
                dest += bits[i] << (num_active - 1 - seen)
                seen += 1
            elif int(x) != bits[i]:
                dest = -1
                break
        if dest >= 0:
            out[dest] += vec[state]
    return out


def build(kron, summation_terms, smart_order, num_cuts):                       # build.c:35-196
    """kron[s] = {kron index: merged vector}; a term with a missing vector is skipped."""
    total = None
    for term in summation_terms:
        if any(term[s] not in kron[s] for s in smart_order):
            continue
        vec = kron[smart_order[0]][term[smart_order[0]]]
        for s in smart_order[1:]:
            vec = np.outer(vec, kron[s][term[s]]).reshape(-1)
        total = vec.copy() if total is None else total + vec
    if total is None:
        length = 1
        for s in smart_order:
            length *= len(next(iter(kron[s].values()))) if kron[s] else 1
        total = np.zeros(length)
    return total * 0.5 ** num_cuts


def reorder(full_qubits, unordered, complete_path_map, subcircuits, layer_schedule,
            threshold=1e-10):                                                  # main.py:141-196
    out_qubits = {}
    for input_qubit in complete_path_map:
        last = complete_path_map[input_qubit][-1]
        out_qubits.setdefault(last["subcircuit_idx"], []).append(
            (last["subcircuit_qubit"], full_qubits.index(input_qubit)))
    for s in out_qubits:
        ordered = sorted(out_qubits[s], key=lambda x: subcircuits[s].qubits.index(x[0]), reverse=True)
        out_qubits[s] = [x[1] for x in ordered]
    fmt, state = [], []
    for s in layer_schedule["smart_order"]:
        if s in out_qubits:
            fmt += out_qubits[s]
            state += layer_schedule["subcircuit_state"][s]
    num_active, num_merged = state.count("active"), state.count("merged")
    labelled = {}
    for ctr, p in enumerate(unordered):
        avg = p / 2 ** num_merged
        for merged_state in itertools.product(["0", "1"], repeat=num_merged):
            act = bin(ctr)[2:].zfill(num_active)
            merged_state = list(merged_state)
            full = []
            for x in state:
                if x == "active":
                    full.append(act[0])
                    act = act[1:]
                elif x == "merged":
                    full.append(merged_state.pop(0))
                else:
                    full.append(x)
            _, full = zip(*sorted(zip(fmt, full), reverse=True))
            labelled[int("".join(full), 2)] = avg
    return {"{:0{}b}".format(k, len(full_qubits)): v for k, v in labelled.items() if v > threshold}


# ---------------------------------------------------------------------------------------------
# whole pipeline (main.py: evaluate :72-84 + post_process :86-139), in memory
# ---------------------------------------------------------------------------------------------
def evaluate(subcircuits, complete_path_map, early_termination=0):
    """-> (counter, kron {s: {kron index: vector}}, summation_terms, num_cuts, intermediates)."""
    O_rho_pairs, combinations = get_combinations(complete_path_map)
    counter = get_counter(subcircuits, O_rho_pairs)
    indexed, measured = {}, {}
    for s, sub in enumerate(subcircuits):
        O_q, rho_q = find_subcircuit_O_rho_qubits(complete_path_map, s)
        combos, indexed[s] = find_all_combinations(O_q, rho_q, sub.qubits)
        measured[s] = {}
        for inits, meas in combos:
            raw = instance_probabilities(sub, inits, meas)
            for mutated in mutate_measurement_basis(meas):                    # evaluator.py:138-140
                measured[s][indexed[s][(tuple(inits), tuple(mutated))]] = measure_instance(raw, mutated)
    kronecker_terms, summation_terms = build_terms(combinations, O_rho_pairs, subcircuits, indexed)
    kron = {s: vertical_collapse(kronecker_terms[s], measured[s], early_termination)
            for s in range(len(subcircuits))}
    inter = {"indexed": indexed, "measured": measured, "kronecker_terms": kronecker_terms}
    return counter, kron, summation_terms, len(O_rho_pairs), inter


def post_process(full_qubits, subcircuits, complete_path_map, counter, kron, summation_terms,
                 num_cuts, qubit_limit=None, recursion_depth=1):
    """-> list of (schedule, reconstructed vector, probability dict) per recursion layer.  The
    zoom target of layer L+1 is the largest bin of the most probable unexplored layer
    (distributor.py:97-117, main.py:452-456)."""
    if qubit_limit is None:
        qubit_limit = sum(c["effective"] for c in counter.values())
    layers, outputs = [], []
    for layer in range(recursion_depth):
        if layer == 0:
            schedule = initialize_dynamic_definition_schedule(counter, qubit_limit)
        else:
            best, best_p = -1, 0.0
            for l, out in enumerate(outputs):
                merged = sum(st.count("merged") for st in layers[l]["subcircuit_state"].values())
                if merged == 0 or out["zoomed_ctr"] == len(out["max_states"]):
                    continue
                p = out["reconstructed_prob"][out["max_states"][out["zoomed_ctr"]]]
                if p > best_p and p > 1e-16:
                    best, best_p = l, p
            if best < 0:
                break
            out = outputs[best]
            schedule = next_dynamic_definition_schedule(best, layers[best],
                                                        out["max_states"][out["zoomed_ctr"]], qubit_limit)
            out["zoomed_ctr"] += 1
        merged = {s: {k: merge(v, schedule["subcircuit_state"][s]) for k, v in kron[s].items()}
                  for s in kron}
        prob = build(merged, summation_terms, schedule["smart_order"], num_cuts)
        order = sorted(range(len(prob)), key=lambda x: prob[x], reverse=True)
        layers.append(schedule)
        outputs.append({"zoomed_ctr": 0, "max_states": order, "reconstructed_prob": prob})
    return [(layers[l], outputs[l]["reconstructed_prob"],
             reorder(full_qubits, outputs[l]["reconstructed_prob"], complete_path_map, subcircuits, layers[l]))
            for l in range(len(layers))]


# ---------------------------------------------------------------------------------------------
# the CKT numpy twin (circuit-knitting-toolbox/.../wire_cutting*.py), float64
# ---------------------------------------------------------------------------------------------
def ckt_attribute_shots(subcircuit_entries, subcircuit_instance_probs):        # wire_cutting.py:252-310
    out = {}
    for s in subcircuit_entries:
        out[s] = {}
        for label in subcircuit_entries[s]:
            entry_idx, kronecker_term = subcircuit_entries[s][label]
            prob = None
            for coefficient, inst in kronecker_term:
                term = coefficient * subcircuit_instance_probs[s][inst]
                prob = term if prob is None else prob + term
            out[s][entry_idx] = prob
    return out


def ckt_build(summation_terms, subcircuit_entry_probs, num_cuts):   # wire_cutting_post_processing.py:340-443
    smart_order = sorted(subcircuit_entry_probs.keys(), key=lambda s: len(subcircuit_entry_probs[s][0]))
    overhead = {"additions": 0, "multiplications": 0}
    total = None
    for term in summation_terms:
        vec = None
        for s in smart_order:
            entry = subcircuit_entry_probs[s][term[s]]
            if vec is None:
                vec = entry
            else:
                vec = np.kron(vec, entry)
                overhead["multiplications"] += len(vec)
        if total is None:
            total = vec.copy()
        else:
            total += vec
            overhead["additions"] += len(total)
    return total / 2 ** num_cuts, smart_order, overhead


def ckt_generate_reconstructed_output(full_qubits, subcircuits, unordered, smart_order,
                                      complete_path_map):          # wire_cutting_verification.py:80-136
    out_qubits = {s: [] for s in smart_order}
    for input_qubit in complete_path_map:
        last = complete_path_map[input_qubit][-1]
        out_qubits[last["subcircuit_idx"]].append((last["subcircuit_qubit"], full_qubits.index(input_qubit)))
    for s in out_qubits:
        ordered = sorted(out_qubits[s], key=lambda x: subcircuits[s].qubits.index(x[0]), reverse=True)
        out_qubits[s] = [x[1] for x in ordered]
    unordered_qubit = []
    for s in smart_order:
        unordered_qubit += out_qubits[s]
    n = len(full_qubits)
    out = np.zeros(len(unordered))
    for state, p in enumerate(unordered):
        bits = bin(state)[2:].zfill(n)
        _, ordered = zip(*sorted(zip(unordered_qubit, bits), reverse=True))
        out[int("".join(ordered), 2)] = p
    return out


_CKT_INIT_EXPANSION = {"zero": ["zero"], "I": ["+zero", "+one"], "X": ["2plus", "-zero", "-one"],
                       "Y": ["2plusI", "-zero", "-one"], "Z": ["+zero", "-one"]}
_CKT_PHYSICAL = {"zero": (1, "zero"), "+zero": (1, "zero"), "+one": (1, "one"), "2plus": (2, "plus"),
                 "-zero": (-1, "zero"), "-one": (-1, "one"), "2plusI": (2, "plusI")}


def ckt_generate_summation_terms(subcircuits, complete_path_map, num_cuts):   # wire_cutting_post_processing.py:225-337
    """(summation_terms, subcircuit_entries, subcircuit_instances) with the CKT numbering: entries
    and instances are numbered in order of first appearance while the 4^num_cuts labels are walked
    with cut 0 as the LEAST significant base-4 digit (get_label :48-68)."""
    pairs = get_pairs(complete_path_map)                                       # :23-46, same walk
    n_sub = len(subcircuits)
    terms, entries, instances = [], {s: {} for s in range(n_sub)}, {s: {} for s in range(n_sub)}
    for idx in range(4 ** num_cuts):
        label = ["IXYZ"[(idx >> (2 * c)) & 3] for c in range(num_cuts)]
        sub_label = {s: {"init": "", "meas": ""} for s in range(n_sub)}        # attribute_label :71-105
        for c, (O_q, rho_q) in zip(label, pairs):
            sub_label[O_q["subcircuit_idx"]]["meas"] += c
            sub_label[rho_q["subcircuit_idx"]]["init"] += c
        term = {}
        for s in range(n_sub):
            key = (sub_label[s]["init"], sub_label[s]["meas"])
            if key not in entries[s]:
                qubits = list(subcircuits[s].qubits)
                init, meas = ["zero"] * len(qubits), ["comp"] * len(qubits)    # fill_label :108-146
                i_ctr = m_ctr = 0
                for O_q, rho_q in pairs:
                    if O_q["subcircuit_idx"] == s:
                        meas[qubits.index(O_q["subcircuit_qubit"])] = sub_label[s]["meas"][m_ctr]
                        m_ctr += 1
                    if rho_q["subcircuit_idx"] == s:
                        init[qubits.index(rho_q["subcircuit_qubit"])] = sub_label[s]["init"][i_ctr]
                        i_ctr += 1
                kron_term = []
                for combo in itertools.product(*[_CKT_INIT_EXPANSION[x] for x in init]):   # get_init_meas :149-189
                    coeff = 1
                    phys = []
                    for x in combo:                                            # convert_to_physical_init :192-222
                        c, name = _CKT_PHYSICAL[x]
                        coeff *= c
                        phys.append(name)
                    inst_key = (tuple(phys), tuple(meas))
                    if inst_key not in instances[s]:
                        instances[s][inst_key] = len(instances[s])
                    kron_term.append((coeff, instances[s][inst_key]))
                entries[s][key] = (len(entries[s]), kron_term)
            term[s] = entries[s][key][0]
        terms.append(term)
    return terms, entries, instances


def ckt_run_subcircuit_instances(subcircuits, subcircuit_instances):   # wire_cutting_evaluation.py:299-380
    """{subcircuit: {instance index: measured probabilities}}: one exact statevector simulation per
    unmutated instance (`I` and `Z` share the circuit, :100-124), then `measure_prob` (:249-296)."""
    out = {}
    for s, sub in enumerate(subcircuits):
        out[s] = {}
        for (init, meas), idx in subcircuit_instances[s].items():
            if idx in out[s]:
                continue
            raw = instance_probabilities(sub, init, meas)                      # raises on a bare 'Z', as :190-193
            for mutated in mutate_measurement_basis(tuple(meas)):
                out[s][subcircuit_instances[s][(init, tuple(mutated))]] = measure_instance(raw, mutated)
    return out


def ckt_reconstruct_full_distribution(full_qubits, subcircuit_instance_probs, cuts):   # wire_cutting.py:149-187
    terms, entries, _ = ckt_generate_summation_terms(cuts["subcircuits"], cuts["complete_path_map"],
                                                     cuts["num_cuts"])
    entry_probs = ckt_attribute_shots(entries, subcircuit_instance_probs)
    unordered, smart_order, _ = ckt_build(terms, entry_probs, cuts["num_cuts"])
    return ckt_generate_reconstructed_output(full_qubits, cuts["subcircuits"], unordered, smart_order,
                                             cuts["complete_path_map"])
