"""Dynamic-definition schedule of the CutQC reconstruction (reference: cutqc/distributor.py).

A schedule says, per subcircuit, what happens to each effective (output) qubit in one
reconstruction pass -- listed most significant first, as `merge.c` reads them:
    'active'  kept as an index bit of the reconstructed vector
    'merged'  summed out (one bin for both values)
    '0' / '1' fixed by an earlier "zoom" into the most probable bin
and the order (`smart_order`) in which the subcircuit vectors are kron-ed (build.c:69-196).
Same rules and tie-breaks as the reference; plus the `merge_%d.txt` / `build_%d.txt` writers of
the file protocol (distributor.py:143-187) for the stage executables in `cutqc/bin/`.
"""
from __future__ import annotations

import copy
import glob
import os
from typing import Dict, List, Optional, Sequence

__all__ = ["distribute_load", "initialize_dynamic_definition_schedule",
           "next_dynamic_definition_schedule", "find_max_recursion_layer", "merge_states",
           "write_files", "find_process_jobs"]


def distribute_load(total_load: int, capacities: Sequence[int]) -> List[int]:
    """Proportional share rounded down, the remainder handed out from the LAST slot backwards
    until each is full (distributor.py:15-27)."""
    cap_sum = sum(capacities)
    assert total_load <= cap_sum
    loads = [int(c / cap_sum * total_load) if cap_sum else 0 for c in capacities]
    left = total_load - sum(loads)
    for slot in range(len(loads) - 1, -1, -1):
        give = min(left, capacities[slot] - loads[slot])
        loads[slot] += give
        left -= give
    assert left == 0
    return loads


def _share(capacities, recursion_qubit):
    return distribute_load(min(sum(capacities), recursion_qubit), capacities)


def initialize_dynamic_definition_schedule(counter: Dict[int, dict], recursion_qubit: int,
                                           verbose: int = 0) -> Dict[int, dict]:
    """Layer 0 (distributor.py:29-60): subcircuits ordered by effective qubits (stable),
    the leading qubits of each active, the rest merged."""
    order = sorted(counter, key=lambda s: counter[s]["effective"])
    active = _share([counter[s]["effective"] for s in order], recursion_qubit)
    layer = {"smart_order": order,
             "subcircuit_state": {s: ["active"] * a + ["merged"] * (counter[s]["effective"] - a)
                                  for s, a in zip(order, active)},
             "upper_bin": None}
    if verbose > 0:
        for k, v in layer.items():
            print(k, v, flush=True)
    return {0: layer}


def next_dynamic_definition_schedule(recursion_layer: int, schedule: dict, state_idx: int,
                                     recursion_qubit: int, verbose: int = 0) -> dict:
    """Zoom into bin `state_idx` of `schedule` (distributor.py:62-95): its active qubits become
    fixed bits (bin index read most significant first, subcircuits in the OLD smart_order), then
    subcircuits are re-sorted by merged-qubit count and the leading merged qubits activated."""
    nxt = copy.deepcopy(schedule)
    states = nxt["subcircuit_state"]
    width = sum(st.count("active") for st in states.values())
    bits = format(state_idx, "b").zfill(width)
    pos = 0
    for s in nxt["smart_order"]:
        for i, x in enumerate(states[s]):
            if x == "active":
                states[s][i] = bits[pos]
                pos += 1
    nxt["smart_order"] = sorted(nxt["smart_order"], key=lambda s: states[s].count("merged"))
    grant = _share([states[s].count("merged") for s in nxt["smart_order"]], recursion_qubit)
    for s, a in zip(nxt["smart_order"], grant):
        for i, x in enumerate(states[s]):
            if a == 0:
                break
            if x == "merged":
                states[s][i] = "active"
                a -= 1
    nxt["upper_bin"] = (recursion_layer, state_idx)
    if verbose > 0:
        for k, v in nxt.items():
            print(k, v, flush=True)
    return nxt


def find_max_recursion_layer(schedules: Dict[int, dict], build_outputs: Dict[int, dict]) -> int:
    """Which finished layer to zoom into next (distributor.py:97-117): the one whose next
    unexplored bin (`max_states[zoomed_ctr]`) is the most probable, > 1e-16, among layers that
    still have merged qubits; -1 when there is none."""
    best, best_p = -1, 0.0
    for layer in sorted(build_outputs):
        out = build_outputs[layer]
        merged = sum(st.count("merged") for st in schedules[layer]["subcircuit_state"].values())
        if merged == 0 or out["zoomed_ctr"] == len(out["max_states"]):
            continue
        p = float(out["reconstructed_prob"][out["max_states"][out["zoomed_ctr"]]])
        if p > best_p and p > 1e-16:
            best, best_p = layer, p
    return best


def merge_states(layer_schedule: dict, subcircuit_idx: int) -> List[int]:
    """Qubit states of one subcircuit in `merge_%d.txt` encoding (distributor.py:157-165):
    -2 merged, -1 active, 0 / 1 zoomed."""
    code = {"merged": -2, "active": -1, "0": 0, "1": 1}
    out = []
    for x in layer_schedule["subcircuit_state"][subcircuit_idx]:
        if x not in code:
            raise Exception("Illegal qubit_state %s" % x)
        out.append(code[x])
    return out


def find_process_jobs(jobs, rank: int, num_threads: int):
    """Contiguous split of `jobs` over ranks, the first len % n ranks taking one more
    (qiskit_helper_functions.non_ibmq_functions.find_process_jobs, used by distributor.py:149,168)."""
    jobs = list(jobs)
    count, extra = divmod(len(jobs), num_threads)
    lo = rank * count + min(rank, extra)
    return jobs[lo:lo + count + (1 if rank < extra else 0)]


def write_files(recursion_layer, num_threads, counter, vertical_collapse_folder,
                dynamic_definition_folder, dynamic_definition_schedule, summation_terms, num_cuts,
                num_subcircuits, verbose=0):
    """merge_%d.txt and build_%d.txt per rank (distributor.py:143-187)."""
    layer = dynamic_definition_schedule[recursion_layer]
    total_active = sum(st.count("active") for st in layer["subcircuit_state"].values())
    for rank in range(num_threads):
        mine = []
        for s in counter:
            files = glob.glob("%s/kron_%d_*.txt" % (vertical_collapse_folder, s))
            mine += find_process_jobs(files, rank, num_threads)
        with open("%s/merge_%d.txt" % (dynamic_definition_folder, rank), "w") as f:
            f.write("num_files_to_merge=%d\n" % len(mine))
            for path in mine:
                s, k = (int(x) for x in os.path.basename(path)[len("kron_"):-len(".txt")].split("_"))
                f.write("subcircuit_idx=%d subcircuit_kron_index=%d num_effective=%d num_active=%d\n"
                        % (s, k, counter[s]["effective"], layer["subcircuit_state"][s].count("active")))
                f.write("".join("%d " % x for x in merge_states(layer, s)) + "\n")
        terms = find_process_jobs(summation_terms, rank, num_threads)
        if verbose > 0:
            print("Rank %d has %d/%d summation terms" % (rank, len(terms), len(summation_terms)),
                  flush=True)
        with open("%s/build_%d.txt" % (dynamic_definition_folder, rank), "w") as f:
            f.write("total_active_qubit=%d num_subcircuits=%d num_summation_terms=%d num_cuts=%d\n"
                    % (total_active, num_subcircuits, len(terms), num_cuts))
            for term in terms:
                f.write("".join("%d,%d " % (s, term[s]) for s in layer["smart_order"]) + "\n")
