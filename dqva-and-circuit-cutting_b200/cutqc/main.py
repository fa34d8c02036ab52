"""CutQC driver, in memory and on the device (reference: cutqc/main.py).

The reference's `CutQC` shuttles every intermediate through `cutqc_data/` text files and four
C executables; cutting itself needs Gurobi (cutqc/cutter.py) and is commented out of its
`__init__` (main.py:27).  Here the object is built from a *cut solution* -- the dict the
reference pickles as `subcircuits.pckl` (cutter.py:541-553: 'circuit', 'subcircuits',
'complete_path_map', 'counter') -- or straight from qsplit fragments (`from_fragments`), and
the stages run as kernels on device-resident arrays:

    evaluate      simulate all instances (engine) -> measure -> vertical_collapse
    post_process  per recursion layer: schedule -> merge -> build -> (zoom) ...
    reorder       bin vector -> {bitstring: probability} in full-circuit order

File-protocol compatible stage executables live in `cutqc/bin/` (see `cutqc/stages.py`).
"""
from __future__ import annotations

import os
import pickle
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from .. import _lib
from .. import recombine as _rc
from ..engine import B200Backend
from . import kernels
from .distributor import (find_max_recursion_layer, initialize_dynamic_definition_schedule,
                          merge_states, next_dynamic_definition_schedule)
from .evaluator import simulate_subcircuits
from .post_process import TermTables

__all__ = ["CutQC", "ReconstructedLayer", "complete_path_map_from_wire_path_map", "get_counter"]


def complete_path_map_from_wire_path_map(wire_path_map):
    """qsplit's `{wire: ((fragment, fragment wire), ...)}` (qsplit_circuit_cutter.py:86-246) in
    cutqc's layout `{wire: [{'subcircuit_idx', 'subcircuit_qubit'}, ...]}` (cutter.py:299-386)."""
    return {wire: [{"subcircuit_idx": int(f), "subcircuit_qubit": q} for f, q in path]
            for wire, path in wire_path_map.items()}


def get_counter(subcircuits, O_rho_pairs):
    """{'effective', 'rho', 'O', 'd'} per subcircuit (cutter.py:450-461)."""
    counter = {s: {"effective": sub.num_qubits, "rho": 0, "O": 0, "d": sub.num_qubits}
               for s, sub in enumerate(subcircuits)}
    for o, r in O_rho_pairs:
        counter[o["subcircuit_idx"]]["effective"] -= 1
        counter[o["subcircuit_idx"]]["O"] += 1
        counter[r["subcircuit_idx"]]["rho"] += 1
    return counter


class ReconstructedLayer:
    """One recursion layer: `prob` (device float vector over the active bins, most significant
    bit = first active qubit of the first subcircuit in `smart_order`) + the bookkeeping that
    places each bin in the full circuit (main.py:141-196)."""

    def __init__(self, prob, schedule, wires, states, num_qubits):
        self.prob, self.schedule = prob, schedule
        self.wires, self.states, self.num_qubits = wires, states, num_qubits
        self.active_wires = [w for w, s in zip(wires, states) if s == "active"]   # MSB first
        self.merged_wires = [w for w, s in zip(wires, states) if s == "merged"]
        self.fixed = {w: int(s) for w, s in zip(wires, states) if s in ("0", "1")}

    @property
    def num_active(self):
        return len(self.active_wires)

    def full_index(self) -> torch.Tensor:
        """int64[2^active]: full-circuit basis state of every bin with merged wires at 0."""
        A = self.num_active
        u = torch.arange(1 << A, dtype=torch.int64, device=self.prob.device)
        base = sum(b << w for w, b in self.fixed.items())
        out = torch.full_like(u, base)
        for j, w in enumerate(self.active_wires):
            out |= ((u >> (A - 1 - j)) & 1) << w
        return out

    def dense(self) -> torch.Tensor:
        """float64[2^n] in full-circuit order; a merged bin is spread uniformly (main.py:174-175)."""
        n = self.num_qubits
        out = torch.zeros(1 << n, dtype=torch.float64, device=self.prob.device)
        idx = self.full_index()
        share = self.prob.to(torch.float64) / (1 << len(self.merged_wires))
        for m in range(1 << len(self.merged_wires)):
            off = sum(((m >> j) & 1) << w for j, w in enumerate(self.merged_wires))
            out[idx | off] = share
        return out

    def to_dict(self, threshold: float = 1e-10) -> Dict[str, float]:
        """main.py:191-195: bins above `threshold`, keys = n-character bitstrings (wire n-1 first)."""
        idx = self.full_index().cpu().numpy()
        share = self.prob.to(torch.float64).cpu().numpy() / (1 << len(self.merged_wires))
        out = {}
        for m in range(1 << len(self.merged_wires)):
            off = sum(((m >> j) & 1) << w for j, w in enumerate(self.merged_wires))
            for i in np.nonzero(share > threshold)[0]:
                out[format(int(idx[i]) | off, "0%db" % self.num_qubits)] = float(share[i])
        return out

    def score(self, graph, threshold: float = 1e-5, k: int = 1, num_anc: int = 0):
        """Fused MIS objective + top-k of this layer without reordering (mis.py:117-134, 232-233);
        needs every output wire active (no merged / zoomed wires)."""
        if self.merged_wires or self.fixed:
            raise _lib.QdcError("score() needs a layer with every output qubit active")
        A = self.num_active
        n_nodes = self.num_qubits - num_anc
        adj = [0] * self.num_qubits
        for a, b in graph.edges():
            adj[int(a)] |= 1 << int(b)
            adj[int(b)] |= 1 << int(a)
        bitpos = [self.active_wires[A - 1 - b] for b in range(A)]
        cb = min(A, 14)
        tabs = _rc.build_tables(A - cb, cb, bitpos, adj, (1 << n_nodes) - 1, device=self.prob.device)
        return _rc.reduce_and_score(self.prob, tabs, threshold=threshold, k=k)


class CutQC:
    """`CutQC(cut_solution, backend)`; see the module docstring.

    evaluate(early_termination=(0,), params=None, param_order=None)
        -> self.kron[et][s]: float64 [n_kron_s, 2^effective_s] (kron_%d_%d.txt contents) and
           self.kron_present[et][s]: bool[n_kron_s] (False = file not written, early termination)
    post_process(early_termination=0, qubit_limit=None, recursion_depth=1) -> [ReconstructedLayer]
    """

    def __init__(self, cut_solution: dict, backend: B200Backend, verbose: int = 0):
        if not isinstance(backend, B200Backend):
            raise _lib.QdcError("CutQC needs a qdc_b200.engine.B200Backend; there is no CPU path")
        self.backend, self.verbose = backend, verbose
        self.circuit = cut_solution["circuit"]
        self.subcircuits = list(cut_solution["subcircuits"])
        self.complete_path_map = cut_solution["complete_path_map"]
        self.terms = TermTables(self.subcircuits, self.complete_path_map)
        self.counter = cut_solution.get("counter") or get_counter(self.subcircuits, self.terms.pairs)
        self.num_cuts = self.terms.num_cuts
        self.measured: List[torch.Tensor] = []
        self.kron: Dict[int, List[torch.Tensor]] = {}
        self.kron_present: Dict[int, List[torch.Tensor]] = {}

    @classmethod
    def from_fragments(cls, circuit, fragments, wire_path_map, backend, verbose: int = 0):
        cpm = complete_path_map_from_wire_path_map(wire_path_map)
        return cls({"circuit": circuit, "subcircuits": list(fragments), "complete_path_map": cpm},
                   backend, verbose)

    # ---- evaluate (main.py:72-84) -----------------------------------------------------------
    def evaluate(self, early_termination: Sequence[int] = (0,), params=None, param_order=None,
                 fused: bool = False):
        """fused=False: the reference's stages (all 4^rho 3^O instances -> measure ->
        vertical_collapse).  fused=True: the kron vectors straight from the amplitudes of the
        2^rho basis-input evolutions (`qdc_sv_pauli_tensor`, basis-input form): the init
        expansion of post_process.py:20-30 is Tr(P_a .) on the rho axes and the signed marginal of
        measure.c is Tr(. P_a) on the O axes, so no instance is simulated or stored
        (subcircuits with at most 4 cut ends; others fall back to the staged path)."""
        if fused:
            return self._evaluate_fused(early_termination, params, param_order)
        sims = simulate_subcircuits(self.backend, self.subcircuits, self.complete_path_map,
                                    params, param_order)
        self.measured = []
        for s, (tab, raw) in enumerate(sims):
            if raw.shape[2] != 1:
                raise _lib.QdcError("CutQC.evaluate takes one parameter vector at a time")
            raw = raw[:, :, 0]                                         # [n_init, 3^O, 2^d]
            parent = torch.as_tensor(tab.parent, device=raw.device)
            rows = raw.index_select(1, parent).reshape(tab.n_instances, 1 << tab.d)
            signs = np.tile(tab.sign_mask, tab.n_init)
            self.measured.append(kernels.measure(rows.contiguous(), tab.keep_mask, signs))
        for et in early_termination:
            self.kron[et], self.kron_present[et] = [], []
            for s, measured in enumerate(self.measured):
                vecs, flags = kernels.vertical_collapse(measured, self.terms.kron_terms[s])
                self.kron[et].append(vecs)
                self.kron_present[et].append(flags.bool() if et == 1 else torch.ones_like(flags).bool())
        return self

    def _evaluate_fused(self, early_termination, params, param_order):
        from ..engine import BASIS_KETS
        from .evaluator import _param_order_of
        order = _param_order_of(self.subcircuits, param_order)
        table = None if params is None else np.atleast_2d(np.asarray(params, dtype=np.float64))
        if table is not None and table.shape[0] != 1:
            raise _lib.QdcError("CutQC.evaluate takes one parameter vector at a time")
        kron = []
        staged = None
        for s, sub in enumerate(self.subcircuits):
            tab, mine = self.terms.tables[s], self.terms.pairs_of[s]
            if len(tab.rho) + len(tab.O) > 4:
                if staged is None:                       # rare: wide cut, use the reference's stages
                    staged = CutQC({"circuit": self.circuit, "subcircuits": self.subcircuits,
                                    "complete_path_map": self.complete_path_map}, self.backend)
                    staged.evaluate((0,), params, param_order)
                kron.append(staged.kron[0][s])
                continue
            rank = {k: len(mine) - 1 - i for i, k in enumerate(mine)}        # base-4 digit of pair k
            rho_pair = [k for k in mine if self.terms.pairs[k][1]["subcircuit_idx"] == s]
            O_pair = [k for k in mine if self.terms.pairs[k][0]["subcircuit_idx"] == s]
            strides = [4 ** rank[k] for k in rho_pair + O_pair]
            zero_mask = ((1 << tab.d) - 1) & ~sum(1 << q for q in tab.rho)
            prog, auto = self.backend.compile_instructions(tab.d, list(sub.to_instructions()), order, zero_mask)
            states = self.backend.run(prog, auto, table, prep_qubits=tab.rho, kets=BASIS_KETS)
            T = self.backend.pauli_tensor(states, prog, 1, len(tab.rho), tab.O,
                                          axis_strides=strides if mine else None,
                                          basis_inputs=bool(tab.rho))
            kron.append(T[0].t().contiguous())                               # [kron index][2^effective]
        self.measured = []
        for et in early_termination:
            self.kron[et] = kron
            self.kron_present[et] = [(v.abs().amax(dim=1) > 1e-16) if et == 1
                                     else torch.ones(v.shape[0], dtype=torch.bool, device=v.device)
                                     for v in kron]
        return self

    # ---- post_process (main.py:86-139, 391-457) ---------------------------------------------
    def _layer(self, schedule, et, out_dtype):
        order = schedule["smart_order"]
        merged, index = [], []
        for s in order:
            merged.append(kernels.merge(self.kron[et][s], merge_states(schedule, s)))
            present = self.kron_present[et][s].cpu().numpy()
            idx = self.terms.term_index[s].copy()
            idx[~present[idx]] = -1                                     # build.c:35-67
            index.append(idx)
        prob = kernels.build(merged, index, self.num_cuts, out_dtype=out_dtype)
        return prob

    def post_process(self, early_termination: int = 0, qubit_limit: Optional[int] = None,
                     recursion_depth: int = 1, out_dtype=torch.float64,
                     dest_folder: Optional[str] = None, resume: bool = False) -> List[ReconstructedLayer]:
        """Dynamic-definition recursion (main.py:86-139 driving distributor.py:186-260).

        `dest_folder` (optional): mirror the recursion state to disk in the reference's layout -
        `meta_data.pckl` and `dynamic_definition_<layer>/build_output.pckl`
        (distributor.py:121-141, 212-215; main.py:452-457; names via helper_fun.get_dirname).
        `resume=True` continues from the state found there, as the reference does when it is
        re-invoked with the next `recursion_layer`: layers already on disk are loaded, not rebuilt."""
        if early_termination not in self.kron:
            raise _lib.QdcError("call evaluate(early_termination=(%d,)) first" % early_termination)
        if qubit_limit is None:
            qubit_limit = sum(c["effective"] for c in self.counter.values())
        schedules: Dict[int, dict] = {}
        outputs: Dict[int, dict] = {}
        layers = []
        first = 0
        if resume:
            if dest_folder is None:
                raise _lib.QdcError("resume=True needs the dest_folder that holds meta_data.pckl")
            schedules, outputs = self._load_state(dest_folder)
            for layer in sorted(outputs):
                layers.append(self._wrap(outputs[layer]["reconstructed_prob"], schedules[layer]))
            first = len(outputs)
        for layer in range(first, recursion_depth):
            if layer == 0:
                schedules = initialize_dynamic_definition_schedule(self.counter, qubit_limit, self.verbose)
            else:
                src = find_max_recursion_layer(schedules, outputs)
                if src < 0:
                    break                                               # distributor.py:123-125
                out = outputs[src]
                schedules[layer] = next_dynamic_definition_schedule(
                    src, schedules[src], int(out["max_states"][out["zoomed_ctr"]]), qubit_limit,
                    self.verbose)
                out["zoomed_ctr"] += 1
                if dest_folder is not None:
                    self._dump_build_output(dest_folder, src, out)       # distributor.py:137-138
            prob = self._layer(schedules[layer], early_termination, out_dtype)
            # main.py:452: bins by decreasing probability (stable, like sorted(..., reverse=True))
            order = torch.sort(prob, descending=True, stable=True).indices
            outputs[layer] = {"zoomed_ctr": 0, "max_states": order, "reconstructed_prob": prob}
            layers.append(self._wrap(prob, schedules[layer]))
            if dest_folder is not None:
                os.makedirs(dest_folder, exist_ok=True)
                with open(os.path.join(dest_folder, "meta_data.pckl"), "wb") as f:
                    pickle.dump({"counter": self.counter, "dynamic_definition_schedule": schedules}, f)
                self._dump_build_output(dest_folder, layer, outputs[layer])
        self.schedules = schedules
        return layers

    @staticmethod
    def _dump_build_output(dest_folder, layer, out):
        folder = os.path.join(dest_folder, "dynamic_definition_%d" % layer)
        os.makedirs(folder, exist_ok=True)
        with open(os.path.join(folder, "build_output.pckl"), "wb") as f:
            pickle.dump({"zoomed_ctr": int(out["zoomed_ctr"]),
                         "max_states": [int(x) for x in out["max_states"].cpu().tolist()],
                         "reconstructed_prob": out["reconstructed_prob"].double().cpu().numpy()}, f)

    def _load_state(self, dest_folder):
        from .helper_fun import read_dict
        meta = read_dict(os.path.join(dest_folder, "meta_data.pckl"))
        if not meta:
            raise _lib.QdcError(f"no meta_data.pckl under {dest_folder}")
        schedules = dict(meta["dynamic_definition_schedule"])
        outputs = {}
        dev = self.backend.device
        for layer in sorted(schedules):
            rec = read_dict(os.path.join(dest_folder, "dynamic_definition_%d" % layer, "build_output.pckl"))
            if not rec:
                raise _lib.QdcError(f"layer {layer} of {dest_folder} has a schedule but no build_output.pckl")
            outputs[layer] = {"zoomed_ctr": int(rec["zoomed_ctr"]),
                              "max_states": torch.as_tensor(np.asarray(rec["max_states"], dtype=np.int64), device=dev),
                              "reconstructed_prob": torch.as_tensor(np.asarray(rec["reconstructed_prob"]), device=dev)}
        return schedules, outputs

    def _wrap(self, prob, layer_schedule) -> ReconstructedLayer:
        full_qubits = list(self.circuit.qubits)
        out_wires: Dict[int, list] = {}
        for wire, path in self.complete_path_map.items():
            last = path[-1]
            qubits = list(self.subcircuits[last["subcircuit_idx"]].qubits)
            out_wires.setdefault(last["subcircuit_idx"], []).append(
                (qubits.index(last["subcircuit_qubit"]), full_qubits.index(wire)))
        wires, states = [], []
        for s in layer_schedule["smart_order"]:
            if s in out_wires:
                wires += [w for _, w in sorted(out_wires[s], reverse=True)]
                states += layer_schedule["subcircuit_state"][s]
        return ReconstructedLayer(prob, layer_schedule, wires, states, len(full_qubits))

    def reorder(self, full_circuit, unordered, complete_path_map, subcircuits, layer_schedule):
        """Reference signature (main.py:141); `unordered` may be a tensor or an array."""
        if not isinstance(unordered, torch.Tensor):
            unordered = torch.as_tensor(np.asarray(unordered, dtype=np.float64))
        return self._wrap(unordered, layer_schedule).to_dict()
