"""Subcircuit instances of the CutQC path (reference: cutqc/evaluator.py).

The reference builds one qiskit circuit per (initialisation, measurement-basis) combination and
simulates them one by one in a process pool (cutqc/main.py:198-240).  Here a subcircuit is
compiled ONCE; the 4^rho initialisations are the engine's preparation batch and the 3^O basis
changes are rows of its parameter table (a `u1(l) . ry(t)` pair per O qubit reproduces I / H /
H.Sdg exactly), so all instances of a subcircuit are one `qdc_sv_run` batch.

Index conventions are the reference's and are kept bit-exact:
  * initialisation labels ('zero', 'one', 'plus', 'plusI') per rho qubit, measurement labels
    ('I', 'X', 'Y') per O qubit, `itertools.product` order, first listed qubit most significant
    (evaluator.py:34-58);
  * `indexed_combinations[(inits, meas)]` numbers the *mutated* combinations, every 'I' also
    standing for 'Z' (evaluator.py:59-80) -- here `InstanceTable.index(init_digits, meas_digits)`.
"""
from __future__ import annotations

import itertools
import math
from typing import Dict, List, Sequence, Tuple

import numpy as np
import torch

from .. import _lib
from ..circuit import Parameter, QuantumCircuit
from ..engine import B200Backend, CUTQC_KETS

__all__ = ["INIT_LABELS", "MEAS_LABELS", "PAULI_LABELS", "find_subcircuit_O_rho_qubits",
           "find_all_combinations", "mutate_measurement_basis", "get_subcircuit_instance",
           "generate_subcircuit_instances", "InstanceTable", "simulate_subcircuits"]

INIT_LABELS = ("zero", "one", "plus", "plusI")          # evaluator.py:36
MEAS_LABELS = ("I", "X", "Y")                           # evaluator.py:35
PAULI_LABELS = ("I", "X", "Y", "Z")                     # post_process.py:61


def find_subcircuit_O_rho_qubits(complete_path_map, subcircuit_idx):
    """Cut ends inside one subcircuit (evaluator.py:20-32): O = a wire leaves here (every path
    element but the last), rho = a wire enters here (every element but the first)."""
    O_qubits, rho_qubits = [], []
    for path in complete_path_map.values():
        for pos, end in enumerate(path):
            if end["subcircuit_idx"] != subcircuit_idx:
                continue
            if pos < len(path) - 1:
                O_qubits.append(end)
            if pos > 0:
                rho_qubits.append(end)
    # the reference appends all O ends of a wire before its rho ends; the two lists are
    # separate, so the order inside each list is the same as here
    return O_qubits, rho_qubits


def mutate_measurement_basis(meas):
    """Every 'I' also yields the 'Z' instance from the same raw data (evaluator.py:69-80)."""
    if "I" not in meas:
        return [meas]
    return list(itertools.product(*[("I", "Z") if x == "I" else (x,) for x in meas]))


class InstanceTable:
    """Instance numbering of one subcircuit as arrays.

    d qubits; `rho` / `O` = subcircuit qubit indices of the cut ends in reference list order.
    Unmutated combination (what is simulated) = init * 3^O + meas3, mutated instance (what
    `measure` produces) = init * 4^O + mut, where `mut` enumerates, for meas3 ascending, the
    I->{I,Z} expansions in product order (evaluator.py:59-66)."""

    def __init__(self, num_qubits: int, rho: Sequence[int], O: Sequence[int]):
        self.d, self.rho, self.O = int(num_qubits), list(rho), list(O)
        nO = len(self.O)
        self.n_init, self.n_meas3, self.n_mut = 4 ** len(self.rho), 3 ** nO, 4 ** nO
        self.mut_of_code = np.zeros(self.n_mut, dtype=np.int64)     # 4-label code -> mut
        self.parent = np.zeros(self.n_mut, dtype=np.int64)          # mut -> meas3
        self.sign_mask = np.zeros(self.n_mut, dtype=np.uint64)      # mut -> qubits with X/Y/Z
        self.labels: List[Tuple[str, ...]] = [()] * self.n_mut      # mut -> labels per O qubit
        mut = 0
        for m3, meas in enumerate(itertools.product(range(3), repeat=nO)):
            free = [j for j, a in enumerate(meas) if a == 0]
            for zs in itertools.product((0, 1), repeat=len(free)):
                lab = list(meas)
                for j, z in zip(free, zs):
                    lab[j] = 3 if z else 0
                code = 0
                for a in lab:
                    code = code * 4 + a
                self.mut_of_code[code] = mut
                self.parent[mut] = m3
                self.sign_mask[mut] = sum(1 << self.O[j] for j, a in enumerate(lab) if a != 0)
                self.labels[mut] = tuple(PAULI_LABELS[a] for a in lab)
                mut += 1
        assert mut == self.n_mut
        self.keep_mask = ((1 << self.d) - 1) & ~sum(1 << q for q in self.O)

    @property
    def effective(self) -> int:
        return self.d - len(self.O)

    @property
    def n_instances(self) -> int:
        return self.n_init * self.n_mut

    def index(self, init_digits: Sequence[int], meas_digits: Sequence[int]) -> int:
        """init_digits[j] in 0..3 (INIT_LABELS) per rho qubit, meas_digits[j] in 0..3
        (PAULI_LABELS) per O qubit -> the reference's `indexed_combinations` value."""
        i = 0
        for a in init_digits:
            i = i * 4 + int(a)
        code = 0
        for a in meas_digits:
            code = code * 4 + int(a)
        return i * self.n_mut + int(self.mut_of_code[code])

    def key(self, instance: int):
        """Inverse of `index` in the reference's dict-key form (tuple(inits), tuple(meas))."""
        i, mut = divmod(int(instance), self.n_mut)
        inits = ["zero"] * self.d
        for j in reversed(range(len(self.rho))):
            i, a = divmod(i, 4)
            inits[self.rho[j]] = INIT_LABELS[a]
        meas = ["comp"] * self.d
        for j, lab in enumerate(self.labels[mut]):
            meas[self.O[j]] = lab
        return tuple(inits), tuple(meas)


def _table_for(subcircuit, O_qubits, rho_qubits) -> InstanceTable:
    qubits = list(subcircuit.qubits)
    return InstanceTable(len(qubits), [qubits.index(q["subcircuit_qubit"]) for q in rho_qubits],
                         [qubits.index(q["subcircuit_qubit"]) for q in O_qubits])


def find_all_combinations(O_qubits, rho_qubits, qubits):
    """Reference-layout view (evaluator.py:34-67): (combinations, indexed_combinations)."""
    class _Q:                                            # only `.qubits` is needed
        pass
    holder = _Q()
    holder.qubits = list(qubits)
    tab = _table_for(holder, O_qubits, rho_qubits)
    combos = []
    for i in range(tab.n_init):
        inits, _ = tab.key(i * tab.n_mut)
        for meas3 in itertools.product(MEAS_LABELS, repeat=len(tab.O)):
            meas = ["comp"] * tab.d
            for q, lab in zip(tab.O, meas3):
                meas[q] = lab
            combos.append((list(inits), meas))
    indexed = {tab.key(k): k for k in range(tab.n_instances)}
    return combos, indexed


_INIT_OPS = {"zero": (), "one": ("x",), "plus": ("h",), "minus": ("x", "h"), "plusI": ("h", "s"),
             "minusI": ("x", "h", "s")}                  # acting order; evaluator.py:89-107
_MEAS_OPS = {"I": (), "comp": (), "X": ("h",), "Y": ("sdg", "h")}     # evaluator.py:109-119


def get_subcircuit_instance(subcircuit_idx, subcircuit, combinations):
    """Explicit per-instance circuits with the reference's keys and shot rule
    (evaluator.py:82-124).  The device path does not need them (see `simulate_subcircuits`);
    they exist for callers that want to run an instance elsewhere."""
    out = {}
    for inits, meas in combinations:
        inst = QuantumCircuit(name=getattr(subcircuit, "name", None))
        inst.qregs, inst.qubits = list(subcircuit.qregs), list(subcircuit.qubits)
        inst.ancillas = list(getattr(subcircuit, "ancillas", []))
        for q, lab in zip(inst.qubits, inits):
            if lab not in _INIT_OPS:
                raise Exception("Illegal initialization : ", lab)
            for g in _INIT_OPS[lab]:
                getattr(inst, g)(q)
        inst.data.extend(op.copy() for op in subcircuit.data)
        for q, lab in zip(inst.qubits, meas):
            if lab not in _MEAS_OPS:
                raise Exception("Illegal measurement basis:", lab)
            for g in _MEAS_OPS[lab]:
                getattr(inst, g)(q)
        shots = min(8192 * 10, max(8192, int(2 ** inst.num_qubits)))
        out[(subcircuit_idx, tuple(inits), tuple(meas))] = {"circuit": inst, "shots": shots}
    return out


def generate_subcircuit_instances(subcircuits, complete_path_map):
    """(circ_dict, all_indexed_combinations) as evaluator.py:10-18."""
    circ_dict, all_indexed = {}, {}
    for s, sub in enumerate(subcircuits):
        O_q, rho_q = find_subcircuit_O_rho_qubits(complete_path_map, s)
        combos, all_indexed[s] = find_all_combinations(O_q, rho_q, sub.qubits)
        circ_dict.update(get_subcircuit_instance(s, sub, combos))
    return circ_dict, all_indexed


# measurement-basis change as parameters: u1(l) then ry(t) on the O qubit
#   I: (0, 0) -> identity;  X: (pi, pi/2) -> ry(pi/2).z = h;  Y: (pi/2, pi/2) -> ry(pi/2).z.sdg = h.sdg
_MEAS_ANGLES = ((0.0, 0.0), (math.pi, math.pi / 2), (math.pi / 2, math.pi / 2))


def _param_order_of(subcircuits, param_order=None):
    if param_order is not None:
        return list(param_order)
    seen = []
    for sub in subcircuits:
        for prm in getattr(sub, "parameters", []):
            if prm not in seen:
                seen.append(prm)
    from ..circuit import _param_sort_key
    return sorted(seen, key=_param_sort_key)


def simulate_subcircuits(backend: B200Backend, subcircuits, complete_path_map, params=None,
                         param_order=None):
    """All instances of every subcircuit on the device.

    Returns a list (per subcircuit) of `(table, raw)`: `raw` float64 [n_init, 3^O, n_sets, 2^d],
    the statevector probabilities the reference writes to raw_%d_%d.txt (evaluator.py:126-155),
    index = qiskit little-endian basis state."""
    if not isinstance(backend, B200Backend):
        raise _lib.QdcError("simulate_subcircuits needs a qdc_b200.engine.B200Backend; there is "
                            "no CPU simulator in this package")
    param_order = _param_order_of(subcircuits, param_order)
    user = np.zeros((1, len(param_order))) if params is None else \
        np.atleast_2d(np.asarray(params, dtype=np.float64))
    n_user = user.shape[0]
    out = []
    for s, sub in enumerate(subcircuits):
        O_q, rho_q = find_subcircuit_O_rho_qubits(complete_path_map, s)
        tab = _table_for(sub, O_q, rho_q)
        insts = list(sub.to_instructions())
        extra = []
        for j, q in enumerate(tab.O):
            lam, theta = Parameter(f"cutqc_meas_l_{s}_{j}"), Parameter(f"cutqc_meas_t_{s}_{j}")
            extra += [lam, theta]
            insts += [("u1", (q,), (lam,), None), ("ry", (q,), (theta,), None)]
        order = list(param_order) + extra
        table = np.zeros((tab.n_meas3, n_user, len(order)))
        table[:, :, :len(param_order)] = user[None]
        for m3, meas in enumerate(itertools.product(range(3), repeat=len(tab.O))):
            for j, a in enumerate(meas):
                table[m3, :, len(param_order) + 2 * j] = _MEAS_ANGLES[a][0]
                table[m3, :, len(param_order) + 2 * j + 1] = _MEAS_ANGLES[a][1]
        zero_mask = ((1 << tab.d) - 1) & ~sum(1 << q for q in tab.rho)
        prog, auto = backend.compile_instructions(tab.d, insts, order, zero_mask)
        states = backend.run(prog, auto, table.reshape(tab.n_meas3 * n_user, len(order)), prep_qubits=tab.rho,
                             kets=CUTQC_KETS)
        raw = backend.probabilities(states, prog)
        del states
        out.append((tab, raw.reshape(tab.n_init, tab.n_meas3, n_user, 1 << tab.d)))
    return out
