"""Wire cutter: `cut_circuit(circuit, cuts) -> (fragments, wire_path_map)`.

Drop-in for the reference's qsplit/qsplit_circuit_cutter.py:86-246 without qiskit's DAGCircuit
or retworkx.  Same contract:

* `cuts` is a list of `(wire, op_number)`: the wire is severed after `op_number` operations
  have acted on it (barriers not counted, :127), everything downstream moves to a fresh wire
  of a new register named `"<old names>_new"` (:98-101); cuts are processed from the largest
  `op_number` to the smallest (:132).
* fragments are the connected components of the operation graph (:23-46) re-indexed densely
  (:50-76); qubits never acted on form one extra trivial fragment (:207-225).
* `wire_path_map[circuit_wire] = ((frag_idx, frag_wire), ...)` lists the fragment wires the
  circuit wire passes through, in order (:227-241).

Where the reference's ordering is implementation-defined we fix a canonical one (documented
deviation): fragments are ordered by the smallest extended-circuit wire they
contain (what retworkx's `weakly_connected_components` yields for a DAG whose input nodes were
added in wire order), and fragment-local wires are numbered in ascending extended-circuit order
(the reference iterates a Python `set` of Qubit objects, :52, which has no defined order).
`united_axis_permutation` makes the recombined distribution independent of both choices.

Pure host-side integer bookkeeping; no amplitudes are touched here.
"""
from __future__ import annotations

from typing import Dict, List, Sequence, Tuple

from ..circuit import Instruction, QuantumCircuit, QuantumRegister, Qubit

__all__ = ["cut_circuit"]


class _DSU:
    def __init__(self, n):
        self.p = list(range(n))

    def find(self, a):
        while self.p[a] != a:
            self.p[a] = self.p[self.p[a]]
            a = self.p[a]
        return a

    def union(self, a, b):
        ra, rb = self.find(a), self.find(b)
        if ra != rb:
            self.p[max(ra, rb)] = min(ra, rb)


def cut_circuit(circuit: QuantumCircuit, cuts: Sequence[Tuple[Qubit, int]],
                qreg_name: str = "q", creg_name: str = "c"):
    assert all(isinstance(wire, Qubit) for wire, _ in cuts)

    circuit_wires: List[Qubit] = list(circuit.qubits) + list(circuit.clbits)
    ops: List[Instruction] = [inst.copy() for inst in circuit.data if inst.name != "barrier"]

    # extended wire list: circuit wires, then one new wire per cut
    ext_wires: List[Qubit] = list(circuit.qubits)
    if len(cuts) > 0:
        old_reg_names = []
        for wire in circuit_wires:
            if wire.register.prefix not in old_reg_names:
                old_reg_names.append(wire.register.prefix)
        new_register = QuantumRegister(len(cuts), "_".join(old_reg_names) + "_new")
        new_wires = iter(new_register)
        ext_wires += list(new_register)

    stitches: Dict[Qubit, Qubit] = {}
    for cut_wire, cut_location in sorted(cuts, key=lambda cut: -cut[1]):
        wire_ops = [op for op in ops if cut_wire in op.qargs]
        if not 0 <= cut_location < len(wire_ops):
            raise IndexError(f"cut location {cut_location} out of range for {cut_wire} "
                             f"({len(wire_ops)} operations)")
        new_wire = next(new_wires)
        # every operation on this wire from the cut onward is a DAG descendant of the cut
        # node, so it is re-pointed at the new wire (:175-178)
        for op in wire_ops[cut_location:]:
            op.qargs[op.qargs.index(cut_wire)] = new_wire
        stitches = {(start if start != cut_wire else new_wire):
                    (end if end != cut_wire else new_wire)
                    for start, end in stitches.items()}                  # :181-183
        stitches[cut_wire] = new_wire                                     # :186

    # connected components of the operation graph
    pos = {w: i for i, w in enumerate(ext_wires)}
    dsu = _DSU(len(ext_wires))
    used = [False] * len(ext_wires)
    for op in ops:
        idxs = [pos[q] for q in op.qargs]
        for i in idxs:
            used[i] = True
        for i in idxs[1:]:
            dsu.union(idxs[0], i)
    roots = sorted({dsu.find(i) for i in range(len(ext_wires)) if used[i]})
    comp_wires = [[ext_wires[i] for i in range(len(ext_wires))
                   if used[i] and dsu.find(i) == r] for r in roots]

    subcircuits: List[QuantumCircuit] = []
    subgraph_wire_maps: List[Dict[Qubit, Qubit]] = []
    for wires in comp_wires:
        new_qbits = QuantumRegister(len(wires), qreg_name)
        frag = QuantumCircuit(new_qbits)
        register_map = {old: new for old, new in zip(wires, new_qbits)}
        wire_set = set(wires)
        for op in ops:
            if op.qargs[0] in wire_set:
                frag.data.append(Instruction(op.name, [register_map[q] for q in op.qargs],
                                             list(op.params), op.matrix))
        subcircuits.append(frag)
        subgraph_wire_maps.append(register_map)

    unused_qubits = [q for q in circuit.qubits if not used[pos[q]]]       # :207-225
    if unused_qubits:
        print("WARNING: some qubits are entirely unused")
        print("unused qubits:", unused_qubits)
        qreg = QuantumRegister(len(unused_qubits), qreg_name)
        subcircuits.append(QuantumCircuit(qreg))
        subgraph_wire_maps.append({old: new for old, new in zip(unused_qubits, qreg)})

    bit_path_map = {w: [w] for w in circuit_wires}                        # :229-232
    for path in bit_path_map.values():
        while path[-1] in stitches:
            path.append(stitches[path[-1]])

    subcirc_wire_map = {ext: (idx, sub)
                        for idx, wire_map in enumerate(subgraph_wire_maps)
                        for ext, sub in wire_map.items()}
    wire_path_map = {w: tuple(subcirc_wire_map[x] for x in path)
                     for w, path in bit_path_map.items()}
    return subcircuits, wire_path_map
