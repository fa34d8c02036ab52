"""ORACLE (test infrastructure, not product code): CPU restatement of the reference's
qsplit fragment-tomography -> Choi model -> recombination path.

Follows qsplit/qsplit_mlrecon_methods.py of the reference function by function (file:line cited
on each def) in the exact-probability (infinite-shot) limit: where the reference stores integer
`counts` sampled by qiskit-aer's `qasm_simulator`, this oracle stores the exact outcome
probabilities |amp|^2 of the same tomography circuit (scale cancels: `direct_fragment_model`
is linear and `recombine_fragment_models` normalises).

Third-party arithmetic not under /root/reference and restated here from the published
definitions:
  * qiskit-ignis 0.5.1 `process_tomography_circuits` with `prep_basis="SIC"`:
    prep S0 = nothing, S1 = u3(t, pi, 0), S2 = u3(t, pi/3, 0), S3 = u3(t, -pi/3, 0),
    t = -2 atan(sqrt 2); measurement X: h, Y: sdg then h, Z: nothing; every qubit measured.
  * ignis `pauli_preparation_matrix` / `sicpovm_preparation_matrix`: projectors onto those
    states.
  * tensornetwork 0.4.5 `contractors.greedy` == plain einsum of the connected network.
  * scipy.linalg.lstsq (SVD least squares, cond=1e-8), scipy.linalg.eigh: called directly.

Pinned by: test_cutting.py:7-40 known answer {000: .5, 111: .5}; "recombined == uncut
statevector distribution" on every scenario in tests/; see tests/test_oracle_golden.py.
parity unpinned at the intermediate-data level (the reference holds no golden counts / Choi
blocks).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/reference legs may import this.
"""
from __future__ import annotations

import itertools

import numpy
import scipy.linalg

from . import statevector as sv

prep_state_keys = {"Pauli": ["Zp", "Zm", "Xp", "Yp"],
                   "SIC": ["S0", "S1", "S2", "S3"]}
meas_state_keys = {"Pauli": ["Zp", "Zm", "Xp", "Xm", "Yp", "Ym"],
                   "SIC": ["S0", "S1", "S2", "S3"]}

_SIC_THETA = -2 * numpy.arctan(numpy.sqrt(2))
_SIC_PREP = {"S0": None,
             "S1": ("u3", (_SIC_THETA, numpy.pi, 0.0)),
             "S2": ("u3", (_SIC_THETA, numpy.pi / 3, 0.0)),
             "S3": ("u3", (_SIC_THETA, -numpy.pi / 3, 0.0))}


# ---------------------------------------------------------------------------------------
# wire bookkeeping (qsplit_mlrecon_methods.py:85-113, 681-691)
# ---------------------------------------------------------------------------------------

def get_frag_num(wire_path_map):                       # :85-88
    return len(set(frag_wire[0] for path in wire_path_map.values() for frag_wire in path))


def identify_stitches(wire_path_map):                  # :91-101
    circuit_wires = list(wire_path_map.keys())
    stitches = {}
    for wire in circuit_wires:
        init_wires = wire_path_map[wire][1:]
        exit_wires = wire_path_map[wire][:-1]
        stitches.update({exit_wire: init_wire
                         for init_wire, exit_wire in zip(init_wires, exit_wires)})
    return stitches


def identify_frag_targets(wire_path_map):              # :104-113
    stitches = identify_stitches(wire_path_map)
    frag_targets = [{"meas": tuple(), "prep": tuple()}
                    for _ in range(get_frag_num(wire_path_map))]
    for meas_frag_qubit, prep_frag_qubit in stitches.items():
        meas_frag, meas_qubit = meas_frag_qubit
        prep_frag, prep_qubit = prep_frag_qubit
        frag_targets[meas_frag]["meas"] += (meas_qubit,)
        frag_targets[prep_frag]["prep"] += (prep_qubit,)
    return frag_targets


def fragment_cuts(wire_path_map):                      # :681-687
    cuts = [{"prep": 0, "meas": 0} for _ in range(get_frag_num(wire_path_map))]
    for cut_meas, cut_prep in identify_stitches(wire_path_map).items():
        cuts[cut_meas[0]]["meas"] += 1
        cuts[cut_prep[0]]["prep"] += 1
    return cuts


def fragment_variants(wire_path_map):                  # :689-691
    return sum(4 ** c["prep"] * 3 ** c["meas"] for c in fragment_cuts(wire_path_map))


def _qubit_index(qubit):
    return qubit if type(qubit) is int else qubit.index


def _frag_insts(fragment):
    """(n, plain-tuple instruction list) of a fragment given as an object with
    `.num_qubits` / `.to_instructions()` or as an `(n, instructions)` pair."""
    if hasattr(fragment, "to_instructions"):
        return fragment.num_qubits, fragment.to_instructions()
    return fragment


# ---------------------------------------------------------------------------------------
# tomography in the exact limit (qsplit_mlrecon_methods.py:116-232)
# ---------------------------------------------------------------------------------------

def partial_tomography(fragment, prep_qubits, meas_qubits):          # :116-162
    """Exact outcome distribution of every tomography variant of one fragment.

    Returns {(prep_label, meas_label): ndarray(2^n)} with labels over the *cut* qubits only,
    prep_label in {S0..S3}^p (itertools.product order, :148), meas_label in {Z,X,Y}^m (:149).
    Non-cut qubits are prepared in S0 = |0> and measured in Z (:136-146).
    """
    n, insts = _frag_insts(fragment)
    prep_qubits = list(map(_qubit_index, prep_qubits or []))
    meas_qubits = list(map(_qubit_index, meas_qubits or []))
    prep_labels = list(itertools.product(prep_state_keys["SIC"], repeat=len(prep_qubits)))
    meas_labels = list(itertools.product(["Z", "X", "Y"], repeat=len(meas_qubits)))
    out = {}
    for prep_label in prep_labels:
        pre = []
        for q, lab in zip(prep_qubits, prep_label):
            if _SIC_PREP[lab] is not None:
                name, params = _SIC_PREP[lab]
                pre.append((name, (q,), params, None))
        state = sv.simulate(n, pre + list(insts))
        for meas_label in meas_labels:
            post = []
            for q, lab in zip(meas_qubits, meas_label):
                if lab == "X":
                    post.append(("h", (q,), (), None))
                elif lab == "Y":
                    post.append(("sdg", (q,), (), None))
                    post.append(("h", (q,), (), None))
            rotated = state.copy()
            for inst in post:
                rotated = sv.apply_instruction(rotated, n, inst)
            out[(prep_label, meas_label)] = sv.probabilities(rotated)
    return out


def organize_tomography_data(variant_probs, n, prep_qubits, meas_qubits,
                             prep_basis="SIC", drop_below=0.0):       # :169-213
    """{final bits: {(prep_label, meas_state): prob}}, built like the reference builds
    {final bits: {(prep_label, meas_state): counts}}.  Outcomes with probability
    <= `drop_below` are treated as never observed (an Aer counts dict has no zero entries)."""
    prep_qubits = list(map(_qubit_index, prep_qubits or []))
    meas_qubits = list(map(_qubit_index, meas_qubits or []))

    def _split_bits(bits):                                            # :178-183
        mid_bits = "".join([bits[len(bits) - idx - 1] for idx in meas_qubits])
        fin_bits = "".join([bit for pos, bit in enumerate(bits)
                            if len(bits) - pos - 1 not in meas_qubits])
        return mid_bits, fin_bits

    organized_data = {}
    for (prep_label, meas_label), probs in variant_probs.items():
        for x in range(1 << n):
            counts = float(probs[x])
            if counts <= drop_below:
                continue
            bits = format(x, f"0{n}b") if n > 0 else ""
            meas_bits, final_bits = _split_bits(bits)
            meas_state = tuple(basis + ("p" if outcome == "0" else "m")
                               for basis, outcome in zip(meas_label, meas_bits))
            count_label = (prep_label, meas_state)
            if final_bits not in organized_data:
                organized_data[final_bits] = {}
            organized_data[final_bits][count_label] = counts

    prep_labels = itertools.product(prep_state_keys[prep_basis], repeat=len(prep_qubits))
    meas_states = itertools.product(meas_state_keys["Pauli"], repeat=len(meas_qubits))
    count_labels = list(itertools.product(prep_labels, meas_states))
    for bits in organized_data.keys():                                # :204-211
        if len(organized_data[bits]) == len(count_labels):
            continue
        for count_label in count_labels:
            if count_label not in organized_data[bits]:
                organized_data[bits][count_label] = 0
    return organized_data


def collect_fragment_data(fragments, wire_path_map, drop_below=0.0):  # :216-232
    frag_targets = identify_frag_targets(wire_path_map)
    data = []
    for idx, fragment in enumerate(fragments):
        n, _ = _frag_insts(fragment)
        raw = partial_tomography(fragment, frag_targets[idx].get("prep"),
                                 frag_targets[idx].get("meas"))
        data.append(organize_tomography_data(raw, n, frag_targets[idx].get("prep"),
                                             frag_targets[idx].get("meas"),
                                             drop_below=drop_below))
    return data


# ---------------------------------------------------------------------------------------
# Choi models (qsplit_mlrecon_methods.py:239-356)
# ---------------------------------------------------------------------------------------

def _ket(label):
    """State vector prepared by the ignis circuit for `label`."""
    if label in ("S0", "Zp"):
        return numpy.array([1, 0], dtype=complex)
    if label == "Zm":
        return numpy.array([0, 1], dtype=complex)
    if label == "Xp":
        return numpy.array([1, 1], dtype=complex) / numpy.sqrt(2)
    if label == "Xm":
        return numpy.array([1, -1], dtype=complex) / numpy.sqrt(2)
    if label == "Yp":
        return numpy.array([1, 1j], dtype=complex) / numpy.sqrt(2)
    if label == "Ym":
        return numpy.array([1, -1j], dtype=complex) / numpy.sqrt(2)
    if label in ("S1", "S2", "S3"):
        return sv.u3_matrix(*_SIC_PREP[label][1])[:, 0]
    raise ValueError(f"label not recognized: {label}")


def label_to_matrix(label):                                           # :239-252
    if label == "I":
        return numpy.eye(2)
    if label[0] in ["X", "Y", "Z"]:
        if len(label) == 1:
            return label_to_matrix(label + "p") - label_to_matrix(label + "m")
        ket = _ket(label)
        return numpy.outer(ket, ket.conj())
    if label[0] == "S":
        ket = _ket(label)
        return numpy.outer(ket, ket.conj())
    raise ValueError(f"label not recognized: {label}")


def target_labels_to_matrix(targets):                                 # :255-266
    try:
        prep_labels, meas_labels = targets["prep"], targets["meas"]
    except (TypeError, KeyError, IndexError):
        prep_labels, meas_labels = targets
    prep_matrix = numpy.array(1)
    meas_matrix = numpy.array(1)
    for label in prep_labels:
        prep_matrix = numpy.kron(label_to_matrix(label), prep_matrix)
    for label in meas_labels:
        meas_matrix = numpy.kron(label_to_matrix(label), meas_matrix)
    return numpy.kron(prep_matrix.T, meas_matrix)


def direct_fragment_model(tomography_data, discard_poor_data=False, rank_cutoff=1e-8):  # :273-317
    if type(tomography_data) is list:
        return [direct_fragment_model(d, discard_poor_data, rank_cutoff)
                for d in tomography_data]
    choi_matrix = {}
    for final_bits, fixed_bit_data in tomography_data.items():
        prep_meas_states, state_counts = zip(*fixed_bit_data.items())
        prep_labels, meas_labels = zip(*prep_meas_states)
        prep_qubit_num = len(prep_labels[0])
        meas_qubit_num = len(meas_labels[0])
        if discard_poor_data:
            if len(fixed_bit_data) < 4 ** (prep_qubit_num + meas_qubit_num):
                continue
        cut_qubit_num = prep_qubit_num + meas_qubit_num
        choi_trace = sum(state_counts) / (2 ** prep_qubit_num * 3 ** meas_qubit_num)
        state_matrix = numpy.array([target_labels_to_matrix(states).flatten()
                                    for states in prep_meas_states] +
                                   [numpy.eye(2 ** cut_qubit_num).flatten()])
        state_counts = numpy.array(list(state_counts) + [choi_trace])
        choi_fit = scipy.linalg.lstsq(state_matrix.conj(), state_counts,
                                      cond=rank_cutoff)[0]
        choi_matrix[final_bits] = choi_fit.reshape(2 ** cut_qubit_num, 2 ** cut_qubit_num)
    return choi_matrix


def to_projector(vector):                                             # :45-46
    return numpy.outer(numpy.conjugate(vector), vector)


def maximum_likelihood_model(choi_matrix):                            # :320-356
    if type(choi_matrix) is list:
        return [maximum_likelihood_model(mat) for mat in choi_matrix]
    choi_eigs, choi_vecs = {}, {}
    for final_bits, choi_block in choi_matrix.items():
        choi_eigs[final_bits], choi_vecs[final_bits] = scipy.linalg.eigh(choi_block)
    all_eigs = numpy.concatenate(list(choi_eigs.values()))
    eig_order = numpy.argsort(all_eigs)
    sorted_eigs = all_eigs[eig_order]
    dim = len(sorted_eigs)
    for idx in range(dim):
        val = sorted_eigs[idx]
        if val >= 0:
            break
        sorted_eigs[idx] = 0
        sorted_eigs[idx + 1:] += val / (dim - (idx + 1))
    reverse_order = numpy.arange(dim)[numpy.argsort(eig_order)]
    all_eigs = sorted_eigs[reverse_order]
    num_blocks = len(choi_eigs)
    block_size = dim // num_blocks
    all_eigs = numpy.reshape(all_eigs, (num_blocks, block_size))
    choi_eigs = {bits: all_eigs[idx, :]
                 for idx, bits in enumerate(choi_eigs.keys())
                 if not numpy.count_nonzero(all_eigs[idx, :]) == 0}
    # NB the reference's to_projector(v) = outer(conj(v), v) = conj(|v><v|); it is applied to
    # eigenvectors of the block as stored, so the rebuilt block is the conjugate of the PSD
    # projection (qsplit_mlrecon_methods.py:45-46, 354-356).  Restated as written.
    return {bits: sum(val * to_projector(choi_vecs[bits][:, idx])
                      for idx, val in enumerate(vals) if val > 0)
            for bits, vals in choi_eigs.items()}


# ---------------------------------------------------------------------------------------
# bit permutation of the united distribution (qsplit_mlrecon_methods.py:483-536)
# ---------------------------------------------------------------------------------------

def _identify_init_exit_wires(wire_path_map, num_fragments):          # :483-498
    all_init_wires, all_exit_wires = set(), set()
    for path in wire_path_map.values():
        all_init_wires.update(path[1:])
        all_exit_wires.update(path[:-1])
    init_wires = tuple([{wire for idx, wire in all_init_wires if idx == frag_idx}
                        for frag_idx in range(num_fragments)])
    exit_wires = tuple([{wire for idx, wire in all_exit_wires if idx == frag_idx}
                        for frag_idx in range(num_fragments)])
    return init_wires, exit_wires


def _united_wire_order(wire_path_map, frag_wires):                    # :501-505
    _, exit_wires = _identify_init_exit_wires(wire_path_map, len(frag_wires))
    return [(frag_idx, wire)
            for frag_idx, wires in enumerate(frag_wires)
            for wire in wires if wire not in exit_wires[frag_idx]]


def _frag_output_wire_map(wire_path_map):                             # :508-509
    return {path[-1]: wire for wire, path in wire_path_map.items()}


def _axis_permutation(old_wire_order, new_wire_order, wire_map=None): # :514-518
    if wire_map is None:
        wire_map = {wire: wire for wire in old_wire_order}
    output_wire_order = [wire_map[wire] for wire in old_wire_order]
    wire_permutation = [output_wire_order.index(wire) for wire in new_wire_order]
    return [len(new_wire_order) - 1 - idx for idx in wire_permutation][::-1]


def _get_all_wires(wire_path_map):                                    # :521-529
    circuit_wires = list(wire_path_map.keys())
    all_frag_wires = set(wire for path in wire_path_map.values() for wire in path)
    frag_num = max(frag_wire[0] for frag_wire in all_frag_wires) + 1
    frag_wires = [sorted([frag_wire[1] for frag_wire in all_frag_wires if frag_wire[0] == idx],
                         key=lambda qq: qq.index)
                  for idx in range(frag_num)]
    return circuit_wires, frag_wires


def united_axis_permutation(wire_path_map):                           # :532-536
    circuit_wires, frag_wires = _get_all_wires(wire_path_map)
    output_wires = _united_wire_order(wire_path_map, frag_wires)
    output_wire_map = _frag_output_wire_map(wire_path_map)
    return _axis_permutation(output_wires, circuit_wires, output_wire_map)


# ---------------------------------------------------------------------------------------
# recombination (qsplit_mlrecon_methods.py:363-472)
# ---------------------------------------------------------------------------------------

def _choi_to_tensor(matrix, prep_qubits):                             # :428-440
    qubits = (len(bin(matrix.shape[0])) - 3)
    tensor = matrix.reshape((2,) * 2 * qubits)
    prep_axes_bra = reversed(range(prep_qubits))
    meas_axes_ket = reversed(range(prep_qubits, qubits))
    prep_axes_ket = reversed(range(qubits, qubits + prep_qubits))
    meas_axes_bra = reversed(range(qubits + prep_qubits, 2 * qubits))
    prep_axes = numpy.array(list(zip(prep_axes_bra, prep_axes_ket)), dtype=int).flatten()
    meas_axes = numpy.array(list(zip(meas_axes_ket, meas_axes_bra)), dtype=int).flatten()
    tensor = tensor.transpose(list(prep_axes) + list(meas_axes))
    return tensor.reshape((4,) * qubits)


def _network_subscripts(wire_path_map):
    """einsum axis labels per fragment implementing the `^` connections of :444-453."""
    stitches = identify_stitches(wire_path_map)
    frag_targets = identify_frag_targets(wire_path_map)
    labels = [[None] * (len(t["prep"]) + len(t["meas"])) for t in frag_targets]
    for lab, (meas_frag_qubit, prep_frag_qubit) in enumerate(stitches.items()):
        meas_frag, meas_qubit = meas_frag_qubit
        prep_frag, prep_qubit = prep_frag_qubit
        meas_qubit_idx = frag_targets[meas_frag]["meas"].index(meas_qubit)
        prep_qubit_idx = frag_targets[prep_frag]["prep"].index(prep_qubit)
        prep_axis = prep_qubit_idx
        meas_axis = len(frag_targets[meas_frag]["prep"]) + meas_qubit_idx
        labels[meas_frag][meas_axis] = lab
        labels[prep_frag][prep_axis] = lab
    return labels


def _recombine_using_networks(frag_models, wire_path_map):            # :412-461
    frag_num = len(frag_models)
    frag_targets = identify_frag_targets(wire_path_map)
    final_bit_pieces = [list(choi.keys()) for choi in frag_models]
    bit_permutation = united_axis_permutation(wire_path_map)
    labels = _network_subscripts(wire_path_map)

    combined_dist = {}
    for frag_bits in itertools.product(*final_bit_pieces):
        joined_bits = "".join(frag_bits[::-1])
        final_bits = "".join([joined_bits[idx] for idx in bit_permutation])
        operands = []
        for choi, bits, idx in zip(frag_models, frag_bits, range(frag_num)):
            tensor = _choi_to_tensor(choi[bits], len(frag_targets[idx]["prep"]))
            operands += [tensor, labels[idx]]
        val = numpy.einsum(*operands, []).real
        try:
            combined_dist[final_bits] += val
        except KeyError:
            combined_dist[final_bits] = val
    return combined_dist


def _recombine_using_insertions(frag_models, wire_path_map):          # :363-409
    frag_num = len(frag_models)
    stitches = identify_stitches(wire_path_map)
    frag_targets = identify_frag_targets(wire_path_map)
    final_bit_pieces = [list(choi.keys()) for choi in frag_models]
    bit_permutation = united_axis_permutation(wire_path_map)

    combined_dist = {}
    for stitch_ops in itertools.product(["I", "Z", "X", "Y"], repeat=len(stitches)):
        frag_ops = {idx: {"prep": {}, "meas": {}} for idx in range(frag_num)}
        for stitch_op, stitch_qubits in zip(stitch_ops, stitches.items()):
            meas_frag_qubit, prep_frag_qubit = stitch_qubits
            meas_frag, meas_qubit = meas_frag_qubit
            prep_frag, prep_qubit = prep_frag_qubit
            meas_idx = frag_targets[meas_frag]["meas"].index(meas_qubit)
            prep_idx = frag_targets[prep_frag]["prep"].index(prep_qubit)
            frag_ops[meas_frag]["meas"][meas_idx] = stitch_op
            frag_ops[prep_frag]["prep"][prep_idx] = stitch_op
        frag_ops = [frag_ops[idx] for idx in range(frag_num)]
        frag_labels = [{"prep": tuple(o["prep"][i] for i in range(len(o["prep"]))),
                        "meas": tuple(o["meas"][i] for i in range(len(o["meas"])))}
                       for o in frag_ops]
        frag_mats = list(map(target_labels_to_matrix, frag_labels))
        for frag_bits in itertools.product(*final_bit_pieces):
            joined_bits = "".join(frag_bits[::-1])
            final_bits = "".join([joined_bits[idx] for idx in bit_permutation])
            frag_vals = [mat.flatten().conj() @ choi[bits].flatten()
                         for choi, bits, mat in zip(frag_models, frag_bits, frag_mats)]
            val = numpy.prod(frag_vals).real
            try:
                combined_dist[final_bits] += val
            except KeyError:
                combined_dist[final_bits] = val
    return combined_dist


def recombine_fragment_models(*args, method="network", **kwargs):     # :463-472
    if method == "network":
        recombination_method = _recombine_using_networks
    elif method == "insertion":
        recombination_method = _recombine_using_insertions
    else:
        raise ValueError("recombination method {method} not recognized")
    combined_dist = recombination_method(*args, **kwargs)
    combined_norm = sum(combined_dist.values())
    return {bits: val / combined_norm for bits, val in combined_dist.items()}


def sim_with_cutting(fragments, wire_path_map, mode="likely", drop_below=0.0):
    """utils/cutting_funcs.py:167-205 in the exact limit."""
    frag_data = collect_fragment_data(fragments, wire_path_map, drop_below=drop_below)
    direct_models = direct_fragment_model(frag_data)
    if mode == "direct":
        models = direct_models
    elif mode == "likely":
        models = maximum_likelihood_model(direct_models)
    else:
        raise Exception('Unknown recombination mode:', mode)
    return recombine_fragment_models(models, wire_path_map)


# ---------------------------------------------------------------------------------------
# array ("GEMM") form of the same contraction, for sizes where the per-bitstring Python loop
# of :422 is infeasible.  Same numbers as above, float64.
# ---------------------------------------------------------------------------------------

PAULI = {"I": numpy.eye(2, dtype=complex),
         "X": numpy.array([[0, 1], [1, 0]], dtype=complex),
         "Y": numpy.array([[0, -1j], [1j, 0]], dtype=complex),
         "Z": numpy.diag([1.0 + 0j, -1.0])}
PAULI_ORDER = "IXYZ"


def sic_frame_coefficients():
    """c[P, s] with Pauli_P = sum_s c[P, s] |S_s><S_s|  (real 4x4)."""
    basis = numpy.array([label_to_matrix(f"S{s}").flatten() for s in range(4)]).T  # [4, s]
    out = numpy.zeros((4, 4))
    for i, p in enumerate(PAULI_ORDER):
        sol = numpy.linalg.solve(basis, PAULI[p].flatten())
        assert numpy.allclose(sol.imag, 0, atol=1e-12)
        out[i] = sol.real
    return out


def fragment_pauli_tensor(fragment, prep_qubits, meas_qubits):
    """Real tensor R[b, a_1..a_q] (a in I,X,Y,Z; prep axes then meas axes) of one fragment:
    R[b, (Pp),(Pm)] = sum_{mu} sign(Pm, mu) * L(b, mu | rho = Pp, basis(Pm)), i.e. the
    Pauli-basis coordinates of the Choi block Lambda_b.  b indexes the fragment's final bits
    as an integer (bit k of b <-> k-th non-meas qubit in ascending local order)."""
    n, insts = _frag_insts(fragment)
    prep_qubits = list(map(_qubit_index, prep_qubits or []))
    meas_qubits = list(map(_qubit_index, meas_qubits or []))
    p, m = len(prep_qubits), len(meas_qubits)
    raw = partial_tomography(fragment, prep_qubits, meas_qubits)
    out_qubits = [q for q in range(n) if q not in meas_qubits]
    o = len(out_qubits)
    idx = numpy.arange(1 << n)
    b_of_x = numpy.zeros(1 << n, dtype=numpy.int64)
    for k, q in enumerate(out_qubits):
        b_of_x |= ((idx >> q) & 1) << k
    coef = sic_frame_coefficients()
    R = numpy.zeros((1 << o,) + (4,) * (p + m))
    for pp in itertools.product(range(4), repeat=p):
        for pm in itertools.product(range(4), repeat=m):
            meas_label = tuple("Z" if PAULI_ORDER[a] in "IZ" else PAULI_ORDER[a] for a in pm)
            sign = numpy.ones(1 << n)
            for q, a in zip(meas_qubits, pm):
                if PAULI_ORDER[a] != "I":
                    sign = sign * (1 - 2 * ((idx >> q) & 1))
            acc = numpy.zeros(1 << n)
            for ss in itertools.product(range(4), repeat=p):
                c = 1.0
                for a, s in zip(pp, ss):
                    c *= coef[a, s]
                if c == 0.0:
                    continue
                prep_label = tuple(f"S{s}" for s in ss)
                acc += c * raw[(prep_label, meas_label)]
            R[(slice(None),) + pp + pm] = numpy.bincount(b_of_x, weights=acc * sign,
                                                         minlength=1 << o)
    return R


def recombine_arrays(pauli_tensors, wire_path_map):
    """P~ over the united index (fragment F-1 most significant ... fragment 0 least), i.e.
    2^-K * einsum over stitched Pauli axes, then `united_axis_permutation` applied to bits.
    Returns ndarray(2^N) indexed by the full-circuit integer (bit i <-> circuit wire i),
    normalised like :471-472."""
    labels = _network_subscripts(wire_path_map)
    K = len(identify_stitches(wire_path_map))
    F = len(pauli_tensors)
    operands, out_labels = [], []
    for f in reversed(range(F)):
        blab = 40 + f
        operands += [pauli_tensors[f], [blab] + labels[f]]
        out_labels.append(blab)
    united = numpy.einsum(*operands, out_labels) / 2 ** K
    widths = [int(numpy.log2(t.shape[0])) for t in pauli_tensors]
    N = sum(widths)
    united = united.reshape(-1)
    perm = united_axis_permutation(wire_path_map)
    # key char j of final = joined char perm[j]; joined char c <-> united bit N-1-c;
    # final char j <-> circuit bit N-1-j
    src_bit_of_dst = [N - 1 - perm[N - 1 - i] for i in range(N)]   # circuit bit i <- united bit
    tensor = united.reshape((2,) * N)                               # axis a <-> united bit N-1-a
    axes = [N - 1 - src_bit_of_dst[N - 1 - a] for a in range(N)]    # out axis a <-> circuit bit N-1-a
    full = numpy.ascontiguousarray(tensor.transpose(axes)).reshape(-1)
    return full / full.sum()
