"""Fragment tomography -> fragment models -> recombination, on the B200.

Same names, argument meaning and error behaviour as the reference's
qsplit/qsplit_mlrecon_methods.py (file:line cited per function); what changes is where the
numbers are made:

* `collect_fragment_data`   (:216-232)  the 4^p preparation variants of every fragment are ONE
  batched launch sequence of the statevector engine (csrc/statevec.cu) instead of 4^p * 3^m
  separately assembled qiskit circuits; the 3^m measurement variants are never simulated -- all
  of them are functions of the same final state.
* `direct_fragment_model`   (:273-317)  the per-bitstring least-squares inversion is, for exact
  (infinite-shot) data, a fixed linear map: the kernel `qdc_sv_pauli_tensor` produces the Choi
  blocks in the Pauli basis straight from the amplitudes.
* `maximum_likelihood_model`(:320-356)  per-block eigen-decomposition + the global eigenvalue
  clipping as this package's own kernels (csrc/ml_project.cu: Cholesky PSD test, one-warp Jacobi,
  soft-threshold by Newton steps, rebuild); a sign flip per tensor component on exact data.
* `recombine_fragment_models`(:463-472) the `itertools.product` loop over all fragment
  bitstrings becomes one FP64 tensor-core contraction over the cut-basis index with the
  bit permutation of `united_axis_permutation` (:532-536) folded into the scoring tables
  (csrc/contract.cu, csrc/mis_reduce.cu).

Objects returned are `Mapping`s with the reference's dict layouts (materialised lazily, for the
sizes where a Python dict is sensible) that keep their device tensors for the next stage.
The reference samples `shots` outcomes per variant; by default this engine evaluates the
infinite-shot limit (`shots` only scales the numbers shown by the dict views);
`collect_fragment_data(sample=True)` draws the counts on the device (csrc/shot_sampler.cu).
"""
from __future__ import annotations

import itertools
import os
from collections.abc import Mapping
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from .. import _lib
from .. import recombine as _rc
from ..engine import B200Backend, BASIS_KETS, SIC_FRAME, SIC_KETS

__all__ = ["get_frag_num", "identify_stitches", "identify_frag_targets", "fragment_cuts",
           "fragment_variants", "united_axis_permutation", "united_bit_wires",
           "collect_fragment_data", "direct_fragment_model", "maximum_likelihood_model",
           "recombine_fragment_models", "FragmentData", "FragmentModel", "DeviceDistribution",
           "label_to_matrix", "target_labels_to_matrix"]

prep_state_keys = {"Pauli": ["Zp", "Zm", "Xp", "Yp"], "SIC": ["S0", "S1", "S2", "S3"]}
meas_state_keys = {"Pauli": ["Zp", "Zm", "Xp", "Xm", "Yp", "Ym"], "SIC": ["S0", "S1", "S2", "S3"]}


# ---------------------------------------------------------------------------------------------
# wire bookkeeping: pure integer / object-identity logic, same results as the reference
# ---------------------------------------------------------------------------------------------

def get_frag_num(wire_path_map):                                  # :85-88
    return len(set(fw[0] for path in wire_path_map.values() for fw in path))


def identify_stitches(wire_path_map):                             # :91-101
    """{(exit fragment, exit wire): (init fragment, init wire)} in wire_path_map key order."""
    stitches = {}
    for wire in wire_path_map.keys():
        path = wire_path_map[wire]
        stitches.update({exit_wire: init_wire
                         for init_wire, exit_wire in zip(path[1:], path[:-1])})
    return stitches


def identify_frag_targets(wire_path_map):                         # :104-113
    frag_targets = [{"meas": tuple(), "prep": tuple()} for _ in range(get_frag_num(wire_path_map))]
    for (meas_frag, meas_qubit), (prep_frag, prep_qubit) in identify_stitches(wire_path_map).items():
        frag_targets[meas_frag]["meas"] += (meas_qubit,)
        frag_targets[prep_frag]["prep"] += (prep_qubit,)
    return frag_targets


def fragment_cuts(wire_path_map):                                 # :681-687
    cuts = [{"prep": 0, "meas": 0} for _ in range(get_frag_num(wire_path_map))]
    for cut_meas, cut_prep in identify_stitches(wire_path_map).items():
        cuts[cut_meas[0]]["meas"] += 1
        cuts[cut_prep[0]]["prep"] += 1
    return cuts


def fragment_variants(wire_path_map):                             # :689-691
    return sum(4 ** c["prep"] * 3 ** c["meas"] for c in fragment_cuts(wire_path_map))


def _qidx(q):
    return q if isinstance(q, int) else q.index


def _get_all_wires(wire_path_map):                                # :521-529
    circuit_wires = list(wire_path_map.keys())
    all_frag_wires = set(w for path in wire_path_map.values() for w in path)
    frag_num = max(fw[0] for fw in all_frag_wires) + 1
    frag_wires = [sorted([fw[1] for fw in all_frag_wires if fw[0] == idx], key=_qidx)
                  for idx in range(frag_num)]
    return circuit_wires, frag_wires


def united_axis_permutation(wire_path_map):                       # :532-536 (+ :483-518)
    """Permutation taking the joined fragment bitstring (fragment F-1 first ... fragment 0 last,
    exit wires removed) to the full-circuit bitstring, as a list of source character indices."""
    circuit_wires, frag_wires = _get_all_wires(wire_path_map)
    exit_wires = [set() for _ in frag_wires]
    for path in wire_path_map.values():
        for idx, wire in path[:-1]:
            exit_wires[idx].add(wire)
    output_wires = [(f, w) for f, wires in enumerate(frag_wires) for w in wires
                    if w not in exit_wires[f]]                     # :501-505
    output_wire_map = {path[-1]: wire for wire, path in wire_path_map.items()}   # :508-509
    output_wire_order = [output_wire_map[w] for w in output_wires]               # :514-518
    wire_permutation = [output_wire_order.index(w) for w in circuit_wires]
    n = len(circuit_wires)
    return [n - 1 - idx for idx in wire_permutation][::-1]


def united_bit_wires(wire_path_map):
    """bitpos[u] = index (in wire_path_map key order = full-circuit bit) of the circuit wire
    carried by bit u of the united integer  sum_f b_f << (width of fragments < f),  where bit k of
    b_f is the k-th non-exit wire of fragment f in ascending wire index.  This is
    `united_axis_permutation` re-expressed on integers (checked against it in the tests)."""
    circuit_wires, frag_wires = _get_all_wires(wire_path_map)
    exit_wires = [set() for _ in frag_wires]
    for path in wire_path_map.values():
        for idx, wire in path[:-1]:
            exit_wires[idx].add(wire)
    output_wire_map = {path[-1]: wire for wire, path in wire_path_map.items()}
    pos = {w: i for i, w in enumerate(circuit_wires)}
    bitpos, widths = [], []
    for f, wires in enumerate(frag_wires):
        outs = [w for w in wires if w not in exit_wires[f]]
        widths.append(len(outs))
        bitpos += [pos[output_wire_map[(f, w)]] for w in outs]
    return bitpos, widths


# ---------------------------------------------------------------------------------------------
# label matrices (:239-266) and the Pauli <-> Choi change of basis
# ---------------------------------------------------------------------------------------------
_PAULI = {"I": np.eye(2, dtype=complex), "X": np.array([[0, 1], [1, 0]], dtype=complex),
          "Y": np.array([[0, -1j], [1j, 0]]), "Z": np.diag([1.0 + 0j, -1.0])}
_PAULI_ORDER = "IXYZ"                       # digit order of the kernel's kappa index


def _ket(label):
    if label in ("S0", "Zp"):
        return np.array([1, 0], dtype=complex)
    if label == "Zm":
        return np.array([0, 1], dtype=complex)
    if label in ("Xp", "Xm"):
        return np.array([1, 1 if label == "Xp" else -1], dtype=complex) / np.sqrt(2)
    if label in ("Yp", "Ym"):
        return np.array([1, 1j if label == "Yp" else -1j], dtype=complex) / np.sqrt(2)
    if label in ("S1", "S2", "S3"):
        return SIC_KETS[int(label[1])]
    raise ValueError(f"label not recognized: {label}")


def label_to_matrix(label):                                       # :239-252
    if label == "I":
        return np.eye(2)
    if label in ("X", "Y", "Z"):
        return label_to_matrix(label + "p") - label_to_matrix(label + "m")
    ket = _ket(label)
    return np.outer(ket, ket.conj())


def target_labels_to_matrix(targets):                             # :255-266
    try:
        prep_labels, meas_labels = targets["prep"], targets["meas"]
    except (TypeError, KeyError, IndexError):
        prep_labels, meas_labels = targets
    prep_matrix, meas_matrix = np.array(1), np.array(1)
    for label in prep_labels:
        prep_matrix = np.kron(label_to_matrix(label), prep_matrix)
    for label in meas_labels:
        meas_matrix = np.kron(label_to_matrix(label), meas_matrix)
    return np.kron(prep_matrix.T, meas_matrix)


_BASIS_CACHE: Dict = {}


def _pauli_choi_basis(n_prep, n_meas):
    """B[kappa] = P_prep^T (x) P_meas for kappa = (prep digits..., meas digits...) row-major,
    complex [4^q, 2^q, 2^q]; Tr(B[k] B[k']) = 2^q delta, and the kernel's tensor entry is
    T[b, kappa] = Tr(B[kappa] Lambda_b)  (the reference's `counts = conj(state_matrix) @ choi`,
    :301-312, extended linearly from projectors to Paulis)."""
    key = (n_prep, n_meas)
    if key not in _BASIS_CACHE:
        mats = []
        for digits in itertools.product(range(4), repeat=n_prep + n_meas):
            labels = [_PAULI_ORDER[d] for d in digits]
            mats.append(target_labels_to_matrix((labels[:n_prep], labels[n_prep:])))
        dim = 2 ** (n_prep + n_meas)
        _BASIS_CACHE[key] = np.array(mats, dtype=complex).reshape(-1, dim, dim)
    return _BASIS_CACHE[key]


def _choi_from_tensor(T: torch.Tensor, n_prep, n_meas) -> torch.Tensor:
    """Lambda[b] = 2^-q sum_kappa T[b, kappa] B[kappa]   ([n_b, 2^q, 2^q] complex128)."""
    q = n_prep + n_meas
    B = torch.as_tensor(_pauli_choi_basis(n_prep, n_meas), device=T.device)
    return torch.einsum("bk,kij->bij", T.to(torch.complex128), B) / (2 ** q)


def _tensor_from_choi(L: torch.Tensor, n_prep, n_meas) -> torch.Tensor:
    """T[b, kappa] = Re Tr(B[kappa] Lambda[b])."""
    B = torch.as_tensor(_pauli_choi_basis(n_prep, n_meas), device=L.device)
    return torch.einsum("kij,bji->bk", B, L).real.contiguous()


def _bits(value, width):
    return format(value, f"0{width}b") if width > 0 else ""


# ---------------------------------------------------------------------------------------------
# collect_fragment_data (:116-232)
# ---------------------------------------------------------------------------------------------
def _frag_insts(fragment):
    if hasattr(fragment, "to_instructions"):
        return fragment.num_qubits, fragment.to_instructions()
    n, insts = fragment
    return int(n), list(insts)


_MEAS_SUFFIX = {"Z": (), "X": ("h",), "Y": ("sdg", "h")}          # ignis measurement circuits


class FragmentData(Mapping):
    """Tomography data of ONE fragment: the final states of its 4^p preparation variants
    (x n_sets parameter vectors), resident on the device -- or, when the fragment has at most 4
    cut ends, of the 2^p computational-basis inputs those variants are linear combinations of
    (`basis_inputs`): the exact-limit Choi blocks are the same, from 2^p simulations.

    As a Mapping it is the reference's `organize_tomography_data` layout (:169-213),
    `{final_bits: {(prep_label, meas_state): counts}}` for parameter set 0, with
    counts = shots * probability; building that view simulates the 3^m measurement variants
    too, so it is meant for fragments small enough for a Python dict."""

    def __init__(self, backend, fragment, prep_qubits, meas_qubits, states, prog, auto,
                 n_sets, shots, cut_labels, params, param_order, prep_basis="SIC",
                 basis_inputs=False):
        self.backend, self.fragment = backend, fragment
        self.prep_qubits, self.meas_qubits = tuple(prep_qubits), tuple(meas_qubits)
        self.states, self.prog, self.auto = states, prog, auto
        self.n_sets, self.shots = n_sets, shots
        self.cut_labels = cut_labels          # {"prep": [stitch numbers], "meas": [...]}
        self.params, self.param_order, self.prep_basis = params, param_order, prep_basis
        self.basis_inputs = basis_inputs      # states = 2^p basis-input evolutions, not 4^p SIC variants
        self._dict = None

    @property
    def n_qubits(self):
        return self.prog.n_qubits

    def variant_probabilities(self, set_index=0):
        """{(prep_label, meas_label): ndarray(2^n)} -- `partial_tomography` (:116-162) in the
        exact limit, every one of the 4^p * 3^m circuits simulated on the device."""
        n, insts = _frag_insts(self.fragment)
        p, m = len(self.prep_qubits), len(self.meas_qubits)
        out = {}
        prep_labels = list(itertools.product(prep_state_keys["SIC"], repeat=p))
        zero_mask = ((1 << n) - 1) & ~sum(1 << q for q in self.prep_qubits)
        for meas_label in itertools.product(["Z", "X", "Y"], repeat=m):
            suffix = [(g, (q,), (), None) for q, lab in zip(self.meas_qubits, meas_label)
                      for g in _MEAS_SUFFIX[lab]]
            prog, auto = self.backend.compile_instructions(n, insts + suffix, self.param_order,
                                                           zero_mask)
            states = self.backend.run(prog, auto, self.params, prep_qubits=self.prep_qubits)
            probs = self.backend.probabilities(states, prog).reshape(len(prep_labels), -1, 1 << n)
            probs = probs[:, set_index].cpu().numpy()
            for i, prep_label in enumerate(prep_labels):
                out[(prep_label, meas_label)] = probs[i]
        return out

    def _materialise(self):
        if self._dict is not None:
            return self._dict
        n = self.n_qubits
        meas = list(self.meas_qubits)
        organized = {}
        for (prep_label, meas_label), probs in self.variant_probabilities().items():
            for x in np.nonzero(probs > 0.0)[0]:
                bits = _bits(int(x), n)
                mid = "".join(bits[n - q - 1] for q in meas)                     # :179-183
                fin = "".join(b for pos, b in enumerate(bits) if n - pos - 1 not in meas)
                state = tuple(basis + ("p" if o == "0" else "m") for basis, o in zip(meas_label, mid))
                organized.setdefault(fin, {})[(prep_label, state)] = float(probs[x]) * self.shots
        prep_labels = list(itertools.product(prep_state_keys[self.prep_basis],
                                             repeat=len(self.prep_qubits)))
        meas_states = list(itertools.product(meas_state_keys["Pauli"], repeat=len(meas)))
        labels = list(itertools.product(prep_labels, meas_states))
        for fin in organized:                                                    # :204-211
            if len(organized[fin]) != len(labels):
                for lab in labels:
                    organized[fin].setdefault(lab, 0)
        self._dict = organized
        return organized

    def __getitem__(self, key):
        return self._materialise()[key]

    def __iter__(self):
        return iter(self._materialise())

    def __len__(self):
        return len(self._materialise())


def collect_fragment_data(fragments, wire_path_map, shots=1, tomography_backend=None,
                          prep_basis="SIC", monitor_jobs=False, params=None, param_order=None,
                          linear_inputs=True, sample=False, seed=None):
    """Simulate every preparation variant of every fragment (:216-232).

    `tomography_backend` must be a `B200Backend` (there is no other simulator here and no CPU
    fallback).  Extensions: `params` [n_sets, n_params] + `param_order` evaluate a batch of
    parameter vectors of unbound fragments at once; `linear_inputs` (default) simulates the 2^p
    basis inputs instead of the 4^p SIC preparations when a fragment has at most 4 cut ends
    (same Choi blocks by linearity; the dict view still simulates every variant); `sample=True`
    draws `shots` outcomes per variant circuit (x every parameter vector) on the device instead of
    taking the infinite-shot limit, `seed` makes the draws reproducible."""
    if not isinstance(tomography_backend, B200Backend):
        raise _lib.QdcError("collect_fragment_data needs a qdc_b200.engine.B200Backend as "
                            "tomography_backend; there is no CPU simulator in this package")
    if prep_basis != "SIC":
        raise _lib.QdcError("only the reference's default prep_basis='SIC' is implemented")
    be = tomography_backend
    frag_targets = identify_frag_targets(wire_path_map)
    stitches = list(identify_stitches(wire_path_map).items())
    if param_order is None:
        seen = []
        for frag in fragments:
            for prm in getattr(frag, "parameters", []):
                if prm not in seen:
                    seen.append(prm)
        from ..circuit import _param_sort_key
        param_order = sorted(seen, key=_param_sort_key)
    data = []
    for idx, fragment in enumerate(fragments):
        n, insts = _frag_insts(fragment)
        prep = [_qidx(q) for q in frag_targets[idx]["prep"]]
        meas = [_qidx(q) for q in frag_targets[idx]["meas"]]
        labels = {"prep": [k for q in frag_targets[idx]["prep"]
                           for k, (_, init) in enumerate(stitches) if init == (idx, q)],
                  "meas": [k for q in frag_targets[idx]["meas"]
                           for k, (exit_, _) in enumerate(stitches) if exit_ == (idx, q)]}
        zero_mask = ((1 << n) - 1) & ~sum(1 << q for q in prep)
        prog, auto = be.compile_instructions(n, insts, param_order, zero_mask)
        if sample:
            data.append(SampledFragmentData(be, fragment, prep, meas, None, prog, auto, 1, shots,
                                            labels, params, param_order, prep_basis,
                                            seed=None if seed is None else seed + idx))
            continue
        basis = bool(prep) and len(prep) + len(meas) <= 4 and linear_inputs
        states = be.run(prog, auto, params, prep_qubits=prep, kets=BASIS_KETS if basis else SIC_KETS)
        n_sets = states.shape[0] // ((2 if basis else 4) ** len(prep))
        data.append(FragmentData(be, fragment, prep, meas, states, prog, auto, n_sets, shots,
                                 labels, params, param_order, prep_basis, basis))
    return data


# ---------------------------------------------------------------------------------------------
# fragment models (:273-356)
# ---------------------------------------------------------------------------------------------
class FragmentModel(Mapping):
    """Choi blocks of one fragment in the Pauli basis: T[set, b, kappa] float64 on the device,
    kappa = (prep axes..., meas axes...) row-major, b = value of the non-measured qubits.
    As a Mapping: the reference's `{final_bits: Lambda (2^q x 2^q complex)}` for set 0, blocks
    that are identically zero dropped (an outcome never observed has no key, :196-201, 350-352)."""

    def __init__(self, T, n_prep, n_meas, cut_labels, scale=1.0):
        self.T, self.n_prep, self.n_meas = T, n_prep, n_meas
        self.cut_labels = cut_labels
        self.scale = scale                   # dict view multiplies by this (shots)
        self._dict = None

    @property
    def n_sets(self):
        return self.T.shape[0]

    @property
    def out_bits(self):
        return int(self.T.shape[1]).bit_length() - 1

    def choi_blocks(self, set_index=0) -> torch.Tensor:
        return _choi_from_tensor(self.T[set_index], self.n_prep, self.n_meas) * self.scale

    def _materialise(self):
        if self._dict is None:
            blocks = self.choi_blocks().cpu().numpy()
            keep = np.nonzero(np.abs(blocks).reshape(blocks.shape[0], -1).max(axis=1) > 0)[0]
            w = self.out_bits
            self._dict = {_bits(int(b), w): blocks[b] for b in keep}
        return self._dict

    def __getitem__(self, key):
        return self._materialise()[key]

    def __iter__(self):
        return iter(self._materialise())

    def __len__(self):
        return len(self._materialise())


_FIT_CACHE: Dict = {}


def _fit_matrix(n_prep, n_meas, rank_cutoff=1e-8, prep_basis="SIC"):
    """(labels, W): `scipy.linalg.lstsq(state_matrix.conj(), counts, cond)` (:301-312) has the same
    design matrix for every bitstring (labels zero-filled to the full set, :204-211), so the fit
    is counts @ W^T with W = pinv(design) folded with the trace row."""
    key = (n_prep, n_meas, rank_cutoff, prep_basis)
    if key not in _FIT_CACHE:
        prep_labels = list(itertools.product(prep_state_keys[prep_basis], repeat=n_prep))
        meas_states = list(itertools.product(meas_state_keys["Pauli"], repeat=n_meas))
        labels = list(itertools.product(prep_labels, meas_states))
        q = n_prep + n_meas
        design = np.array([target_labels_to_matrix(lab).flatten() for lab in labels]
                          + [np.eye(2 ** q).flatten()]).conj()
        pinv = np.linalg.pinv(design, rcond=rank_cutoff)             # [4^q, n_labels + 1]
        # trace row = sum(counts) / (2^p 3^m)  (:306-309)
        W = pinv[:, :-1] + pinv[:, -1:] / (2 ** n_prep * 3 ** n_meas)
        _FIT_CACHE[key] = (labels, W)
    return _FIT_CACHE[key]


def _model_from_counts(data: Mapping, device, rank_cutoff) -> FragmentModel:
    """Reference-layout counts dict -> model: one GEMM on the device for all bitstrings."""
    keys = list(data.keys())
    first = list(next(iter(data.values())).keys())
    n_prep, n_meas = len(first[0][0]), len(first[0][1])
    q = n_prep + n_meas
    labels, W = _fit_matrix(n_prep, n_meas, rank_cutoff)
    counts = np.array([[data[k].get(lab, 0) for lab in labels] for k in keys], dtype=float)
    width = len(keys[0])
    C_d = torch.zeros(1 << width, len(labels), dtype=torch.complex128, device=device)
    idx = torch.as_tensor([int(k, 2) if width else 0 for k in keys], device=device)
    C_d[idx] = torch.as_tensor(counts, device=device).to(torch.complex128)
    blocks = (C_d @ torch.as_tensor(W, device=device).T).reshape(-1, 2 ** q, 2 ** q)
    T = _tensor_from_choi(blocks, n_prep, n_meas)
    return FragmentModel(T.unsqueeze(0), n_prep, n_meas, None)


class SampledFragmentData(FragmentData):
    """Finite-shot tomography data: every one of the 4^p 3^m variant circuits is simulated (x all
    parameter sets) and `shots` outcomes are drawn from each on the device by this package's own
    sampler (`qdc_sv_sample_counts`: tile masses -> scan -> one warp per shot, Philox counters), which
    files them straight into the organised table - what the reference's qasm_simulator jobs
    followed by `organize_tomography_data` produce (qsplit_mlrecon_methods.py:20-42, 169-232).
    `counts` [n_sets, final bits b, label] is that table as a device tensor; the Mapping view is the
    reference's dict for parameter set 0."""

    def __init__(self, *args, seed=None, **kwargs):
        super().__init__(*args, **kwargs)
        import ctypes as C
        lib = _lib.load()
        be = self.backend
        n, insts = _frag_insts(self.fragment)
        p, m = len(self.prep_qubits), len(self.meas_qubits)
        dev = be.device
        if seed is None:
            seed = int.from_bytes(os.urandom(8), "little")
        labels, _ = _fit_matrix(p, m)
        col = {lab: i for i, lab in enumerate(labels)}
        prep_labels = list(itertools.product(prep_state_keys["SIC"], repeat=p))
        rank = {q: i for i, q in enumerate(sorted(self.meas_qubits))}   # kernel: measured value in ascending qubit order
        meas_mask = sum(1 << q for q in self.meas_qubits)
        zero_mask = ((1 << n) - 1) & ~sum(1 << q for q in self.prep_qubits)
        table = None if self.params is None else np.atleast_2d(np.asarray(self.params, dtype=np.float64))
        n_sets = 1 if table is None else table.shape[0]
        self.n_sets = n_sets
        n_init = len(prep_labels)
        counts = torch.zeros(n_sets, 1 << (n - m), len(labels), dtype=torch.float64, device=dev)
        ws = torch.empty(max(1, int(lib.qdc_sv_sample_workspace_doubles(n, n_init * n_sets))),
                         dtype=torch.float64, device=dev)
        if n_init * n_sets > 65535:
            raise _lib.QdcError("sample=True: at most 65535 (preparation, parameter set) states per basis choice")
        for k, meas_label in enumerate(itertools.product(["Z", "X", "Y"], repeat=m)):
            suffix = [(g, (q,), (), None) for q, lab in zip(self.meas_qubits, meas_label)
                      for g in _MEAS_SUFFIX[lab]]
            prog, auto = be.compile_instructions(n, insts + suffix, self.param_order, zero_mask)
            states = be.run(prog, auto, table, prep_qubits=self.prep_qubits)      # [4^p * n_sets, 2^n]
            cols = np.empty((n_init, 1 << m), dtype=np.int32)
            for i, prep_label in enumerate(prep_labels):
                for v in range(1 << m):
                    state = tuple(basis + ("m" if (v >> rank[q]) & 1 else "p")
                                  for basis, q in zip(meas_label, self.meas_qubits))
                    cols[i, v] = col[(prep_label, state)]
            d_cols = torch.as_tensor(cols, device=dev)
            with torch.cuda.device(dev):
                _lib.check(lib.qdc_sv_sample_counts(_lib.ptr(states), n, n_init, n_sets, C.c_uint64(prog.flip_mask),
                                                    int(self.shots), C.c_uint64(int(seed) & (2 ** 64 - 1)),
                                                    C.c_uint64(k), C.c_uint64(meas_mask), _lib.ptr(d_cols),
                                                    _lib.ptr(counts), counts.shape[2], _lib.ptr(ws),
                                                    _lib.stream_ptr(None)), "qdc_sv_sample_counts")
            del states
        self.counts = counts
        self.labels = labels

    def _materialise(self):
        if self._dict is None:
            n, m = self.n_qubits, len(self.meas_qubits)
            host = self.counts[0].cpu().numpy()
            seen = np.nonzero(host.sum(axis=1) > 0)[0]
            self._dict = {_bits(int(b), n - m): {lab: float(host[b, i]) for i, lab in enumerate(self.labels)}
                          for b in seen}
        return self._dict


def direct_fragment_model(tomography_data, discard_poor_data=False, rank_cutoff=1e-8):
    """Choi blocks by linear inversion (:273-317)."""
    if type(tomography_data) is list:
        return [direct_fragment_model(d, discard_poor_data, rank_cutoff) for d in tomography_data]
    if isinstance(tomography_data, SampledFragmentData):
        d = tomography_data
        p, m = len(d.prep_qubits), len(d.meas_qubits)
        _, W = _fit_matrix(p, m, rank_cutoff)
        blocks = (d.counts.to(torch.complex128) @ torch.as_tensor(W, device=d.counts.device).T)
        T = _tensor_from_choi(blocks.reshape(-1, 2 ** (p + m), 2 ** (p + m)), p, m) / d.shots
        return FragmentModel(T.reshape(d.counts.shape[0], d.counts.shape[1], -1), p, m, d.cut_labels, d.shots)
    if isinstance(tomography_data, FragmentData):
        d = tomography_data
        T = d.backend.pauli_tensor(d.states, d.prog, d.n_sets, len(d.prep_qubits), d.meas_qubits,
                                   frame=SIC_FRAME, basis_inputs=d.basis_inputs)
        model = FragmentModel(T, len(d.prep_qubits), len(d.meas_qubits), d.cut_labels, d.shots)
        model.exact = True                   # built from simulated amplitudes: every Choi block is a Gram matrix
        return model
    if not torch.cuda.is_available():
        raise _lib.QdcError("direct_fragment_model needs a CUDA device; there is no CPU fallback")
    return _model_from_counts(tomography_data, torch.device("cuda"), rank_cutoff)


_ML_TABLES = {}


def _ml_tables(n_prep, n_meas, dev):
    """Sparse forms of the Pauli <-> Choi-matrix change of basis for `qdc_ml_project`, derived from
    `_pauli_choi_basis` (so the axis and transposition conventions live in one place): every entry
    (i, j) of a block touches exactly D = 2^q tensor components and every component D entries.
      fwd: Lambda[i, j]  = sum_t T[fwd_k[i D + j, t]] * fwd_c[i D + j, t]         (2^-q folded in)
      inv: T'[kappa]     = Re sum_t inv_c[kappa, t] * L'[j_t, i_t],  inv_ij = (i << 8) | j"""
    key = (n_prep, n_meas, str(dev))
    if key not in _ML_TABLES:
        B = _pauli_choi_basis(n_prep, n_meas)
        K, D, _ = B.shape
        q = n_prep + n_meas
        fwd_k = np.zeros((D * D, D), dtype=np.int32)
        fwd_c = np.zeros((D * D, D, 2))
        for i in range(D):
            for j in range(D):
                ks = np.nonzero(B[:, i, j])[0]
                assert len(ks) == D
                fwd_k[i * D + j] = ks
                c = B[ks, i, j] / 2 ** q
                fwd_c[i * D + j, :, 0], fwd_c[i * D + j, :, 1] = c.real, c.imag
        inv_ij = np.zeros((K, D), dtype=np.int32)
        inv_c = np.zeros((K, D, 2))
        for k in range(K):
            ii, jj = np.nonzero(B[k])
            assert len(ii) == D
            inv_ij[k] = (ii << 8) | jj
            c = B[k, ii, jj]
            inv_c[k, :, 0], inv_c[k, :, 1] = c.real, c.imag
        _ML_TABLES[key] = tuple(torch.as_tensor(x, device=dev).contiguous() for x in (fwd_k, fwd_c, inv_ij, inv_c))
    return _ML_TABLES[key]


def maximum_likelihood_model(choi_matrix, force_full=False, assume_psd=None):
    """Nearest positive-semidefinite model (:320-356) on the device (`qdc_ml_project`,
    csrc/ml_project.cu): every Choi block is eigen-decomposed by a one-warp Jacobi, ALL eigenvalues
    of the fragment are clipped together -- the reference's sequential walk (:333-343) is
    soft-thresholding lambda' = max(lambda - t0, 0) with t0 the root of sum max(lambda - t, 0) =
    sum lambda, found by Newton steps over grid-wide reductions -- and the blocks are rebuilt as the
    complex conjugate of V diag V^+, as the reference's `to_projector` does (:45-46, 354-356).  When a
    Cholesky test finds every block positive semidefinite to 1e-14 of its trace (exact,
    infinite-shot data), t0 = 0 and only the conjugate is taken; `force_full=True` skips that test.
    `assume_psd` skips it the other way: models that `direct_fragment_model` built from exactly
    simulated amplitudes are Gram matrices, positive semidefinite by construction, and the reference's
    projection changes them by rounding noise only; they carry `exact = True` and take this path by
    default (`assume_psd=False` runs the test anyway).
    `model.ml_info[s]` holds (needed the full path, Newton steps, t0) of parameter set s on the
    device.  No host synchronisation."""
    if type(choi_matrix) is list:
        return [maximum_likelihood_model(m, force_full, assume_psd) for m in choi_matrix]
    if not isinstance(choi_matrix, FragmentModel):
        raise _lib.QdcError("maximum_likelihood_model expects the FragmentModel returned by "
                            "direct_fragment_model (device-resident Choi blocks)")
    model = choi_matrix
    if not model.T.is_cuda:
        raise _lib.QdcError("maximum_likelihood_model runs on the GPU only (qdc_ml_project): the model's "
                            "tensor is on " + str(model.T.device) + "; there is no CPU fallback")
    lib = _lib.load()
    dev = model.T.device
    q = model.n_prep + model.n_meas
    if q > 4:
        raise _lib.QdcError(f"maximum_likelihood_model: {q} cut ends on one fragment (blocks of {2 ** q} x {2 ** q}); "
                            "the device kernel takes at most 4")
    D = 2 ** q
    T = model.T.contiguous()
    n_sets, n_b, K = T.shape
    assert K == D * D
    fwd_k, fwd_c, inv_ij, inv_c = _ml_tables(model.n_prep, model.n_meas, dev)
    if assume_psd is None:
        assume_psd = bool(getattr(model, "exact", False))
    mode = 1 if force_full else (-1 if assume_psd else 0)
    out = torch.empty_like(T)
    ws_bytes = int(lib.qdc_ml_workspace_bytes(n_b, D))
    info = []
    with torch.cuda.device(dev):
        for s in range(n_sets):
            ws = torch.empty(ws_bytes, dtype=torch.uint8, device=dev)
            _lib.check(lib.qdc_ml_project(_lib.ptr(T[s]), n_b, D, K, _lib.ptr(fwd_k), _lib.ptr(fwd_c),
                                          _lib.ptr(inv_ij), _lib.ptr(inv_c), _lib.ptr(out[s]), K,
                                          _lib.ptr(ws), mode, _lib.stream_ptr()),
                       "qdc_ml_project")
            info.append(ws[:40])                         # MlState: need_full, done, counter, steps, t, total, |total|
    result = FragmentModel(out, model.n_prep, model.n_meas, model.cut_labels, model.scale)
    result.ml_info = info
    return result


def ml_info_host(model, set_index=0):
    """(needed the full path, Newton steps, t0) of `maximum_likelihood_model`'s last run (reads back)."""
    raw = model.ml_info[set_index].cpu().numpy()
    ints = raw[:16].view(np.int32)
    return bool(ints[0]), int(ints[3]), float(raw[16:24].view(np.float64)[0])


# ---------------------------------------------------------------------------------------------
# recombination (:363-472)
# ---------------------------------------------------------------------------------------------
class DeviceDistribution(Mapping):
    """Recombined distribution of one parameter set.  As a Mapping: the reference's
    `{bitstring: probability}` (full-circuit order, char j <-> wire N-1-j, ancillas first),
    read from the device array `P` (float64[2^N], index bit i <-> circuit wire i) on first use.
    `score` holds the fused MIS reduction / top-k result when the backend knows the graph."""

    def __init__(self, n_wires, P=None, score=None, norm=None):
        self.n_wires, self.P, self.score, self.norm = n_wires, P, score, norm
        self._dict = None

    def _materialise(self):
        if self._dict is None:
            if self.P is None:
                raise _lib.QdcError("this distribution was reduced on the fly and never stored "
                                    "(too large for HBM as a dense array); use .score")
            p = self.P.cpu().numpy()
            nz = np.nonzero(p)[0]
            self._dict = {_bits(int(x), self.n_wires): float(p[x]) for x in nz}
        return self._dict

    def __getitem__(self, key):
        return self._materialise()[key]

    def __iter__(self):
        return iter(self._materialise())

    def __len__(self):
        return len(self._materialise())


def _sorted_axes(model: FragmentModel):
    """T reshaped so that its cut axes are in ascending stitch-number order."""
    labels = list(model.cut_labels["prep"]) + list(model.cut_labels["meas"])
    q = len(labels)
    order = sorted(range(q), key=lambda a: labels[a])
    T = model.T
    if order != list(range(q)):
        shp = T.shape
        T = T.reshape(shp[0], shp[1], *([4] * q)).permute(0, 1, *[2 + a for a in order])
        T = T.reshape(shp[0], shp[1], -1).contiguous()
    return T, [labels[a] for a in order]


def _row_operand(models, sorted_T, sorted_labels, col_labels, s):
    """Combine fragments F-1 .. 1 of parameter set `s` into A[(b_{F-1} .. b_1), kappa_0] with
    kappa_0 over the cuts that touch fragment 0 (ascending stitch number), summing over every cut
    internal to these fragments.  Pre-contraction of small tensors: a torch einsum (library call)
    on the device; the 2^N-sized contraction that follows is the hand-written kernel."""
    F = len(models)
    if F == 2 and list(sorted_labels[1]) == list(col_labels):
        return sorted_T[1][s if sorted_T[1].shape[0] > 1 else 0]
    operands, out = [], []
    for f in range(F - 1, 0, -1):
        T = sorted_T[f][s if sorted_T[f].shape[0] > 1 else 0]
        q = len(sorted_labels[f])
        operands += [T.reshape(T.shape[0], *([4] * q)), [40 + f] + list(sorted_labels[f])]
        out.append(40 + f)
    A = torch.einsum(*operands, out + list(col_labels))
    return A.reshape(-1, 4 ** len(col_labels)).contiguous()


def recombine_fragment_models(frag_models, wire_path_map, method="network", backend=None,
                              graph=None, threshold=1e-5, topk=1, store=None, compact="auto",
                              contraction=None):
    """Full-circuit distribution from fragment models (:463-472): sum over the 4^K cut-basis
    index of the product of fragment tensors, bits permuted to circuit order, normalised.

    Returns one `DeviceDistribution` (a list of them for a batch of parameter sets).
    Extensions: `graph` (or `backend.graph`) fuses the MIS objective and top-k into the
    contraction; `store=False` never materialises the 2^N array (required beyond ~2^31 outputs);
    `compact` ("auto": from 2^24 outputs on) restricts the contraction to the fragment outcomes
    that occur at all (all-zero tensor rows are dropped, exactly; see recombine.compact_operands);
    `contraction` picks the contraction kernel: "f64" (FP64 DMMA, the default) or "i8" / "i8:S"
    (S radix-256 slices on the int8 tensor cores, stated error bound in include/qdc_b200.h;
    needs store=False and 4^K <= 256)."""
    if method not in ("network", "insertion"):
        raise ValueError("recombination method {method} not recognized")     # :469 (sic)
    if not all(isinstance(m, FragmentModel) and m.cut_labels is not None for m in frag_models):
        raise _lib.QdcError("recombine_fragment_models expects FragmentModel objects produced by "
                            "direct_fragment_model / maximum_likelihood_model from FragmentData")
    F = len(frag_models)
    dev = frag_models[0].T.device
    K = len(identify_stitches(wire_path_map))
    bitpos, widths = united_bit_wires(wire_path_map)
    N = len(bitpos)
    col_bits, row_bits = widths[0], N - widths[0]
    if graph is None and backend is not None:
        graph = backend.graph
    n_wires = len(wire_path_map)
    if graph is not None:
        n_nodes = graph.number_of_nodes()
        adj = [sum(1 << int(v) for v in graph.neighbors(i)) for i in range(n_nodes)]
        adj += [0] * (n_wires - n_nodes)
        node_mask = (1 << n_nodes) - 1
    else:
        adj, node_mask = [0] * n_wires, 0
    tables = _rc.build_tables(row_bits, col_bits, bitpos, adj, node_mask, device=dev)
    if store is None:
        store = N <= 30
    sorted_T, sorted_labels = zip(*[_sorted_axes(m) for m in frag_models])
    col_labels = sorted_labels[0]
    n_sets = max(m.n_sets for m in frag_models)
    results = []
    for s in range(n_sets):
        Bt = sorted_T[0][s if sorted_T[0].shape[0] > 1 else 0]
        if F == 1:
            A = torch.ones(1, 1, dtype=torch.float64, device=dev)
        else:
            A = _row_operand(frag_models, sorted_T, sorted_labels, col_labels, s)
        A, Bt = _rc.pad_kdim(A), _rc.pad_kdim(Bt)
        tabs = tables
        if compact is True or (compact == "auto" and N >= 24):
            A, Bt, tabs, _ = _rc.compact_operands(A, Bt, tables)
        c_method, c_slices = _rc.resolve_contraction(contraction)
        if c_method == "i8" and (store or A.shape[1] > 256):
            if contraction is not None:
                raise _lib.QdcError("contraction='i8' needs store=False and at most 256 cut-basis terms")
            c_method = "f64"                              # environment default: fall back to the exact kernel
        res = _rc.contract_and_score(A, Bt, tabs, scale=0.5 ** K, threshold=threshold, k=topk,
                                     store="circuit" if store else None, method=c_method,
                                     slices=c_slices)
        results.append(DeviceDistribution(n_wires, res.P, res, res.norm))
    return results[0] if n_sets == 1 else results
