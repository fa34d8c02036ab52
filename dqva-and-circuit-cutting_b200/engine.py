"""`B200Backend` -- the object passed where the reference passes its Aer backend
(`sim_with_cutting(fragments, wire_path_map, frag_shots, backend)`, utils/cutting_funcs.py:167;
`collect_fragment_data(..., tomography_backend)`, qsplit/qsplit_mlrecon_methods.py:216).

Host-side only: lowers gate lists to the C-ABI's `qdc_gate` records, caches compiled schedules,
owns device buffers through torch, and calls the statevector kernels of csrc/statevec.cu
(`qdc_sv_compile/run/probs/pauli_tensor`).  There is no CPU path: every method raises `QdcError`
if `libqdc_b200.so` is missing.

Batching: a run simulates `n_init` initial-state variants x `n_sets` parameter vectors of ONE
circuit structure at once (the batch dimension the reference never forms: it calls
`qiskit.execute` on 4^p * 3^m separately assembled circuits per COBYLA evaluation).
"""
from __future__ import annotations

import contextlib
import ctypes as C
import math
from dataclasses import dataclass, field
from typing import List, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _lib
from .circuit import ParameterExpression

__all__ = ["B200Backend", "CompiledCircuit", "QdcGate", "lower_instructions", "SIC_KETS",
           "SIC_FRAME", "CUTQC_KETS", "CUTQC_FRAME", "BASIS_KETS"]

# gate kinds of include/qdc_b200.h
G_ID, G_X, G_Y, G_Z, G_H, G_S, G_SDG, G_T, G_TDG, G_RX, G_RY, G_RZ, G_U1, G_U2 = range(14)
_FIXED = {"x": G_X, "y": G_Y, "z": G_Z, "h": G_H, "s": G_S, "sdg": G_SDG, "t": G_T, "tdg": G_TDG}
_ROT = {"rx": G_RX, "ry": G_RY, "rz": G_RZ, "u1": G_U1}
SV_FUSE_ANCILLA = 1


class QdcGate(C.Structure):
    """ctypes mirror of `qdc_gate` (include/qdc_b200.h)."""
    _fields_ = [("kind", C.c_int32), ("target", C.c_int32), ("ctrl_mask", C.c_uint64),
                ("ctrl_val", C.c_uint64), ("param_slot", C.c_int32), ("reserved", C.c_int32),
                ("scale", C.c_double), ("offset", C.c_double), ("m", C.c_double * 8)]


# ---- tomography constants --------------------------------------------------------------------
# qiskit-ignis 0.5.1 SIC preparation circuits (used by partial_tomography with prep_basis="SIC",
# qsplit_mlrecon_methods.py:116-162): S0 = |0>, S_k = u3(theta, phi_k, 0)|0> with
# theta = -2 atan(sqrt 2), phi = pi, pi/3, -pi/3.  u3(t, p, l)|0> = (cos t/2, e^{ip} sin t/2).
def _sic_kets():
    th = -2.0 * math.atan(math.sqrt(2.0))
    kets = [np.array([1.0, 0.0], dtype=complex)]
    for phi in (math.pi, math.pi / 3, -math.pi / 3):
        kets.append(np.array([math.cos(th / 2), np.exp(1j * phi) * math.sin(th / 2)]))
    return np.array(kets)


_PAULIS = [np.eye(2, dtype=complex), np.array([[0, 1], [1, 0]], dtype=complex),
           np.array([[0, -1j], [1j, 0]]), np.diag([1.0 + 0j, -1.0])]          # I, X, Y, Z


def _frame(kets):
    """c[a, s] with Pauli_a = sum_s c[a, s] |k_s><k_s| (real for both preparation sets)."""
    basis = np.array([np.outer(k, k.conj()).reshape(-1) for k in kets]).T       # [4, s]
    out = np.zeros((4, 4))
    for a, P in enumerate(_PAULIS):
        sol = np.linalg.solve(basis, P.reshape(-1))
        assert np.abs(sol.imag).max() < 1e-12
        out[a] = sol.real
    return out


SIC_KETS = _sic_kets()
SIC_FRAME = _frame(SIC_KETS)
# cutqc initialisations zero, one, plus, plusI (cutqc/evaluator.py:36, :96-109); their frame is
# the table of cutqc/post_process.py:20-30: I = zero + one, X = 2 plus - zero - one, ...
CUTQC_KETS = np.array([[1, 0], [0, 1], [1 / math.sqrt(2), 1 / math.sqrt(2)],
                       [1 / math.sqrt(2), 1j / math.sqrt(2)]], dtype=complex)
CUTQC_FRAME = _frame(CUTQC_KETS)
# computational-basis inputs |0>, |1>: by linearity the 2^p basis evolutions carry everything the
# 4^p preparation variants do (see B200Backend.pauli_tensor, basis_inputs=True)
BASIS_KETS = np.array([[1, 0], [0, 1]], dtype=complex)


def _code_table(kets) -> np.ndarray:
    """[n_codes][re0, im0, re1, im1] as qdc_sv_run expects."""
    kets = np.asarray(kets, dtype=complex)
    return np.ascontiguousarray(np.stack([kets[:, 0].real, kets[:, 0].imag, kets[:, 1].real,
                                          kets[:, 1].imag], axis=1))


# ---- lowering --------------------------------------------------------------------------------
def _u3(theta, phi, lam):
    c, s = math.cos(theta / 2), math.sin(theta / 2)
    return np.array([[c, -np.exp(1j * lam) * s],
                     [np.exp(1j * phi) * s, np.exp(1j * (phi + lam)) * c]], dtype=complex)


def lower_instructions(instructions, param_order: Sequence = ()):
    """Plain-tuple instructions `(name, qubits, params, matrix)` -> (gate records, auto values).

    Symbolic angles (`ParameterExpression`) use the slot of their Parameter in `param_order`;
    numeric rotation angles are turned into extra slots after the symbolic ones ("auto" slots,
    values returned in order) so that circuits differing only in bound angles share one
    compiled schedule."""
    n_sym = len(param_order)
    slot_of = {p: i for i, p in enumerate(param_order)}
    recs, auto = [], []

    def angle(p):
        if isinstance(p, ParameterExpression):
            return slot_of[p.param], float(p.scale), float(p.offset)
        auto.append(float(p))
        return n_sym + len(auto) - 1, 1.0, 0.0

    def rec(kind, target, cmask=0, cval=0, ang=None, mat=None):
        slot, scale, offset = (-1, 0.0, 0.0) if ang is None else ang
        m = [0.0] * 8
        if mat is not None:
            mat = np.asarray(mat, dtype=complex).reshape(2, 2)
            m = [mat[0, 0].real, mat[0, 0].imag, mat[0, 1].real, mat[0, 1].imag,
                 mat[1, 0].real, mat[1, 0].imag, mat[1, 1].real, mat[1, 1].imag]
        recs.append((kind, int(target), int(cmask), int(cval), slot, scale, offset, tuple(m)))

    for name, qs, params, matrix in instructions:
        if name in ("barrier", "Id", "id"):
            continue
        if name in _FIXED:
            rec(_FIXED[name], qs[0])
        elif name in _ROT:
            rec(_ROT[name], qs[0], ang=angle(params[0]))
        elif name == "u3":
            if any(isinstance(p, ParameterExpression) for p in params):
                raise _lib.QdcError("u3 with unbound parameters is not supported")
            rec(G_U2, qs[0], mat=_u3(*map(float, params)))
        elif name == "cx":
            rec(G_X, qs[1], 1 << qs[0], 1 << qs[0])
        elif name == "cz":
            rec(G_Z, qs[1], 1 << qs[0], 1 << qs[0])
        elif name == "crx":
            rec(G_RX, qs[1], 1 << qs[0], 1 << qs[0], ang=angle(params[0]))
        elif name == "mcx":
            mask = sum(1 << q for q in qs[:-1])
            rec(G_X, qs[-1], mask, mask)
        elif name == "zcrx":
            mask = sum(1 << q for q in qs[:-1])
            rec(G_RX, qs[-1], mask, 0, ang=angle(params[0]))
        elif name == "unitary" and len(qs) == 1:
            rec(G_U2, qs[0], mat=matrix)
        else:
            raise _lib.QdcError(f"gate '{name}' on {len(qs)} qubits is outside the engine's gate "
                                "set (one-target gates with 0/1 controls)")
    return recs, auto


def _pack_gates(recs):
    arr = (QdcGate * max(len(recs), 1))()
    for g, (kind, target, cmask, cval, slot, scale, offset, m) in zip(arr, recs):
        g.kind, g.target, g.ctrl_mask, g.ctrl_val = kind, target, cmask, cval
        g.param_slot, g.scale, g.offset = slot, scale, offset
        for k in range(8):
            g.m[k] = m[k]
    return arr


@dataclass
class CompiledCircuit:
    """A compiled gate schedule (opaque `qdc_sv_program*`) plus what is needed to run it."""
    handle: C.c_void_p
    n_qubits: int
    n_sym: int                      # symbolic parameter slots (columns the caller provides)
    n_auto: int                     # numeric-angle slots appended by the lowering
    n_groups: int
    n_ops: int
    n_matrices: int
    tile_bits: int
    flip_mask: int

    def __del__(self):
        try:
            if self.handle:
                _lib.load().qdc_sv_free(self.handle)
                self.handle = None
        except Exception:
            pass

    def dump(self):
        """Schedule as numpy arrays (host-side tests of the compiler)."""
        lib = _lib.load()
        ng, no = self.n_groups, max(self.n_ops, 1)
        first = np.zeros(ng + 1, dtype=np.int32)
        lmask = np.zeros(ng, dtype=np.uint64)
        kind = np.zeros(no, dtype=np.int32)
        target = np.zeros(no, dtype=np.int32)
        cmask = np.zeros(no, dtype=np.uint64)
        cval = np.zeros(no, dtype=np.uint64)
        src = np.zeros(no, dtype=np.int32)
        conj = np.zeros(no, dtype=np.int32)
        _lib.check(lib.qdc_sv_dump(self.handle, *[a.ctypes.data_as(C.c_void_p) for a in
                                                  (first, lmask, kind, target, cmask, cval, src,
                                                   conj)]), "qdc_sv_dump")
        n = self.n_ops
        return {"group_first_op": first, "group_local_mask": lmask, "kind": kind[:n],
                "target": target[:n], "ctrl_mask": cmask[:n], "ctrl_val": cval[:n],
                "src_gate": src[:n], "conj_x": conj[:n]}


class B200Backend:
    """Batched complex128 statevector engine on one B200 (`cuda:<index>`)."""

    name = "qdc_b200_statevector"

    def __init__(self, device=None, tile_bits: int = 0, coalesce_bits: int = -1,
                 fuse_ancilla: bool = True, graph=None, max_state_bytes: int = 64 << 30):
        self.device = torch.device(device if device is not None else "cuda")
        self.tile_bits = tile_bits
        self.coalesce_bits = coalesce_bits
        self.fuse_ancilla = fuse_ancilla
        self.graph = graph                     # optional networkx graph for fused MIS scoring
        self.max_state_bytes = int(max_state_bytes)
        self._cache = {}
        _lib.load()                            # fail loudly, there is no CPU path

    # ---- compile --------------------------------------------------------------------------
    def compile_instructions(self, n_qubits: int, instructions, param_order: Sequence = (),
                             zero_mask: Optional[int] = None) -> Tuple[CompiledCircuit, List[float]]:
        """Returns (compiled schedule, auto-slot values of THIS instruction list)."""
        lib = _lib.load()
        recs, auto = lower_instructions(instructions, param_order)
        if zero_mask is None:
            zero_mask = (1 << n_qubits) - 1
        flags = SV_FUSE_ANCILLA if self.fuse_ancilla else 0
        key = (n_qubits, tuple(recs), len(param_order), zero_mask, self.tile_bits,
               self.coalesce_bits, flags)
        prog = self._cache.get(key)
        if prog is None:
            handle = C.c_void_p()
            gates = _pack_gates(recs)
            # the program's op and matrix tables are uploaded to the CURRENT device: make it this backend's
            # (host-only planning on a CPU "device" uploads nothing and needs no context)
            with (torch.cuda.device(self.device) if torch.device(self.device).type == "cuda" else contextlib.nullcontext()):
                _lib.check(lib.qdc_sv_compile(n_qubits, C.cast(gates, C.c_void_p), len(recs),
                                              C.c_uint64(zero_mask), self.tile_bits, self.coalesce_bits,
                                              flags, C.byref(handle)), "qdc_sv_compile")
            nq, ng, no, nm, tb = (C.c_int() for _ in range(5))
            flip = C.c_uint64()
            _lib.check(lib.qdc_sv_info(handle, C.byref(nq), C.byref(ng), C.byref(no), C.byref(nm),
                                       C.byref(tb), C.byref(flip)), "qdc_sv_info")
            prog = CompiledCircuit(handle, n_qubits, len(param_order), len(auto), ng.value,
                                   no.value, nm.value, tb.value, flip.value)
            self._cache[key] = prog
        return prog, auto

    def compile(self, circuit, param_order: Optional[Sequence] = None,
                zero_mask: Optional[int] = None):
        if param_order is None:
            param_order = list(circuit.parameters)
        return self.compile_instructions(circuit.num_qubits, circuit.to_instructions(),
                                         param_order, zero_mask)

    # ---- run ------------------------------------------------------------------------------
    def _param_table(self, prog: CompiledCircuit, auto: Sequence[float], params) -> torch.Tensor:
        if params is None:
            params = np.zeros((1, prog.n_sym))
        if isinstance(params, torch.Tensor):
            p = params.to(device=self.device, dtype=torch.float64)
        else:
            p = torch.as_tensor(np.asarray(params, dtype=np.float64), device=self.device)
        if p.dim() == 1:
            p = p.reshape(1, -1)
        if p.dim() != 2 or p.shape[1] != prog.n_sym:
            raise _lib.QdcError(f"parameter table must be [n_sets, {prog.n_sym}], got {tuple(p.shape)}")
        if prog.n_auto:
            a = torch.as_tensor(np.asarray(auto, dtype=np.float64), device=self.device)
            p = torch.cat([p, a.expand(p.shape[0], -1)], dim=1)
        return p.contiguous()

    def run(self, prog: CompiledCircuit, auto: Sequence[float] = (), params=None,
            prep_qubits: Sequence[int] = (), kets=None, init_lo: int = 0,
            n_init: Optional[int] = None, out: Optional[torch.Tensor] = None,
            stream=None) -> torch.Tensor:
        """Simulate `n_init` preparation variants x `n_sets` parameter rows.  Returns the
        amplitudes, complex128 [n_init * n_sets, 2^n] (batch = init * n_sets + set), stored in
        the schedule's flip frame (index x ^ prog.flip_mask holds basis state x)."""
        lib = _lib.load()
        table = self._param_table(prog, auto, params)
        n_sets = table.shape[0]
        n_prep = len(prep_qubits)
        if n_prep and kets is None:
            kets = SIC_KETS
        n_codes = len(kets) if n_prep else 1
        if n_init is None:
            n_init = n_codes ** n_prep
        n_batch = n_init * n_sets
        dim = 1 << prog.n_qubits
        if out is None:
            out = torch.empty(n_batch, dim, dtype=torch.complex128, device=self.device)
        assert out.is_contiguous() and out.numel() == n_batch * dim
        mats = torch.empty(max(n_sets * prog.n_matrices * 8, 1), dtype=torch.float64,
                           device=self.device)
        pq = (C.c_int32 * max(n_prep, 1))(*[int(q) for q in prep_qubits])
        code = _code_table(kets) if n_prep else np.zeros((1, 4))
        with torch.cuda.device(self.device):
            _lib.check(lib.qdc_sv_run(prog.handle, _lib.ptr(out), init_lo, n_init, n_sets,
                                      _lib.ptr(table), table.shape[1], pq, n_prep,
                                      code.ctypes.data_as(C.POINTER(C.c_double)), n_codes,
                                      _lib.ptr(mats), _lib.stream_ptr(stream)), "qdc_sv_run")
        return out

    def run_scored(self, prog: CompiledCircuit, auto: Sequence[float], params, adj: Sequence[int], node_mask: int,
                   scratch: Optional[torch.Tensor] = None, keep_state: bool = False, stream=None) -> torch.Tensor:
        """The uncut cost in one pass (`qdc_sv_run_score`): simulate every parameter row and reduce
        [sum p, sum p |x| [x independent], sum p |x|] inside the last gate group, so the final state is
        never re-read and (keep_state=False) never written.  Returns float64 [n_sets, 3] on the device.
        `scratch`: complex128 [n_sets, 2^n] working buffer (allocated when missing)."""
        lib = _lib.load()
        table = self._param_table(prog, auto, params)
        n_sets = table.shape[0]
        dim = 1 << prog.n_qubits
        if scratch is None:
            scratch = torch.empty(n_sets, dim, dtype=torch.complex128, device=self.device)
        assert scratch.is_contiguous() and scratch.numel() >= n_sets * dim
        mats = torch.empty(max(n_sets * prog.n_matrices * 8, 1), dtype=torch.float64, device=self.device)
        d_adj = torch.as_tensor(np.asarray([int(a) for a in adj], dtype=np.uint64).view(np.int64), device=self.device)
        assert d_adj.numel() >= prog.n_qubits
        ws = torch.empty(int(lib.qdc_sv_score_workspace_bytes(prog.handle, n_sets)) // 8 + 1, dtype=torch.float64,
                         device=self.device)
        sums = torch.empty(n_sets, 3, dtype=torch.float64, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(lib.qdc_sv_run_score(prog.handle, _lib.ptr(scratch), n_sets, _lib.ptr(table), table.shape[1],
                                            _lib.ptr(mats), _lib.ptr(d_adj), C.c_uint64(int(node_mask)),
                                            1 if keep_state else 0, _lib.ptr(ws), _lib.ptr(sums),
                                            _lib.stream_ptr(stream)), "qdc_sv_run_score")
        return sums

    def probabilities(self, states: torch.Tensor, prog_or_flip, n_qubits: Optional[int] = None,
                      stream=None) -> torch.Tensor:
        """|amp|^2 with the flip frame undone: float64 [batch, 2^n], index = basis state."""
        lib = _lib.load()
        flip = prog_or_flip.flip_mask if isinstance(prog_or_flip, CompiledCircuit) else int(prog_or_flip)
        if n_qubits is None:
            n_qubits = prog_or_flip.n_qubits
        n_batch = states.numel() >> n_qubits
        probs = torch.empty(n_batch, 1 << n_qubits, dtype=torch.float64, device=states.device)
        with torch.cuda.device(states.device):
            _lib.check(lib.qdc_sv_probs(_lib.ptr(states), n_qubits, n_batch, C.c_uint64(flip),
                                        _lib.ptr(probs), _lib.stream_ptr(stream)), "qdc_sv_probs")
        return probs

    def pauli_tensor(self, states: torch.Tensor, prog: CompiledCircuit, n_sets: int, n_prep: int,
                     meas_qubits: Sequence[int], frame=None, axis_strides=None,
                     ld: Optional[int] = None, out: Optional[torch.Tensor] = None,
                     stream=None, basis_inputs: bool = False) -> torch.Tensor:
        """T[set, b, kappa]: the fragment's Choi blocks in the Pauli basis (float64).

        basis_inputs=False: `states` are the 4^p preparation variants (`run(..., kets=SIC_KETS)`
        or CUTQC_KETS) and `frame` their dual frame.  basis_inputs=True: `states` are the 2^p
        evolutions of the computational-basis inputs (`run(..., kets=BASIS_KETS)`); since every
        preparation is a linear combination of them, T[b, a] = Tr(P_a M_b) with (M_b)_{x'x} =
        conj(psi_x'(b)) psi_x(b) -- the same numbers from 2^p instead of 4^p simulations
        (p + m <= 4)."""
        lib = _lib.load()
        n, m = prog.n_qubits, len(meas_qubits)
        n_axes = n_prep + m
        n_kappa = 4 ** n_axes
        if ld is None:
            ld = n_kappa
        n_b = 1 << (n - m)
        if out is None:
            alloc = torch.zeros if ld != n_kappa or axis_strides is not None else torch.empty
            out = alloc(n_sets, n_b, ld, dtype=torch.float64, device=states.device)
        if frame is None:
            frame = SIC_FRAME
        fr = np.ascontiguousarray(np.asarray(frame, dtype=np.float64).reshape(16))
        mq = (C.c_int32 * max(m, 1))(*[int(q) for q in meas_qubits])
        st = None
        if axis_strides is not None:
            st = (C.c_int64 * max(n_axes, 1))(*[int(s) for s in axis_strides])
        with torch.cuda.device(states.device):
            _lib.check(lib.qdc_sv_pauli_tensor(
                _lib.ptr(states), n, n_sets, C.c_uint64(prog.flip_mask), n_prep,
                2 if basis_inputs else 4,
                fr.ctypes.data_as(C.POINTER(C.c_double)), mq, m, st, _lib.ptr(out), ld,
                _lib.stream_ptr(stream)), "qdc_sv_pauli_tensor")
        return out
