"""Device wrappers of the CutQC-style stage kernels (csrc/cutqc.cu) -- in-memory counterparts of
the reference's measure.c / vertical_collapse.c / merge.c / build.c.  Host-side glue only: torch
owns the buffers, every number comes from the C-ABI (`qdc_cutqc_*`, `qdc_khatri_rao`,
`qdc_contract_*`).  No CPU fallback."""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence

import numpy as np
import torch

from .. import _lib
from .. import recombine as _rc

__all__ = ["measure", "vertical_collapse", "merge", "build"]


def _dev(device=None):
    if not torch.cuda.is_available():
        raise _lib.QdcError("the cutqc stage kernels need a CUDA device; there is no CPU fallback")
    return torch.device(device if device is not None else "cuda")


def measure(raw: torch.Tensor, keep_mask: int, sign_masks: Sequence[int], stream=None) -> torch.Tensor:
    """raw [n_inst, 2^d] (float32 or float64, device) -> measured [n_inst, 2^eff] float64.
    keep_mask: qubits measured in 'comp'; sign_masks[i]: qubits of instance i measured in a
    basis other than 'I' / 'comp' (cutqc/measure.c:123-161)."""
    lib = _lib.load()
    assert raw.is_cuda and raw.is_contiguous() and raw.dtype in (torch.float32, torch.float64)
    n_inst, length = raw.shape
    d = int(length).bit_length() - 1
    n_eff = bin(keep_mask).count("1")
    signs = torch.as_tensor(np.asarray(sign_masks, dtype=np.uint64).view(np.int64), device=raw.device)
    out = torch.empty(n_inst, 1 << n_eff, dtype=torch.float64, device=raw.device)
    with torch.cuda.device(raw.device):
        _lib.check(lib.qdc_cutqc_measure(_lib.ptr(raw), 0 if raw.dtype == torch.float64 else 1, d,
                                         n_inst, C.c_uint64(keep_mask), _lib.ptr(signs),
                                         _lib.ptr(out), _lib.stream_ptr(stream)), "qdc_cutqc_measure")
    return out


def vertical_collapse(measured: torch.Tensor, terms: Sequence[Sequence], stream=None):
    """measured [n_inst, len] float64; terms[t] = [(coefficient, instance row), ...]
    (cutqc/vertical_collapse.c:27-104).  Returns (kron [n_terms, len], nonzero flags int32)."""
    lib = _lib.load()
    assert measured.is_cuda and measured.is_contiguous() and measured.dtype == torch.float64
    dev = measured.device
    ptr = np.zeros(len(terms) + 1, dtype=np.int32)
    for t, term in enumerate(terms):
        ptr[t + 1] = ptr[t] + len(term)
    coeff = np.array([c for term in terms for c, _ in term], dtype=np.int32)
    inst = np.array([i for term in terms for _, i in term], dtype=np.int32)
    d_ptr, d_coeff, d_inst = (torch.as_tensor(a, device=dev) for a in (ptr, coeff, inst))
    length = measured.shape[1]
    out = torch.empty(len(terms), length, dtype=torch.float64, device=dev)
    flags = torch.zeros(len(terms), dtype=torch.int32, device=dev)
    with torch.cuda.device(dev):
        for lo in range(0, len(terms), 65535):                   # grid.y limit
            hi = min(lo + 65535, len(terms))
            _lib.check(lib.qdc_cutqc_collapse(_lib.ptr(measured), length, hi - lo,
                                              C.c_void_p(d_ptr.data_ptr() + 4 * lo),
                                              _lib.ptr(d_inst), _lib.ptr(d_coeff),
                                              C.c_void_p(out.data_ptr() + 8 * lo * length),
                                              C.c_void_p(flags.data_ptr() + 4 * lo),
                                              _lib.stream_ptr(stream)), "qdc_cutqc_collapse")
    return out, flags


def merge(vectors: torch.Tensor, qubit_states: Sequence[int], stream=None) -> torch.Tensor:
    """vectors [n_vec, 2^n_eff] float64; qubit_states as in merge_%d.txt (cutqc/merge.c:42-56),
    MSB first: -2 merged, -1 active, 0/1 zoomed.  Returns [n_vec, 2^active]."""
    lib = _lib.load()
    assert vectors.is_cuda and vectors.is_contiguous() and vectors.dtype == torch.float64
    n_eff = len(qubit_states)
    assert vectors.shape[1] == 1 << n_eff
    active = zoom = zval = 0
    for q, st in enumerate(qubit_states):
        bit = 1 << (n_eff - 1 - q)
        if st == -1:
            active |= bit
        elif st in (0, 1):
            zoom |= bit
            zval |= bit if st == 1 else 0
        elif st != -2:
            raise ValueError(f"illegal qubit state {st}")
    out = torch.empty(vectors.shape[0], 1 << bin(active).count("1"), dtype=torch.float64,
                      device=vectors.device)
    with torch.cuda.device(vectors.device):
        _lib.check(lib.qdc_cutqc_merge(_lib.ptr(vectors), vectors.shape[0], n_eff,
                                       C.c_uint64(active), C.c_uint64(zoom), C.c_uint64(zval),
                                       _lib.ptr(out), _lib.stream_ptr(stream)), "qdc_cutqc_merge")
    return out


def _operand(lib, vecs: torch.Tensor, idx: Sequence[int], kdim: int, sp) -> torch.Tensor:
    d_idx = torch.as_tensor(np.asarray(idx, dtype=np.int32), device=vecs.device)
    A = torch.empty(vecs.shape[1], kdim, dtype=torch.float64, device=vecs.device)
    _lib.check(lib.qdc_cutqc_gather(_lib.ptr(vecs), vecs.shape[1], _lib.ptr(d_idx), len(idx), kdim,
                                    _lib.ptr(A), sp), "qdc_cutqc_gather")
    return A


def build(kron: Sequence[torch.Tensor], term_index: Sequence[Sequence[int]], num_cuts: int,
          out_dtype=torch.float64, stream=None, contraction: Optional[str] = None) -> torch.Tensor:
    """reconstructed = 0.5^num_cuts * sum_t kron(v_0[t], v_1[t], ...)  (cutqc/build.c:69-196).

    kron[s]: [n_kron_s, 2^active_s] float64 vectors of subcircuit s, in build order (first =
    most significant, the reference's `smart_order`); `contraction`: "f64" | "i8[:S]" as in
    `recombine.resolve_contraction` (None: the QDC_CONTRACTION environment default; the int8
    kernel needs at most 256 summation terms).  term_index[s][t] = row of kron[s] used by
    summation term t, or -1 when that kron file does not exist (the term is skipped,
    build.c:35-67).  Returns the 2^total_active vector."""
    lib = _lib.load()
    S = len(kron)
    assert S >= 1 and len(term_index) == S
    n_terms = len(term_index[0])
    dev = kron[0].device
    sp = _lib.stream_ptr(stream)
    idx = np.asarray(term_index, dtype=np.int64)
    idx[:, (idx < 0).any(axis=0)] = -1                    # a term missing anywhere is zero everywhere
    kdim = max(16, (n_terms + 15) // 16 * 16)
    with torch.cuda.device(dev):
        ops = [_operand(lib, kron[s], idx[s], kdim, sp) for s in range(S)]
        if S == 1:
            ones = torch.zeros(1, kdim, dtype=torch.float64, device=dev)
            ones[0, :n_terms] = 1.0
            ops.append(ones)
        while len(ops) > 2:                               # fold the leading operands (Khatri-Rao)
            a0, a1 = ops[0], ops[1]
            W = torch.empty(a0.shape[0] * a1.shape[0], kdim, dtype=torch.float64, device=dev)
            _lib.check(lib.qdc_khatri_rao(_lib.ptr(a0), a0.shape[0], _lib.ptr(a1), a1.shape[0], kdim,
                                          _lib.ptr(W), sp), "qdc_khatri_rao")
            ops = [W] + ops[2:]
        A, Bt = ops
        rb = int(A.shape[0]).bit_length() - 1
        cb = int(Bt.shape[0]).bit_length() - 1
        tabs = _rc.build_tables(rb, cb, list(range(rb + cb)), [0] * (rb + cb), 0, device=dev)
        if (contraction is None and "QDC_CONTRACTION" not in os.environ and out_dtype == torch.float32 and 128 <= kdim <= 256
                and A.shape[0] * Bt.shape[0] >= 1 << 26):
            # float32 outputs (the reference's build.c is float32): 5 exact int8 slices bound the error at 1e-9 of the
            # largest product, far below float32 rounding, and the integer-scoring / FP32-fold epilogue is 2x the
            # FP64 DMMA kernel at 2^28 outputs and 256 terms (profiles/r2_time_build.log); below 2^26 outputs the
            # fixed cost of slicing the operands outweighs it, and with few terms the FP64 kernel is as fast
            contraction = "i8:5"
        method, slices = _rc.resolve_contraction(contraction)
        if method == "i8" and kdim > 256:
            if contraction is not None:
                raise _lib.QdcError("contraction='i8' takes at most 256 summation terms")
            method = "f64"
        res = _rc.contract_and_score(A, Bt, tabs, scale=0.5 ** num_cuts, threshold=float("inf"),
                                     k=1, normalize=False, store="united", out_dtype=out_dtype,
                                     cand_cap=1, stream=stream, method=method, slices=slices)
    return res.P.reshape(-1)
