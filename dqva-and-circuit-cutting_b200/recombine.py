"""Device pipeline for the 4^K contraction (K5) fused with MIS scoring / top-k (K8).

Host-side orchestration only: torch owns the device buffers and the stream, every number is
produced by the kernels in csrc/contract.cu and csrc/mis_reduce.cu through the C-ABI.

Replaces `recombine_fragment_models` (qsplit/qsplit_mlrecon_methods.py:463-472) followed by
`strip_ancillas` / `avg_cost` / `top_probs` (mis.py:114-134, 232-233) of the reference.
"""
from __future__ import annotations

import os
import ctypes as C
from dataclasses import dataclass
from typing import List, Optional, Sequence

import numpy as np
import torch

from . import _lib

__all__ = ["MisTables", "ScoreResult", "build_tables", "contract_and_score", "reduce_and_score",
           "compact_operands", "pad_kdim", "PreparedOperand", "prepare_i8", "i8_error_bounds"]


def _dev(device):
    return torch.device(device if device is not None else "cuda")


@dataclass
class MisTables:
    """Per-row / per-column scoring tables for a united index u = row * 2^col_bits + col."""
    row_bits: int
    col_bits: int
    row_x: torch.Tensor
    row_forb: torch.Tensor
    row_wok: torch.Tensor
    col_x: torch.Tensor
    col_wok: torch.Tensor
    bitpos: tuple
    node_mask: int


def build_tables(row_bits: int, col_bits: int, bitpos: Sequence[int], adj: Sequence[int],
                 node_mask: int, device=None, stream=None, index_flip: int = 0) -> MisTables:
    """bitpos[b] = full-circuit wire of united bit b (b < col_bits: column bits);
    adj[i] = neighbour mask of wire i; node_mask = wires that are graph nodes;
    index_flip: entry u of the scored array holds united index u ^ index_flip."""
    lib = _lib.load()
    dev = _dev(device)
    nb = row_bits + col_bits
    assert len(bitpos) == nb
    i64 = dict(dtype=torch.int64, device=dev)
    row_x = torch.empty(1 << row_bits, **i64)
    row_forb = torch.empty(1 << row_bits, **i64)
    row_wok = torch.empty(1 << row_bits, dtype=torch.int32, device=dev)
    col_x = torch.empty(1 << col_bits, **i64)
    col_wok = torch.empty(1 << col_bits, dtype=torch.int32, device=dev)
    h_bitpos = (C.c_int * max(nb, 1))(*[int(b) for b in bitpos])
    n_wires = len(adj)
    h_adj = (C.c_uint64 * max(n_wires, 1))(*[int(a) for a in adj])
    with torch.cuda.device(dev):
        _lib.check(lib.qdc_mis_build_tables(row_bits, col_bits, h_bitpos, h_adj, n_wires,
                                            C.c_uint64(node_mask), C.c_uint64(index_flip),
                                            _lib.ptr(row_x),
                                            _lib.ptr(row_forb), _lib.ptr(row_wok),
                                            _lib.ptr(col_x), _lib.ptr(col_wok),
                                            _lib.stream_ptr(stream)), "qdc_mis_build_tables")
    return MisTables(row_bits, col_bits, row_x, row_forb, row_wok, col_x, col_wok,
                     tuple(int(b) for b in bitpos), int(node_mask))


@dataclass
class ScoreResult:
    """Device-resident result of one scoring pass; `.get()` copies the few scalars to host."""
    sums: torch.Tensor          # [sum p, sum p*w*indset, sum p*w]
    top_x: torch.Tensor         # int64[k]  (circuit-order integer, -1 = none)
    top_p: torch.Tensor         # float64[k]
    cand_count: torch.Tensor    # int64[1]
    cand_cap: int
    P: Optional[torch.Tensor] = None
    norm: Optional[torch.Tensor] = None
    _slices: Optional[int] = None           # int8 method: slices used (None: chosen on the device, see `slices`)
    tol: Optional[float] = None
    block: Optional[torch.Tensor] = None    # int8 method: the result block all of the above are views of
    _host: Optional[np.ndarray] = None

    def _block_host(self) -> np.ndarray:
        """The int8 path keeps sums, norm, slice count, bounds, candidate count and top-k in ONE
        device block: one device -> host copy serves every accessor."""
        if self._host is None:
            self._host = self.block.cpu().numpy()
        return self._host

    @property
    def slices(self) -> Optional[int]:
        """Slices per operand the int8 contraction used; with slices="auto" the count was chosen on
        the device and this reads it back (a synchronisation)."""
        if self._slices is None and self.block is not None:
            self._slices = int(self._block_host()[4])
        return self._slices

    @property
    def error_bound(self) -> Optional[dict]:
        """A-posteriori bounds of the int8 contraction for the slice count in use (the plan kernel's,
        i8_prep.cu): {"slices", "tol", "per_output", "cost"}."""
        if self.block is None:
            return None
        h = self._block_host()
        return {"slices": int(h[4]), "tol": self.tol, "per_output": float(h[5]), "cost": float(h[6])}

    def get(self):
        if self.block is not None:
            h = self._block_host()
            k = self.top_x.numel()
            sums, n_cand = h[0:3], int(h[7:8].view(np.int64)[0])
            # h[9]: outputs the integer filter flagged (a superset, re-evaluated exactly on the device);
            # h[10]: candidates above the threshold that found no room because the flagged list was full
            if int(h[10:11].view(np.int64)[0]) > 0:
                raise _lib.QdcError(f"top-k candidate buffer overflow: {int(h[9:10].view(np.int64)[0])} outputs were flagged "
                                    f"for exact re-evaluation and did not fit the candidate arrays; raise the threshold or pass a larger cand_cap")
            xs, ps = h[16:16 + k].view(np.int64), h[16 + k:16 + 2 * k]
        else:
            sums = self.sums.cpu().numpy()
            n_cand = int(self.cand_count.item())
            xs = self.top_x.cpu().numpy()
            ps = self.top_p.cpu().numpy()
        if n_cand > self.cand_cap:
            raise _lib.QdcError(f"top-k candidate buffer overflow: {n_cand} entries above the "
                                f"threshold, capacity {self.cand_cap}")
        top = [(int(x), float(p)) for x, p in zip(xs, ps) if x != -1]
        return {"sum_p": float(sums[0]), "sum_pw_indset": float(sums[1]),
                "sum_pw": float(sums[2]), "top": top, "n_candidates": n_cand}


def pad_kdim(t: torch.Tensor, multiple: int = 16) -> torch.Tensor:
    """Zero-pad the contraction axis (last) of a [rows, kdim] operand to a multiple of 16."""
    k = t.shape[-1]
    if k % multiple == 0:
        return t.contiguous()
    out = torch.zeros(t.shape[0], (k + multiple - 1) // multiple * multiple, dtype=t.dtype,
                      device=t.device)
    out[:, :k] = t
    return out


class _Workspace:
    """Scratch buffers reused across calls (per device)."""
    _cache = {}

    @classmethod
    def get(cls, dev, kdim, cand_cap, k):
        lib = _lib.load()
        key = (str(dev), kdim, cand_cap, k)
        ws = cls._cache.get(key)
        if ws is None:
            f64 = dict(dtype=torch.float64, device=dev)
            ws = {
                "partials": torch.empty(3 * lib.qdc_reduce_partials(), **f64),
                "colsums": torch.empty(max(lib.qdc_norm_workspace_doubles(max(kdim, 1)), 1), **f64),
                "cand_x": torch.empty(max(cand_cap, 1), dtype=torch.int64, device=dev),
                "cand_p": torch.empty(max(cand_cap, 1), **f64),
            }
            cls._cache[key] = ws
        return ws

    @classmethod
    def i8_buffer(cls, dev, tag, n_bytes):
        """256-byte aligned scratch of at least n_bytes for the int8 path (prepared operands,
        column-sum partials), reused across calls on the same stream; grows, never shrinks."""
        key = (str(dev), "i8", tag, torch.cuda.current_stream(dev).cuda_stream)
        buf = cls._cache.get(key)
        if buf is None or buf.numel() < n_bytes + 256:
            buf = torch.empty(n_bytes + 256, dtype=torch.uint8, device=dev)
            cls._cache[key] = buf
        off = (-buf.data_ptr()) % 256
        return buf[off:off + n_bytes]


def resolve_contraction(spec: Optional[str]):
    """"f64" | "i8" | "i8:S" (S slices, 2..6) | "i8:auto" (slice count from the a-posteriori error
    bounds, see contract_and_score) -> (method, slices).  None reads the environment
    variable QDC_CONTRACTION (default "f64": FP64 DMMA, no approximation of any kind)."""
    if spec is None:
        spec = os.environ.get("QDC_CONTRACTION", "f64")
    if spec == "f64":
        return "f64", 6
    if spec.startswith("i8"):
        if spec == "i8:auto":
            return "i8", "auto"
        slices = int(spec[3:]) if len(spec) > 2 else 6
        if spec[2:3] not in ("", ":") or not 2 <= slices <= 6:
            raise ValueError(f"contraction {spec!r}: expected 'i8' or 'i8:S' with S in 2..6")
        return "i8", slices
    raise ValueError(f"contraction {spec!r} not recognized ('f64', 'i8', 'i8:S')")


I8_LSH_MAX = 20          # csrc/i8_layout.cuh
I8_GROUP = 32            # adjacent columns sharing one integer unit
I8_FOLD_COLS = 16        # columns per rounding of the epilogue's integer fold


def _pow2_ceiling_exponents(X: torch.Tensor) -> torch.Tensor:
    """Per row the exponent e with |x| 2^(7 - e) <= 127 that i8_prep_kernel applies (float64; -inf for
    a zero row): mx = m 2^e, m in [0.5, 1), one more when the 20 leading mantissa bits of m reach 127/128."""
    mx = X.abs().amax(dim=1)
    m, e = torch.frexp(mx)
    frac = torch.floor((m * 2.0 - 1.0) * 2.0 ** 20)
    e = e.to(torch.float64) + (frac >= 0xFC000).to(torch.float64)
    return torch.where(mx > 0, e, torch.full_like(e, float("-inf")))


def i8_error_bounds(A: torch.Tensor, Bt: torch.Tensor, tables: MisTables, scale: float,
                    norm: Optional[torch.Tensor], row_lo: int = 0, n_rows: Optional[int] = None):
    """A-posteriori bounds of the int8-sliced contraction for THESE operands, per slice count -- a
    torch restatement of what `i8_plan_kernel` (csrc/i8_prep.cu) computes on the device; the product
    path never calls it (tests and notebooks do).

    Per output |err| <= 2^(e_r + f_c) * scale / norm * kdim * (S + 1.1) * 2^(-8 S), e_r / f_c the
    exponents the slicing applies: per row of A its own; per column its own down to 2^-20 of the
    largest of its 32-column group (smaller columns are sliced at that floor).  The MIS cost only
    sums outputs whose row and column are themselves independent sets, each weighted by at most
    the number of graph nodes, so
        |cost error| <= n_nodes * (sum_{r ok} 2^e_r) * scale / norm * 2^(-8 S) *
                        ((sum_{c ok} 2^f_c) * kdim * (S + 1.1) + 2^f_max * 2^-27 * ceil(n_cols / 16))
    (the last term: the epilogue rounds each 16-column sum to the unit of the largest column
    exponent).  Returns {S: {"per_output": ..., "cost": ...}} for S = 2..6."""
    if n_rows is None:
        n_rows = A.shape[0] - row_lo
    kd = A.shape[1]
    ea = _pow2_ceiling_exponents(A[row_lo:row_lo + n_rows])
    fb = _pow2_ceiling_exponents(Bt)
    n_cols = Bt.shape[0]
    pad = (-n_cols) % I8_GROUP
    fb_pad = torch.cat([fb, fb.new_full((pad,), float("-inf"))]).view(-1, I8_GROUP)
    floor = fb_pad.amax(dim=1, keepdim=True) - I8_LSH_MAX
    fb = torch.where(torch.isinf(fb_pad), fb_pad, torch.maximum(fb_pad, floor)).reshape(-1)[:n_cols]
    pa, pb = torch.exp2(ea), torch.exp2(fb)                  # exp2(-inf) = 0: zero rows contribute nothing
    r_ok = ((tables.row_wok[row_lo:row_lo + n_rows] >> 8) & 1).to(torch.float64)
    c_ok = ((tables.col_wok >> 8) & 1).to(torch.float64)
    f = torch.full((), float(scale), dtype=torch.float64, device=A.device)
    if norm is not None:
        f = f / norm.reshape(()).abs()
    stats = torch.stack([pa.max() * pb.max() * f, (pa * r_ok).sum() * (pb * c_ok).sum() * f,
                         (pa * r_ok).sum() * pb.max() * f]).cpu().tolist()
    n_nodes = bin(int(tables.node_mask)).count("1")
    fold = 0.5 * ((n_cols + I8_FOLD_COLS - 1) // I8_FOLD_COLS) * 2.0 ** (-6 - I8_LSH_MAX)
    out = {}
    for S in range(2, 7):
        g = kd * (S + 1.1) * 2.0 ** (-8 * S)
        out[S] = {"per_output": stats[0] * g,
                  "cost": n_nodes * (stats[1] * g + stats[2] * fold * 2.0 ** (-8 * S))}
    return out


@dataclass
class PreparedOperand:
    """An FP64 operand of the int8 contraction after `prepare_i8`: radix-256 slices in the tensor
    cores' core-matrix order, per-row scales, scoring tables, column sums and error-bound statistics
    in one device buffer (csrc/i8_layout.cuh).  Prepare once, contract many times: a fragment whose
    tensor does not change between parameter sets is not sliced again."""
    buf: torch.Tensor           # uint8, 256-byte aligned
    src: torch.Tensor           # the FP64 rows it was made from (flagged candidates are re-evaluated exactly)
    n_rows: int
    kdim: int
    slices: int                 # slices stored (a contraction may use fewer)
    column_role: bool


def prepare_i8(X: torch.Tensor, wok: Optional[torch.Tensor], xbits: Optional[torch.Tensor] = None, *,
               slices: int = 6, column_role: bool = False, stream=None,
               out: Optional[torch.Tensor] = None) -> PreparedOperand:
    """One pass over a row-major float64 operand `X[n_rows, kdim]` (kdim <= 256): `qdc_i8_prepare`.
    wok: the rows' `weight | ok << 8` table (MisTables.row_wok / col_wok); xbits: col_x for the
    column operand (column_role=True), which is the operand streamed through the tiles."""
    lib = _lib.load()
    assert X.dtype == torch.float64 and X.is_contiguous() and X.dim() == 2
    n_rows, kdim = X.shape
    n_bytes = int(lib.qdc_i8_operand_bytes(n_rows, kdim, slices))
    if n_bytes <= 0:
        raise _lib.QdcError(f"int8 contraction: unsupported operand shape {tuple(X.shape)} / {slices} slices "
                            "(kdim <= 256, slices 2..6)")
    if out is None:
        raw = torch.empty(n_bytes + 256, dtype=torch.uint8, device=X.device)
        off = (-raw.data_ptr()) % 256
        out = raw[off:off + n_bytes]
    assert out.numel() >= n_bytes and out.data_ptr() % 256 == 0
    with torch.cuda.device(X.device):
        _lib.check(lib.qdc_i8_prepare(_lib.ptr(X), n_rows, kdim, slices, 1 if column_role else 0,
                                      _lib.ptr(wok), _lib.ptr(xbits), _lib.ptr(out),
                                      _lib.stream_ptr(stream)), "qdc_i8_prepare")
    return PreparedOperand(out, X, n_rows, kdim, slices, column_role)


def _default_cap(total: int, threshold: float) -> int:
    if threshold > 0:
        bound = int(2.0 / threshold) + 1024
    else:
        bound = 1 << 22
    return int(min(total, bound, 1 << 24))


def contract_and_score(A, Bt, tables: MisTables, scale: float = 1.0,
                       threshold: float = 1e-5, k: int = 1, normalize: bool = True,
                       row_lo: int = 0, n_rows: Optional[int] = None,
                       store: Optional[str] = None, out_dtype=torch.float64,
                       out: Optional[torch.Tensor] = None, cand_cap: Optional[int] = None,
                       stream=None, kernel_events: Optional[list] = None,
                       packed_cache: Optional[dict] = None, method: str = "f64",
                       slices=6, norm: Optional[torch.Tensor] = None, tol: float = 1e-10) -> ScoreResult:
    """P~ = scale * A @ Bt.T over rows [row_lo, row_lo + n_rows) of A, scored on the fly.

    A: [2^row_bits, kdim] float64 cuda, Bt: [2^col_bits, kdim]; kdim % 16 == 0 (see pad_kdim).
    store: None (never materialise), "united" (slab [n_rows, n_cols]) or "circuit"
    (full-circuit order, length 2^(row_bits+col_bits), scatter).
    kernel_events: if a list, a (start, end) CUDA event pair bracketing the contraction kernel
    alone (recorded on the launching stream) is appended to it.
    packed_cache: optional dict reused across calls (FP64 method); the caller promises that an
    operand with the same storage, shape and version counter has the same contents.
    norm: a 1-element float64 device tensor to divide by instead of the total of THIS call's
    operands (a rank that holds only its row slab passes `dist.sharded_norm`).
    method: "f64" (FP64 DMMA, `qdc_contract_f64`) or "i8" (exact radix-256 slicing on the tcgen05
    int8 tensor cores, `qdc_i8_prepare` + `qdc_contract_i8_prepared`; kdim <= 256, store None or
    "united"; error bound in include/qdc_b200.h).  With "i8", A and Bt may be `PreparedOperand`s
    (`prepare_i8`: A with the slab's rows only, Bt with column_role=True).
    slices: 2..6, or "auto" = the smallest count whose a-posteriori bounds (per output and on the
    MIS cost, computed on the device by the plan kernel; `ScoreResult.error_bound`) are both below
    `tol` for these operands (6 if none is).  No host synchronisation either way.
    stream: a torch.cuda.Stream; temporaries are allocated under it.
    """
    if stream is not None and stream != torch.cuda.current_stream(stream.device):
        with torch.cuda.stream(stream):
            return contract_and_score(A, Bt, tables, scale, threshold, k, normalize, row_lo, n_rows, store,
                                      out_dtype, out, cand_cap, stream, kernel_events, packed_cache,
                                      method, slices, norm, tol)
    if method == "i8":
        return _contract_i8(A, Bt, tables, scale, threshold, k, normalize, row_lo, n_rows, store, out_dtype,
                            out, cand_cap, stream, kernel_events, slices, norm, tol)
    lib = _lib.load()
    dev = A.device
    assert A.dtype == torch.float64 and Bt.dtype == torch.float64 and A.is_contiguous() \
        and Bt.is_contiguous()
    M, kdim = A.shape
    N = Bt.shape[0]
    assert Bt.shape[1] == kdim and M == tables.row_x.numel() and N == tables.col_x.numel()
    assert store != "united" or (M == 1 << tables.row_bits and N == 1 << tables.col_bits)
    if n_rows is None:
        n_rows = M - row_lo
    if cand_cap is None:
        cand_cap = _default_cap(n_rows * N, threshold)
    ws = _Workspace.get(dev, kdim, cand_cap, k)
    f64 = dict(dtype=torch.float64, device=dev)
    sums = torch.zeros(3, **f64)
    cand_count = torch.zeros(1, dtype=torch.int64, device=dev)
    top_x = torch.empty(k, dtype=torch.int64, device=dev)
    top_p = torch.empty(k, **f64)
    given_norm = norm is not None
    if given_norm:
        assert norm.dtype == torch.float64 and norm.numel() == 1 and norm.device == dev
    elif normalize:
        norm = torch.empty(1, **f64)
    store_mode = {None: 0, "united": 1, "circuit": 2}[store]
    odt = 0 if out_dtype == torch.float64 else 1
    if store_mode and out is None:
        if store_mode == 1:
            out = torch.empty(n_rows, N, dtype=out_dtype, device=dev)
        else:
            out = torch.zeros(1 << (tables.row_bits + tables.col_bits), dtype=out_dtype, device=dev)
    sp = _lib.stream_ptr(stream)
    with torch.cuda.device(dev):
        if normalize and not given_norm:
            _lib.check(lib.qdc_contract_norm(_lib.ptr(A), M, _lib.ptr(Bt), N, kdim, scale,
                                             _lib.ptr(ws["colsums"]), _lib.ptr(norm), sp),
                       "qdc_contract_norm")
        Ap = _packed(lib, A, sp, packed_cache)
        Bp = _packed(lib, Bt, sp, packed_cache)
        if kernel_events is not None:
            ev0 = torch.cuda.Event(enable_timing=True)
            ev1 = torch.cuda.Event(enable_timing=True)
            ev0.record(stream)
        _lib.check(lib.qdc_contract_f64(
            _lib.ptr(Ap), M, _lib.ptr(Bp), row_lo, n_rows, N, kdim, scale, _lib.ptr(norm), threshold,
            _lib.ptr(tables.row_x), _lib.ptr(tables.row_forb), _lib.ptr(tables.row_wok),
            _lib.ptr(tables.col_x), _lib.ptr(tables.col_wok), _lib.ptr(out), store_mode, odt,
            _lib.ptr(sums), _lib.ptr(ws["cand_x"]), _lib.ptr(ws["cand_p"]),
            _lib.ptr(cand_count), cand_cap, _lib.ptr(ws["partials"]), sp), "qdc_contract_f64")
        if kernel_events is not None:
            ev1.record(stream)
            kernel_events.append((ev0, ev1))
        _select_topk(lib, ws, cand_count, cand_cap, k, top_x, top_p, sp)
    return ScoreResult(sums, top_x, top_p, cand_count, cand_cap, out, norm)


def _contract_i8(A, Bt, tables, scale, threshold, k, normalize, row_lo, n_rows, store, out_dtype, out,
                 cand_cap, stream, kernel_events, slices, norm, tol) -> ScoreResult:
    """The int8 tensor-core method of `contract_and_score`.  Launches per call (slices given):
    prepare A slab, [column sums of A's other rows], prepare Bt, plan, contraction, finish -- no
    torch kernel, no host synchronisation; with slices="auto" the contraction is launched once per
    candidate slice count and all but the one the plan kernel chose return at once."""
    lib = _lib.load()
    auto = slices == "auto"
    stored = 6 if auto else int(slices)
    a_prep = A if isinstance(A, PreparedOperand) else None
    b_prep = Bt if isinstance(Bt, PreparedOperand) else None
    dev = (a_prep.buf if a_prep else A).device
    N = b_prep.n_rows if b_prep else Bt.shape[0]
    kdim = a_prep.kdim if a_prep else A.shape[1]
    assert (b_prep.kdim if b_prep else Bt.shape[1]) == kdim and N == tables.col_x.numel()
    if a_prep is not None:
        assert not a_prep.column_role and a_prep.slices >= stored
        stored = a_prep.slices
        n_rows = a_prep.n_rows if n_rows is None else n_rows
        assert n_rows == a_prep.n_rows, "a prepared row operand holds exactly the slab's rows"
        M_total = None
        a_src = a_prep.src
    else:
        assert A.dtype == torch.float64 and A.is_contiguous() and A.shape[0] == tables.row_x.numel()
        M_total = A.shape[0]
        if n_rows is None:
            n_rows = M_total - row_lo
        a_src = A[row_lo:row_lo + n_rows]
    if b_prep is not None:
        assert b_prep.column_role and b_prep.slices == stored, "both operands must store the same number of slices"
    else:
        assert Bt.dtype == torch.float64 and Bt.is_contiguous()
    store_mode = {None: 0, "united": 1}[store] if store in (None, "united") else None
    if store_mode is None:
        raise ValueError("the int8 contraction stores the united slab only (store=None or 'united')")
    if store_mode and out_dtype == torch.float32:
        store_mode = 2
    assert store_mode == 0 or N == 1 << tables.col_bits
    if cand_cap is None:
        cand_cap = _default_cap(n_rows * N, threshold)
    # the scoring-only kernels file every output their integer filter FLAGS (a superset of the candidates) in
    # the candidate arrays and a follow-up kernel evaluates those exactly: the arrays get room for the
    # superset (64 MB at most), `cand_cap` stays the limit on true candidates
    list_cap = cand_cap if store_mode == 1 else max(cand_cap, min(n_rows * N, 1 << 22))
    ws = _Workspace.get(dev, kdim, list_cap, k)
    if store_mode and out is None:
        out = torch.empty(n_rows, N, dtype=out_dtype, device=dev)
    block = torch.empty(16 + 2 * max(k, 1), dtype=torch.float64, device=dev)
    sums, norm_t = block[0:3], block[3:4]
    cand_count = block[7:8].view(torch.int64)
    top_x = block[16:16 + k].view(torch.int64)
    top_p = block[16 + k:16 + 2 * k]
    if norm is not None:
        assert norm.dtype == torch.float64 and norm.numel() == 1 and norm.device == dev
        norm_mode = 2
    else:
        norm_mode = 1 if normalize else 0
    sp = _lib.stream_ptr(stream)
    n_nodes = bin(int(tables.node_mask)).count("1")
    with torch.cuda.device(dev):
        if a_prep is None:
            nb = int(lib.qdc_i8_operand_bytes(n_rows, kdim, stored))
            if nb <= 0:
                raise _lib.QdcError(f"int8 contraction: unsupported shape kdim={kdim}, slices={stored}")
            a_prep = prepare_i8(a_src, tables.row_wok[row_lo:row_lo + n_rows], None, slices=stored,
                                column_role=False, stream=stream, out=_Workspace.i8_buffer(dev, "A", nb))
        extra, n_extra = None, 0
        if norm_mode == 1 and M_total is not None and n_rows < M_total:
            # the normalisation is a property of the WHOLE distribution: column sums of the rows of A
            # outside this rank's slab (a streaming pass, nothing is sliced)
            n_extra = int(lib.qdc_colsum_partials())
            extra = _Workspace.i8_buffer(dev, "extra", n_extra * kdim * 8)
            _lib.check(lib.qdc_colsum(_lib.ptr(A), 0, row_lo, row_lo + n_rows, M_total, kdim,
                                      _lib.ptr(extra), sp), "qdc_colsum")
        if b_prep is None:
            nb = int(lib.qdc_i8_operand_bytes(N, kdim, stored))
            b_prep = prepare_i8(Bt, tables.col_wok, tables.col_x, slices=stored, column_role=True,
                                stream=stream, out=_Workspace.i8_buffer(dev, "B", nb))
        if kernel_events is not None:
            ev0 = torch.cuda.Event(enable_timing=True)
            ev1 = torch.cuda.Event(enable_timing=True)
            ev0.record(stream)
        _lib.check(lib.qdc_contract_i8_prepared(
            _lib.ptr(a_prep.buf), n_rows, _lib.ptr(b_prep.buf), N, kdim, stored, 0 if auto else stored,
            float(tol), n_nodes, float(scale), norm_mode, _lib.ptr(norm), _lib.ptr(extra), n_extra,
            float(threshold), _lib.ptr(a_prep.src), _lib.ptr(b_prep.src), row_lo,
            _lib.ptr(tables.row_x), _lib.ptr(tables.row_forb), _lib.ptr(tables.row_wok),
            _lib.ptr(out), store_mode, _lib.ptr(block), _lib.ptr(ws["cand_x"]), _lib.ptr(ws["cand_p"]),
            list_cap, k, _lib.ptr(top_x), _lib.ptr(top_p), _lib.ptr(ws["partials"]), sp),
            "qdc_contract_i8_prepared")
        if kernel_events is not None:
            ev1.record(stream)
            kernel_events.append((ev0, ev1))
    return ScoreResult(sums, top_x, top_p, cand_count, cand_cap, out, norm_t,
                       _slices=None if auto else stored, tol=tol, block=block)


def compact_operands(A: torch.Tensor, Bt: torch.Tensor, tables: MisTables):
    """Drop the rows of A and of Bt that are identically zero, together with their table entries.

    A fragment outcome that never occurs has an all-zero tensor row (for the DQVA ansatz: every
    assignment of the fragment's nodes that is not an independent set of its subgraph), and every
    output that involves it has probability exactly 0, so the contraction only has to visit the
    product of the two supports.  Exact: nothing is thresholded.  Returns (A', Bt', tables',
    kept_fraction); costs one host synchronisation (the support sizes)."""
    rows = torch.nonzero(A.abs().amax(dim=1) > 0).reshape(-1)
    cols = torch.nonzero(Bt.abs().amax(dim=1) > 0).reshape(-1)
    M, N = A.shape[0], Bt.shape[0]
    frac = (rows.numel() * cols.numel()) / float(M * N)
    if rows.numel() == 0 or cols.numel() == 0:
        rows, cols = rows.new_zeros(1), cols.new_zeros(1)          # keep one (zero) row: sums stay 0
    if cols.numel() > 1 and cols.numel() % 2 == 1:
        # the kernel's epilogue walks columns in pairs: keep one of the all-zero columns as well
        spare = torch.ones(N, dtype=torch.bool, device=cols.device)
        spare[cols] = False
        cols = torch.sort(torch.cat([cols, torch.nonzero(spare).reshape(-1)[:1]])).values
    tabs = MisTables(tables.row_bits, tables.col_bits, tables.row_x[rows].contiguous(),
                     tables.row_forb[rows].contiguous(), tables.row_wok[rows].contiguous(),
                     tables.col_x[cols].contiguous(), tables.col_wok[cols].contiguous(),
                     tables.bitpos, tables.node_mask)
    return A[rows].contiguous(), Bt[cols].contiguous(), tabs, frac


def _packed(lib, X: torch.Tensor, sp, cache: Optional[dict]) -> torch.Tensor:
    """k-group-major copy [kdim/4][rows_pad][4] of a row-major operand (qdc_contract_pack)."""
    key = (X.data_ptr(), tuple(X.shape), X._version)
    if cache is not None and key in cache:
        return cache[key]
    rows, kdim = X.shape
    out = torch.empty(lib.qdc_contract_packed_doubles(rows, kdim), dtype=torch.float64,
                      device=X.device)
    _lib.check(lib.qdc_contract_pack(_lib.ptr(X), rows, kdim, _lib.ptr(out), sp),
               "qdc_contract_pack")
    if cache is not None:
        while len(cache) >= 4:   # both operands of the current and of the previous evaluation
            cache.pop(next(iter(cache)))
        cache[key] = out
    return out


def _select_topk(lib, ws, cand_count, cand_cap, k, top_x, top_p, sp):
    # the candidate count stays on the device: no host sync between scoring and selection
    _lib.check(lib.qdc_topk_select(_lib.ptr(ws["cand_x"]), _lib.ptr(ws["cand_p"]),
                                   _lib.ptr(cand_count), cand_cap, k,
                                   _lib.ptr(top_x), _lib.ptr(top_p), sp), "qdc_topk_select")


def reduce_and_score(P: torch.Tensor, tables: MisTables, threshold: float = 1e-5, k: int = 1,
                     norm: Optional[torch.Tensor] = None, row_lo: int = 0,
                     cand_cap: Optional[int] = None, stream=None,
                     amplitudes: bool = False) -> ScoreResult:
    """Score a materialised slab P[n_rows, 2^col_bits] (united order; float64 or float32).
    amplitudes=True: P is a float64 view of complex128 amplitudes (re, im interleaved, e.g.
    `torch.view_as_real(states)`), scored as |amp|^2 without materialising probabilities."""
    lib = _lib.load()
    dev = P.device
    assert P.is_contiguous() and P.dtype in (torch.float64, torch.float32)
    N = 1 << tables.col_bits
    if amplitudes:
        assert P.dtype == torch.float64
        P2 = P.reshape(-1, N, 2)
    else:
        P2 = P.reshape(-1, N)
    n_rows = P2.shape[0]
    if cand_cap is None:
        cand_cap = _default_cap(n_rows * N, threshold)
    ws = _Workspace.get(dev, 0, cand_cap, k)
    f64 = dict(dtype=torch.float64, device=dev)
    sums = torch.zeros(3, **f64)
    cand_count = torch.zeros(1, dtype=torch.int64, device=dev)
    top_x = torch.empty(k, dtype=torch.int64, device=dev)
    top_p = torch.empty(k, **f64)
    sp = _lib.stream_ptr(stream)
    with torch.cuda.device(dev):
        _lib.check(lib.qdc_mis_reduce(
            _lib.ptr(P2), 2 if amplitudes else (0 if P.dtype == torch.float64 else 1), row_lo,
            n_rows, N,
            _lib.ptr(norm), threshold, _lib.ptr(tables.row_x), _lib.ptr(tables.row_forb),
            _lib.ptr(tables.row_wok), _lib.ptr(tables.col_x), _lib.ptr(tables.col_wok),
            _lib.ptr(sums), _lib.ptr(ws["cand_x"]), _lib.ptr(ws["cand_p"]),
            _lib.ptr(cand_count), cand_cap, _lib.ptr(ws["partials"]), sp), "qdc_mis_reduce")
        _select_topk(lib, ws, cand_count, cand_cap, k, top_x, top_p, sp)
    return ScoreResult(sums, top_x, top_p, cand_count, cand_cap, P, norm)


def penalty_reduce(P: torch.Tensor, n_bits: int, adj: Sequence[int], node_mask: int,
                   index_flip: int = 0, amplitudes: bool = False, stream=None) -> torch.Tensor:
    """[sum p, sum p * #edges(x), sum p * |x|] over 2^n_bits entries (device float64[3]); the
    QAOA+ objective of ansatz/qaoa_plus.py:40-53 is sums[2] - Lambda * sums[1].  `P`: float64 /
    float32 probabilities, or with amplitudes=True a float64 view of complex128 amplitudes."""
    lib = _lib.load()
    dev = P.device
    assert P.is_contiguous() and P.dtype in (torch.float64, torch.float32)
    assert P.numel() == (2 if amplitudes else 1) << n_bits
    ws = torch.empty(lib.qdc_penalty_workspace_bytes(), dtype=torch.uint8, device=dev)
    partials = torch.empty(3 * lib.qdc_reduce_partials(), dtype=torch.float64, device=dev)
    sums = torch.zeros(3, dtype=torch.float64, device=dev)
    h_adj = (C.c_uint64 * max(len(adj), 1))(*[int(a) for a in adj])
    with torch.cuda.device(dev):
        _lib.check(lib.qdc_penalty_reduce(_lib.ptr(P), 2 if amplitudes else (0 if P.dtype == torch.float64 else 1),
                                          n_bits, C.c_uint64(index_flip), h_adj, len(adj),
                                          C.c_uint64(node_mask), _lib.ptr(ws), _lib.ptr(sums),
                                          _lib.ptr(partials), _lib.stream_ptr(stream)),
                   "qdc_penalty_reduce")
    return sums
