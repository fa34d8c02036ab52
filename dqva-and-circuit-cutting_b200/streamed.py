"""Contraction + MIS scoring of HOST operands with the host -> device copies under the tensor cores.

`recombine.contract_and_score` wants both operands in HBM before its first kernel: the normalisation
(and with it every candidate test `p / norm > threshold`) needs the column sums of ALL rows of both
operands.  For operands that start in pinned host memory (the reference hands fragment data over as
numpy arrays, qsplit_mlrecon_methods.py:246-321) that serialises two 268 MB PCIe copies (9-10 ms at
C4) in front of a 38 ms contraction.  In an optimisation loop the norm of step i is known to within
rounding from step i - 1 (it is 1 analytically for exact fragment tensors), so `StreamedContraction`
SPECULATES on a caller-supplied `norm_hint`:

  * the row operand is cut into R row slabs and the column operand into C column chunks; the pieces
    cross PCIe on a copy stream in the order A0 B0 A1 B1 ..., each followed by an event;
  * the compute stream slices a piece when it has arrived (`prepare_i8`, once per piece) and runs
    the tcgen05 contraction of every (slab, chunk) pair whose two pieces are there, dividing by the
    hint and testing candidates against `threshold * SLACK` (SLACK = 1/2): a superset of the true
    candidates as long as the true norm is at least half the hint;
  * every pair's "sum p" is scale / hint x <column sums of its slab, column sums of its chunk> (computed from
    the moments the slicing pass leaves behind, not by the tensor cores), so the true norm is hint x the total of
    those; the R C result blocks are read back in ONE copy, the three sums are rescaled by hint / norm on the
    host (they are linear in 1 / norm) and the candidates re-tested exactly;
  * if the true norm turns out below SLACK * hint (or no hint was given) the operands, which are on
    the device by then, go through the ordinary `contract_and_score`: correct, just not overlapped.

With N ranks every rank streams ITS row slab, copies 1 / N of every column chunk and all-gathers
the chunk over NVLink behind the copy; the blocks of all ranks are all-gathered once at the end.  No result differs from the unstreamed path by more than the rounding of the
final rescale (tests/test_gpu_streamed.py: sums to 1e-13 relative, identical top-k indices).
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import ctypes as C

import numpy as np
import torch
import torch.distributed as dist

from . import _lib
from . import dist as qdist
from . import recombine as rc

__all__ = ["StreamedContraction", "combine_blocks", "SLACK"]

SLACK = 0.5          # candidates are flagged against threshold * SLACK; the hint may overshoot the norm by 1 / SLACK


def _splits(n: int, parts, align: int = 1) -> List[Tuple[int, int]]:
    """[(lo, count)] of consecutive pieces of n rows.  parts: a count (equal pieces) or a sequence of relative sizes,
    e.g. (1, 1, 2, 4): the first pair of pieces is what nothing can overlap with, so it pays to make it small, while
    the later ones amortise the per-launch cost.  Piece boundaries are multiples of `align`."""
    weights = [1] * parts if isinstance(parts, int) else [int(w) for w in parts]
    assert weights and all(w > 0 for w in weights)
    total, acc, cuts = sum(weights), 0, [0]
    for w in weights[:-1]:
        acc += w
        cuts.append(n * acc // total // align * align)
    cuts.append(n)
    return [(a, b - a) for a, b in zip(cuts, cuts[1:])]


def combine_blocks(blocks: np.ndarray, caps, norm_hint: float, threshold: float, k: int) -> dict:
    """Host side of a speculated step: `blocks[i]` is the result block of one (slab, chunk) pair of one rank
    (csrc/i8_layout.cuh: [0..2] sums, [4] slices, [7] candidates, [10] lost, [16..] top-k indices then
    probabilities), all computed with 1 / norm_hint and threshold * SLACK; caps[i] its candidate capacity.
    Every pair's "sum p" is scale / hint x <column sums of its slab, column sums of its chunk>, so over all pairs
    they add up to norm / hint: the true norm needs no extra pass.  Returns the evaluation's dict with
    "streamed": True, or {"streamed": False, "norm": norm} when the speculation has to be refused (norm below
    SLACK x hint: candidates may have been missed; or a pair ran out of candidate room)."""
    norm = float(norm_hint * blocks[:, 0].sum())
    if not (norm >= SLACK * norm_hint):
        return {"streamed": False, "norm": norm}
    f = norm_hint / norm
    cands, n_cand, used = [], 0, []
    for b, cap in zip(blocks, caps):
        n_b = int(b[7:8].view(np.int64)[0])
        if int(b[10:11].view(np.int64)[0]) > 0 or n_b > cap:
            return {"streamed": False, "norm": norm}
        n_cand += n_b
        used.append(int(b[4]))
        xs, ps = b[16:16 + k].view(np.int64), b[16 + k:16 + 2 * k]
        cands += [(int(x), float(p) * f) for x, p in zip(xs, ps) if x != -1 and float(p) * f > threshold]
    sums = blocks[:, 0:3].sum(axis=0) * f
    return {"sum_p": float(sums[0]), "sum_pw_indset": float(sums[1]), "sum_pw": float(sums[2]),
            "top": qdist.merge_topk(cands, k), "n_candidates": n_cand, "norm": norm, "streamed": True, "slices": used}


class StreamedContraction:
    """Reusable plan (staging buffers, sliced-operand buffers, per-pair tables, streams, events) for
    `P~ = scale * A @ Bt.T` scored on the fly, with A [M, kdim] and Bt [N, kdim] float64 in PINNED
    host memory.  `tables`: `recombine.build_tables` for the whole distribution.  row_chunks /
    col_chunks: R and C above as counts of equal pieces, or sequences of relative piece sizes such
    as (1, 1, 2).  In a process group every rank owns rows `dist.row_shard(M, rank, ws)` (pass the
    rank's slab of the pinned row operand to `run`, and the WHOLE pinned column operand: the rank
    reads only its 1 / N of every chunk)."""

    def __init__(self, tables: rc.MisTables, kdim: int, *, row_chunks=2, col_chunks=4, k: int = 1, device=None):
        self.tables = tables
        self.dev = tables.row_x.device if device is None else torch.device(device)
        self.rank, self.ws = qdist.world()
        assert k >= 1
        self.kdim, self.k = kdim, k
        M, N = tables.row_x.numel(), tables.col_x.numel()
        self.M, self.N = M, N
        self.row_lo, self.n_rows = qdist.row_shard(M, self.rank, self.ws)
        self.slabs = [(self.row_lo + lo, n) for lo, n in _splits(self.n_rows, row_chunks)]
        self.chunks = _splits(N, col_chunks, align=self.ws)
        self.R, self.C = len(self.slabs), len(self.chunks)
        if any(n <= 0 for _, n in self.slabs + self.chunks) or any(n % self.ws for _, n in self.chunks):
            raise ValueError("pieces must be non-empty and column chunks must split evenly over the ranks "
                             "(the all-gather's layout)")
        f64 = dict(dtype=torch.float64, device=self.dev)
        self.A_stage = torch.empty(self.n_rows, kdim, **f64)
        self.Bt_stage = torch.empty(N, kdim, **f64)
        self.copy_stream = torch.cuda.Stream(self.dev)
        self.done = torch.cuda.Event()
        t = tables
        self.slab_tabs = [(t.row_x[lo:lo + n].contiguous(), t.row_forb[lo:lo + n].contiguous(),
                           t.row_wok[lo:lo + n].contiguous()) for lo, n in self.slabs]
        self.chunk_tabs = [(t.col_x[lo:lo + n].contiguous(), t.col_wok[lo:lo + n].contiguous()) for lo, n in self.chunks]
        self.own_tabs = rc.MisTables(t.row_bits, t.col_bits, t.row_x[self.row_lo:self.row_lo + self.n_rows].contiguous(),
                                     t.row_forb[self.row_lo:self.row_lo + self.n_rows].contiguous(),
                                     t.row_wok[self.row_lo:self.row_lo + self.n_rows].contiguous(),
                                     t.col_x, t.col_wok, t.bitpos, t.node_mask)
        lib = _lib.load()

        def prepared(n):
            nb = int(lib.qdc_i8_operand_bytes(n, kdim, 6))
            if nb <= 0:
                raise _lib.QdcError(f"int8 contraction: unsupported piece shape ({n}, {kdim})")
            raw = torch.empty(nb + 256, dtype=torch.uint8, device=self.dev)
            off = (-raw.data_ptr()) % 256
            return raw[off:off + nb]
        self.a_bufs = [prepared(n) for _, n in self.slabs]
        self.b_bufs = [prepared(n) for _, n in self.chunks]
        self.colsums = torch.zeros(2, kdim, **f64)
        self.fallbacks = 0
        self.n_blk = 16 + 2 * max(k, 1)
        self.blocks = torch.empty(self.R * self.C, self.n_blk, **f64)
        self.hint_t = torch.ones(1, **f64)
        self._calls = {}

    def _planned(self, scale: float, threshold: float, tol: float, slices):
        """Argument lists of every `qdc_i8_prepare` (one per piece) and `qdc_contract_i8_prepared` (one per
        pair) of a step, built once per (scale, threshold, tol, slices): all buffers are this object's, so a step
        is R + C + R C plain foreign calls -- the smallest pairs run for 0.1-0.6 ms and the host must stay ahead."""
        key = (float(scale), float(threshold), float(tol), slices)
        plan = self._calls.get(key)
        if plan is not None:
            return plan
        auto = slices == "auto"
        stored = 6 if auto else int(slices)
        t, kd, k = self.tables, self.kdim, self.k
        n_nodes = bin(int(t.node_mask)).count("1")
        vp = _lib.ptr
        caps = [rc._default_cap(na * nb, threshold) for _, na in self.slabs for _, nb in self.chunks]
        # room for the outputs the integer filter flags (a superset of the candidates), as in recombine._contract_i8
        list_cap = max(max(cap, min(na * nb, 1 << 22))
                       for cap, (na, nb) in zip(caps, [(na, nb) for _, na in self.slabs for _, nb in self.chunks]))
        ws = rc._Workspace.get(self.dev, kd, list_cap, k)
        prep_a, prep_b, pairs = [], [], {}
        for i, (lo, n) in enumerate(self.slabs):
            piece = self.A_stage[lo - self.row_lo:lo - self.row_lo + n]
            prep_a.append([vp(piece), n, kd, stored, 0, vp(self.slab_tabs[i][2]), None, vp(self.a_bufs[i])])
        for i, (lo, n) in enumerate(self.chunks):
            piece = self.Bt_stage[lo:lo + n]
            prep_b.append([vp(piece), n, kd, stored, 1, vp(self.chunk_tabs[i][1]), vp(self.chunk_tabs[i][0]),
                           vp(self.b_bufs[i])])
        for r, (alo, na) in enumerate(self.slabs):
            a_src = self.A_stage[alo - self.row_lo:alo - self.row_lo + na]
            rx, rf, rw = self.slab_tabs[r]
            for c, (blo, nb) in enumerate(self.chunks):
                blk = self.blocks[r * self.C + c]
                pairs[(r, c)] = [vp(self.a_bufs[r]), na, vp(self.b_bufs[c]), nb, kd, stored, 0 if auto else stored,
                                 float(tol), n_nodes, float(scale), 2, vp(self.hint_t), None, 0, float(threshold),
                                 vp(a_src), vp(self.Bt_stage[blo:blo + nb]), 0, vp(rx), vp(rf), vp(rw), None, 0,
                                 vp(blk), vp(ws["cand_x"]), vp(ws["cand_p"]), list_cap, k,
                                 vp(blk[16:16 + k]), vp(blk[16 + k:16 + 2 * k]), vp(ws["partials"])]
        plan = (prep_a, prep_b, pairs, caps, ws)
        self._calls[key] = plan
        return plan

    # -- copy side ------------------------------------------------------------------------------------------
    def _issue_copies(self, A_pin: torch.Tensor, Bt_pin: torch.Tensor):
        """All pieces, in arrival order A0 B0 A1 B1 ...; returns [(kind, index, event)]."""
        cs = self.copy_stream
        cs.wait_event(self.done)                      # the previous run has finished reading the staging buffers
        order = []
        for i in range(max(self.R, self.C)):
            if i < self.R:
                order.append(("A", i))
            if i < self.C:
                order.append(("B", i))
        arrivals = []
        with torch.cuda.stream(cs):
            for kind, i in order:
                if kind == "A":
                    lo, n = self.slabs[i]
                    lo -= self.row_lo
                    self.A_stage[lo:lo + n].copy_(A_pin[lo:lo + n], non_blocking=True)
                else:
                    lo, n = self.chunks[i]
                    if self.ws == 1:
                        self.Bt_stage[lo:lo + n].copy_(Bt_pin[lo:lo + n], non_blocking=True)
                    else:
                        per = n // self.ws
                        o = lo + self.rank * per
                        own = self.Bt_stage[o:o + per]
                        own.copy_(Bt_pin[o:o + per], non_blocking=True)
                        dist.all_gather_into_tensor(self.Bt_stage[lo:lo + n].view(-1), own.reshape(-1))
                ev = torch.cuda.Event()
                ev.record(cs)
                arrivals.append((kind, i, ev))
        return arrivals

    # -- the step ---------------------------------------------------------------------------------------------
    def run(self, A_pin: torch.Tensor, Bt_pin: torch.Tensor, *, scale: float, threshold: float = 1e-5,
            norm_hint: Optional[float] = None, slices="auto", tol: float = 1e-10) -> dict:
        """One evaluation.  Returns the dict of `ScoreResult.get()` (sum_p, sum_pw_indset, sum_pw, top,
        n_candidates -- here the count above threshold * SLACK * hint / norm, an upper bound of the
        exact count) plus "norm" (feed it back as the next `norm_hint`), "streamed" (False when the
        speculation was refused and the ordinary path ran) and "slices" (per pair)."""
        assert A_pin.shape == (self.n_rows, self.kdim) and Bt_pin.shape == (self.N, self.kdim)
        assert A_pin.is_pinned() and Bt_pin.is_pinned() and A_pin.dtype == torch.float64
        dev, k = self.dev, self.k
        main = torch.cuda.current_stream(dev)
        arrivals = self._issue_copies(A_pin, Bt_pin)
        speculate = norm_hint is not None and norm_hint > 0.0 and np.isfinite(norm_hint)
        lib = _lib.load()
        sp = C.c_void_p(main.cuda_stream)
        if speculate:
            self.hint_t.fill_(norm_hint)
            prep_a, prep_b, pairs, caps, _ = self._planned(scale, threshold * SLACK, tol / (self.R * self.C * self.ws),
                                                          slices)
        else:
            self.colsums.zero_()
        have_a, have_b = [False] * self.R, [False] * self.C
        with torch.cuda.device(dev):
            for kind, i, ev in arrivals:
                main.wait_event(ev)
                if not speculate:                     # column sums of what THIS rank brought; the finish adds the ranks up
                    if kind == "A":
                        lo, n = self.slabs[i]
                        self.colsums[0] += self.A_stage[lo - self.row_lo:lo - self.row_lo + n].sum(dim=0)
                    else:
                        lo, n = self.chunks[i]
                        per = n // self.ws
                        self.colsums[1] += self.Bt_stage[lo + self.rank * per:lo + (self.rank + 1) * per].sum(dim=0)
                    continue
                if kind == "A":
                    _lib.check(lib.qdc_i8_prepare(*prep_a[i], sp), "qdc_i8_prepare")
                    have_a[i] = True
                    ready = [(i, c) for c in range(self.C) if have_b[c]]
                else:
                    _lib.check(lib.qdc_i8_prepare(*prep_b[i], sp), "qdc_i8_prepare")
                    have_b[i] = True
                    ready = [(r, i) for r in range(self.R) if have_a[r]]
                for rc_ in ready:
                    _lib.check(lib.qdc_contract_i8_prepared(*pairs[rc_], sp), "qdc_contract_i8_prepared")
        n_blk = self.n_blk
        flat = self.blocks.view(-1) if speculate else self.colsums.view(-1)
        if self.ws > 1:
            gathered = torch.empty(self.ws * flat.numel(), dtype=torch.float64, device=dev)
            dist.all_gather_into_tensor(gathered, flat)
            flat = gathered
        self.done.record(main)
        h = flat.cpu().numpy().reshape(self.ws, -1)
        if not speculate:
            cols = h.reshape(self.ws, 2, self.kdim).sum(axis=0)
            return self._unstreamed(scale, threshold, slices, tol, float(scale * np.dot(cols[0], cols[1])))
        out = combine_blocks(h.reshape(-1, n_blk), caps * self.ws, norm_hint, threshold, k)
        if out["streamed"]:
            return out
        return self._unstreamed(scale, threshold, slices, tol, out["norm"])

    def _unstreamed(self, scale, threshold, slices, tol, norm) -> dict:
        """The operands are in HBM; the ordinary path with the norm just computed."""
        self.fallbacks += 1
        norm_t = torch.tensor([norm], dtype=torch.float64, device=self.dev)
        res = rc.contract_and_score(self.A_stage, self.Bt_stage, self.own_tabs, scale=scale, threshold=threshold,
                                    k=self.k, method="i8", slices=slices, norm=norm_t, tol=tol / self.ws)
        self.done.record(torch.cuda.current_stream(self.dev))
        if self.ws > 1:
            sums, top = qdist.finish_block(res, self.k)
            out = {"sum_p": float(sums[0]), "sum_pw_indset": float(sums[1]), "sum_pw": float(sums[2]), "top": top,
                   "n_candidates": len(top)}
        else:
            out = res.get()
        out.update(norm=norm, streamed=False, slices=[res.slices])
        return out
