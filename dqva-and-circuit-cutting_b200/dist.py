"""Multi-GPU plumbing for the QDC hot path: one process per GPU, `torch.distributed`.

The path shards without any bulk exchange: fragment variants x parameter sets
are independent, and the reconstructed distribution is partitioned by bitstring prefix (rows of
the row operand).  The only collective is the finish: every rank contributes three partial sums
and its local top-k; one small all_gather makes the global scalars and global top-k available on
every rank.  NCCL on GPUs, gloo in the CPU tests.
"""
from __future__ import annotations

import os
from typing import List, Sequence, Tuple

import torch
import torch.distributed as dist

__all__ = ["init_from_env", "world", "row_shard", "round_robin", "finish", "finish_block", "PendingFinish", "merge_topk",
           "sharded_over_parameter_sets", "sharded_norm", "stage_sharded_operands", "bind_host_near_gpu"]


def init_from_env(backend: str | None = None) -> Tuple[int, int, int]:
    """Initialise the default process group from RANK/WORLD_SIZE/MASTER_* (torchrun) if
    WORLD_SIZE > 1.  Returns (rank, world_size, local_rank)."""
    world_size = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world_size > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        kwargs = {}
        if backend == "nccl":
            torch.cuda.set_device(local_rank)
            kwargs["device_id"] = torch.device("cuda", local_rank)
        dist.init_process_group(backend=backend, rank=rank, world_size=world_size, **kwargs)
        if backend == "nccl":
            bind_host_near_gpu(local_rank)
    return rank, world_size, local_rank


def bind_host_near_gpu(local_rank: int) -> List[int]:
    """Pin this process to the host cores NVML reports as local to GPU `local_rank` (its NUMA node), so
    that the pinned staging buffers allocated afterwards are first-touched on the memory next to that
    GPU's PCIe root: with 8 ranks copying at once, buffers that all sit on one socket share its memory
    controllers and the inter-socket link.  Returns the cores chosen ([] = left alone: NVML missing,
    no GPU, or QDC_BIND_NUMA=0)."""
    if os.environ.get("QDC_BIND_NUMA", "1") == "0" or not torch.cuda.is_available():
        return []
    try:
        import pynvml
        pynvml.nvmlInit()
        pr = torch.cuda.get_device_properties(local_rank)
        bus = f"{pr.pci_domain_id:08x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        h = pynvml.nvmlDeviceGetHandleByPciBusId(bus.encode())
        n_cpu = os.cpu_count() or 1
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (n_cpu + 63) // 64)
        cores = [64 * w + b for w, word in enumerate(words) for b in range(64) if (int(word) >> b) & 1]
        allowed = sorted(set(cores) & os.sched_getaffinity(0))
        if allowed:
            os.sched_setaffinity(0, allowed)
        return allowed
    except Exception:                          # placement is an optimisation, never a requirement
        return []


def world() -> Tuple[int, int]:
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def row_shard(n_rows: int, rank: int, world_size: int) -> Tuple[int, int]:
    """Contiguous prefix slab [lo, lo + count) of the row operand owned by `rank`.
    Rows are the most significant united bits, so a slab is a set of bitstring prefixes."""
    base, rem = divmod(n_rows, world_size)
    lo = rank * base + min(rank, rem)
    return lo, base + (1 if rank < rem else 0)


def round_robin(n_items: int, rank: int, world_size: int) -> List[int]:
    """Indices of the (fragment variant x parameter set) work items owned by `rank`."""
    return list(range(rank, n_items, world_size))


def sharded_norm(A_slab: torch.Tensor, Bt: torch.Tensor, scale: float) -> torch.Tensor:
    """Normalisation of the recombined distribution when every rank holds only ITS rows of the row
    operand: sum_{r,c} P~ = scale * sum_k (sum_r A[r,k]) (sum_c Bt[c,k]), so the ranks all-reduce
    the kdim column sums of their slabs (the one real exchange of the sharded-operand layout:
    kdim doubles per rank).  Returns a 1-element float64 tensor on A_slab's device, the form
    `recombine.contract_and_score(norm=...)` takes."""
    col = A_slab.sum(dim=0, dtype=torch.float64)
    _, ws = world()
    if ws > 1:
        dist.all_reduce(col, op=dist.ReduceOp.SUM)
    return (scale * torch.dot(col, Bt.sum(dim=0, dtype=torch.float64))).reshape(1)


def stage_sharded_operands(A_pin_slab: torch.Tensor, Bt_pin_shard: torch.Tensor, A_stage: torch.Tensor,
                           Bt_stage: torch.Tensor, col_lo: int, scale: float) -> torch.Tensor:
    """Host operands -> device for one rank of a sharded contraction, and the normalisation they imply.
    Every rank owns a slab of rows of the row operand (pinned `A_pin_slab` -> `A_stage`) and 1 / N of
    the rows of the column operand (`Bt_pin_shard` -> `Bt_stage[col_lo : col_lo + len]`), so the
    column operand crosses PCIe once in total; the shards are all-gathered over NVLink while the row
    slab is still arriving over PCIe (the collective is issued asynchronously right after the shard's
    copy), and the normalisation comes from the column sums of what each rank received from the host
    -- its row slab and its column SHARD, not the gathered operand -- in one all-reduce of 2 kdim
    doubles.  Requires equal shards (the all-gather's layout).  Returns the 1-element float64 norm
    tensor `recombine.contract_and_score(norm=...)` takes; on return the current stream has waited
    for both collectives."""
    n_own = Bt_pin_shard.shape[0]
    own = Bt_stage[col_lo:col_lo + n_own]
    own.copy_(Bt_pin_shard, non_blocking=True)
    gather = dist.all_gather_into_tensor(Bt_stage.view(-1), own.reshape(-1), async_op=True)
    A_stage.copy_(A_pin_slab, non_blocking=True)
    cols = torch.cat([A_stage.sum(dim=0, dtype=torch.float64), own.sum(dim=0, dtype=torch.float64)])
    reduce = dist.all_reduce(cols, op=dist.ReduceOp.SUM, async_op=True)
    gather.wait()
    reduce.wait()
    kd = A_stage.shape[1]
    return (scale * torch.dot(cols[:kd], cols[kd:])).reshape(1)


def merge_topk(cands: Sequence[Tuple[int, float]], k: int) -> List[Tuple[int, float]]:
    """Global top-k with the library's order: probability descending, then index ascending."""
    return sorted(cands, key=lambda t: (-t[1], t[0]))[:k]


def finish(sums: torch.Tensor, top_x: torch.Tensor, top_p: torch.Tensor, k: int):
    """All-gather [3 sums | k probs | k indices] from every rank; returns
    (global sums tensor[3], global top-k list) on every rank.  With world_size 1 this is a
    local repack and no collective is issued."""
    rank, ws = world()
    if ws == 1:
        xs, ps = top_x.cpu().tolist(), top_p.cpu().tolist()
        return sums.clone(), merge_topk([(x, p) for x, p in zip(xs, ps) if x != -1], k)
    # float64 carries 2^53 exactly, enough for indices of up to 53 bits
    packed = torch.cat([sums.to(torch.float64), top_p.to(torch.float64),
                        top_x.to(torch.float64)])
    gathered = torch.empty(ws * packed.numel(), dtype=torch.float64, device=packed.device)
    dist.all_gather_into_tensor(gathered, packed)
    g = gathered.view(ws, -1)
    total = g[:, :3].sum(dim=0)
    host = g.cpu()
    cands = []
    for r in range(ws):
        ps = host[r, 3:3 + k].tolist()
        xs = host[r, 3 + k:3 + 2 * k].tolist()
        cands += [(int(x), p) for x, p in zip(xs, ps) if x >= 0]
    return total, merge_topk(cands, k)


class PendingFinish:
    """`finish_block(..., wait=False)`: the all-gather has been issued, nothing has been read back.  `get()`
    copies the gathered blocks to the host (the synchronisation) and returns what `finish_block` returns."""

    def __init__(self, gathered: torch.Tensor, ws: int, k: int, cand_cap: int):
        self.gathered, self.ws, self.k, self.cand_cap = gathered, ws, k, cand_cap

    def get(self):
        import numpy as np
        h = self.gathered.cpu().numpy().reshape(self.ws, -1)
        k, cands = self.k, []
        for r in range(self.ws):
            xs = h[r, 16:16 + k].view(np.int64)
            ps = h[r, 16 + k:16 + 2 * k]
            n_cand = int(h[r, 7:8].view(np.int64)[0])
            if n_cand > self.cand_cap:
                raise RuntimeError(f"rank {r}: top-k candidate buffer overflow ({n_cand} > {self.cand_cap})")
            cands += [(int(x), float(p)) for x, p in zip(xs, ps) if x != -1]
        return h[:, 0:3].sum(axis=0), merge_topk(cands, k)


def finish_block(res, k: int, wait: bool = True):
    """`finish` for a `ScoreResult` of the int8 path, whose sums, candidate count and top-k live in
    ONE device block: the block itself is all-gathered (indices travel as their 64-bit patterns) and
    read back once -- no packing kernels.  Returns (global sums as a numpy array[3], global top-k
    list) on every rank; with wait=False a `PendingFinish` whose `get()` does the read-back, so a loop
    of evaluations whose results are only looked at afterwards never stalls the device on the host."""
    rank, ws = world()
    block = res.block
    if ws == 1:
        gathered = block
    else:
        gathered = torch.empty(ws * block.numel(), dtype=block.dtype, device=block.device)
        dist.all_gather_into_tensor(gathered, block)
    pending = PendingFinish(gathered, ws, k, res.cand_cap)
    return pending.get() if wait else pending


def sharded_over_parameter_sets(fn, table, device=None):
    """Evaluate `fn(rows) -> float array[len(rows)]` for every row of the parameter table with
    the rows dealt round-robin over the ranks (parameter sets are independent: SURVEY.md 8e, the
    C2 batch and the C5 sweep), then all-gather the scalars so that every rank returns the full
    vector in table order.  The only exchange is `ceil(n_sets / world_size)` doubles per rank."""
    import numpy as np
    table = np.atleast_2d(np.asarray(table, dtype=np.float64))
    n_sets = table.shape[0]
    rank, ws = world()
    if ws == 1:
        return np.asarray(fn(table), dtype=np.float64)
    mine = round_robin(n_sets, rank, ws)
    local = np.asarray(fn(table[mine]), dtype=np.float64) if mine else np.zeros(0)
    per = (n_sets + ws - 1) // ws
    if device is None:
        device = torch.device("cuda", torch.cuda.current_device()) \
            if dist.get_backend() == "nccl" else torch.device("cpu")
    buf = torch.zeros(per, dtype=torch.float64, device=device)
    buf[:len(mine)] = torch.as_tensor(local, dtype=torch.float64)
    gathered = torch.empty(ws * per, dtype=torch.float64, device=device)
    dist.all_gather_into_tensor(gathered, buf)
    g = gathered.view(ws, per).cpu().numpy()
    out = np.empty(n_sets)
    for r in range(ws):
        idx = round_robin(n_sets, r, ws)
        out[idx] = g[r, :len(idx)]
    return out
