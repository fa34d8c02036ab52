"""Phase times of the sharded end-to-end step (bench.py's e2e leg at N > 1), each phase bracketed by a
device synchronise so the numbers add up serially; then the step as the bench runs it (overlapped).
Run under torchrun: python -m torch.distributed.run --nproc-per-node N scripts/probe_e2e_phases.py"""
import importlib
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

importlib.import_module("dqva-and-circuit-cutting_b200")
from qdc_b200 import dist as qdist, recombine as rc  # noqa: E402


def main():
    rank, ws, local = qdist.init_from_env()
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    rb = cb = 17
    n = rb + cb
    A_h, Bt_h = bench.make_operands(rb, cb)
    _, adj = bench.make_graph(n)
    bitpos = bench.make_bitpos(n, cb)
    scale = 0.5 ** bench.NUM_CUTS
    M, N = 1 << rb, 1 << cb
    row_lo, n_rows = qdist.row_shard(M, rank, ws)
    c_lo, c_n = qdist.row_shard(N, rank, ws)
    tabs = rc.build_tables(rb, cb, bitpos, adj, (1 << n) - 1, device=dev)
    tabs_slab = rc.MisTables(tabs.row_bits, tabs.col_bits, tabs.row_x[row_lo:row_lo + n_rows].contiguous(),
                             tabs.row_forb[row_lo:row_lo + n_rows].contiguous(),
                             tabs.row_wok[row_lo:row_lo + n_rows].contiguous(), tabs.col_x, tabs.col_wok,
                             tabs.bitpos, tabs.node_mask)
    A_pin = torch.from_numpy(A_h).pin_memory()[row_lo:row_lo + n_rows]
    Bt_pin = torch.from_numpy(Bt_h).pin_memory()[c_lo:c_lo + c_n]
    A_stage = torch.empty(n_rows, bench.KDIM, dtype=torch.float64, device=dev)
    Bt_stage = torch.empty(N, bench.KDIM, dtype=torch.float64, device=dev)
    tol = 1e-10 / ws
    k = 8

    def sync():
        torch.cuda.synchronize()

    phases = {}

    def lap(name, t0):
        sync()
        phases.setdefault(name, []).append((time.perf_counter() - t0) * 1e3)

    for it in range(8):
        dist.barrier()
        sync()
        t = time.perf_counter()
        own = Bt_stage[c_lo:c_lo + c_n]
        own.copy_(Bt_pin, non_blocking=True)
        lap("h2d column shard", t)
        t = time.perf_counter()
        dist.all_gather_into_tensor(Bt_stage.view(-1), own.reshape(-1))
        lap("all_gather", t)
        t = time.perf_counter()
        A_stage.copy_(A_pin, non_blocking=True)
        lap("h2d row slab", t)
        t = time.perf_counter()
        cols = torch.cat([A_stage.sum(dim=0), own.sum(dim=0)])
        lap("column sums", t)
        t = time.perf_counter()
        dist.all_reduce(cols)
        norm = (scale * torch.dot(cols[:256], cols[256:])).reshape(1)
        lap("all_reduce + dot", t)
        t = time.perf_counter()
        res = rc.contract_and_score(A_stage, Bt_stage, tabs_slab, scale=scale, threshold=1e-5, k=k,
                                    method="i8", slices="auto", norm=norm, tol=tol)
        lap("contract_and_score", t)
        t = time.perf_counter()
        out = qdist.finish_block(res, k)
        lap("finish_block", t)

    def step():
        norm = qdist.stage_sharded_operands(A_pin, Bt_pin, A_stage, Bt_stage, c_lo, scale)
        res = rc.contract_and_score(A_stage, Bt_stage, tabs_slab, scale=scale, threshold=1e-5, k=k,
                                    method="i8", slices="auto", norm=norm, tol=tol)
        return qdist.finish_block(res, k)

    for _ in range(3):
        step()
    dist.barrier()
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        got = step()
    e1.record()
    dist.barrier()
    sync()
    tt = torch.tensor([e0.elapsed_time(e1) / 10], dtype=torch.float64, device=dev)
    dist.all_reduce(tt, op=dist.ReduceOp.MAX)
    med = lambda v: sorted(v[2:])[len(v[2:]) // 2]
    print(f"rank {rank}: {len(os.sched_getaffinity(0))} cores, h2d column shard {med(phases['h2d column shard']):.3f} ms "
          f"row slab {med(phases['h2d row slab']):.3f} ms", flush=True)
    dist.barrier()
    if rank == 0:
        total = 0.0
        for name, v in phases.items():
            med = sorted(v[2:])[len(v[2:]) // 2]
            total += med
            print(f"{name:22s} {med:7.3f} ms")
        print(f"{'sum of serial phases':22s} {total:7.3f} ms")
        print(f"overlapped step (max over ranks, device events) {float(tt):.3f} ms; sum_p {got[0][0]:.12f}")
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
