"""e2e step time of streamed.StreamedContraction at C4 for several piece counts against the plain e2e step
(copy both operands, then contract_and_score).  Under torchrun for N > 1."""
import importlib
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

importlib.import_module("dqva-and-circuit-cutting_b200")
from qdc_b200 import dist as qdist, recombine as rc  # noqa: E402
from qdc_b200.streamed import StreamedContraction  # noqa: E402


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
    row_lo, n_rows = qdist.row_shard(1 << rb, rank, ws)
    tabs = rc.build_tables(rb, cb, bitpos, adj, (1 << n) - 1, device=dev)
    A_pin = torch.from_numpy(A_h).pin_memory()[row_lo:row_lo + n_rows]
    Bt_pin = torch.from_numpy(Bt_h).pin_memory()

    def timed(fn, steps=10, warm=3):
        for _ in range(warm):
            fn()
        if ws > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            out = fn()
        e1.record()
        if ws > 1:
            dist.barrier()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / steps], dtype=torch.float64, device=dev)
        if ws > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t), out

    G4, G3 = (1, 1, 2, 4), (1, 1, 2)
    configs = [(1, 1), (2, 4), (G4, 4), (G3, 4), (G4, 2), ((1, 3), 4), ((1, 1, 2), 2), ((1, 3, 4), 4), (2, (1, 1, 2)), (2, 8)] \
        if ws == 1 else [(1, 1), (2, 2), (2, 1), (4, 1), (1, 2), (2, 4)]
    for slices in ("auto",):
        for R, C in configs:
            sc = StreamedContraction(tabs, bench.KDIM, row_chunks=R, col_chunks=C, k=8, device=dev)
            hint = {"n": None}

            def step():
                out = sc.run(A_pin, Bt_pin, scale=scale, threshold=1e-5, norm_hint=hint["n"], slices=slices, tol=1e-10)
                hint["n"] = out["norm"]
                return out
            t, out = timed(step)
            if rank == 0:
                print(f"N={ws} slices={slices} pieces {R} x {C}: {t:.3f} ms per step, fallbacks {sc.fallbacks}, "
                      f"slices used {sorted(set(out['slices']))}, sum_p {out['sum_p']:.12f} cost {out['sum_pw_indset']:.12f}",
                      flush=True)
            del sc
    if ws > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
