"""world_size-2 gloo test of the multi-GPU finish (the only collective on the path) and of the
prefix / round-robin sharding helpers."""
import os
import socket

import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from qdc_b200 import dist as qdist


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, ws, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(ws), LOCAL_RANK=str(rank))
    import importlib, sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    importlib.import_module("dqva-and-circuit-cutting_b200")
    from qdc_b200 import dist as qd
    r, w, _ = qd.init_from_env(backend="gloo")
    k = 3
    sums = torch.tensor([0.25 * (rank + 1), 1.0 * rank, 2.0], dtype=torch.float64)
    top_x = torch.tensor([10 + rank, 20 + rank, -1], dtype=torch.int64)
    top_p = torch.tensor([0.5 - 0.1 * rank, 0.3, float("-inf")], dtype=torch.float64)
    total, top = qd.finish(sums, top_x, top_p, k)
    out[rank] = (total.tolist(), top)
    dist.destroy_process_group()


def test_finish_world_size_2():
    ws = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker, args=(ws, _free_port(), out), nprocs=ws, join=True)
    for rank in range(ws):
        total, top = out[rank]
        assert total == [0.75, 1.0, 4.0]
        assert top == [(10, 0.5), (11, 0.4), (20, 0.3)]


def test_sharding_helpers():
    for n, ws in [(1 << 17, 8), (10, 4), (3, 8), (7, 1)]:
        slabs = [qdist.row_shard(n, r, ws) for r in range(ws)]
        assert slabs[0][0] == 0 and sum(c for _, c in slabs) == n
        for (lo, c), (lo2, _) in zip(slabs, slabs[1:]):
            assert lo + c == lo2
    items = sorted(i for r in range(3) for i in qdist.round_robin(10, r, 3))
    assert items == list(range(10))
    assert qdist.merge_topk([(5, 0.2), (3, 0.2), (9, 0.7)], 2) == [(9, 0.7), (3, 0.2)]


def _worker_sets(rank, ws, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(ws), LOCAL_RANK=str(rank))
    import importlib, sys
    import numpy as np
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    importlib.import_module("dqva-and-circuit-cutting_b200")
    from qdc_b200 import dist as qd
    qd.init_from_env(backend="gloo")
    table = np.arange(7 * 3, dtype=float).reshape(7, 3)
    seen = []

    def fn(rows):
        seen.append(len(rows))
        return rows.sum(axis=1) * 2.0
    res = qd.sharded_over_parameter_sets(fn, table)
    out[rank] = (res.tolist(), seen)
    dist.destroy_process_group()


def test_parameter_sets_sharded_world_size_2():
    """7 parameter sets over 2 ranks: 4 + 3 evaluated locally, every rank ends with all 7."""
    ws = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker_sets, args=(ws, _free_port(), out), nprocs=ws, join=True)
    want = [2.0 * sum(range(3 * i, 3 * i + 3)) for i in range(7)]
    assert out[0][0] == want and out[1][0] == want
    assert out[0][1] == [4] and out[1][1] == [3]


def _worker_norm(rank, ws, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(ws), LOCAL_RANK=str(rank))
    import importlib, sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    importlib.import_module("dqva-and-circuit-cutting_b200")
    from qdc_b200 import dist as qd
    qd.init_from_env(backend="gloo")
    g = torch.Generator().manual_seed(5)
    A = torch.randn(37, 16, dtype=torch.float64, generator=g)
    Bt = torch.randn(21, 16, dtype=torch.float64, generator=g)
    lo, cnt = qd.row_shard(A.shape[0], rank, ws)
    norm = qd.sharded_norm(A[lo:lo + cnt], Bt, 0.25)
    out[rank] = (float(norm.item()), float(0.25 * (A @ Bt.T).sum().item()))
    dist.destroy_process_group()


def test_sharded_norm_world_size_2():
    """Every rank holds only its row slab; the all-reduced column sums give the total of the
    whole product on every rank."""
    ws = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker_norm, args=(ws, _free_port(), out), nprocs=ws, join=True)
    for rank in range(ws):
        got, want = out[rank]
        assert abs(got - want) < 1e-12 * max(1.0, abs(want))
    assert out[0][0] == out[1][0]


def _worker_stage(rank, ws, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(ws), LOCAL_RANK=str(rank))
    import importlib, sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    importlib.import_module("dqva-and-circuit-cutting_b200")
    from qdc_b200 import dist as qd
    qd.init_from_env(backend="gloo")
    g = torch.Generator().manual_seed(11)
    A = torch.randn(38, 16, dtype=torch.float64, generator=g)
    Bt = torch.randn(24, 16, dtype=torch.float64, generator=g)
    lo, cnt = qd.row_shard(A.shape[0], rank, ws)
    c_lo, c_n = qd.row_shard(Bt.shape[0], rank, ws)
    A_stage = torch.empty(cnt, 16, dtype=torch.float64)
    Bt_stage = torch.full_like(Bt, float("nan"))
    norm = qd.stage_sharded_operands(A[lo:lo + cnt], Bt[c_lo:c_lo + c_n], A_stage, Bt_stage, c_lo, 0.25)
    out[rank] = (float(norm.item()), float(0.25 * (A @ Bt.T).sum().item()),
                 bool(torch.equal(Bt_stage, Bt)), bool(torch.equal(A_stage, A[lo:lo + cnt])))
    dist.destroy_process_group()


def test_stage_sharded_operands_world_size_2():
    """Each rank brings its row slab and half of the column operand; afterwards every rank holds the
    whole column operand (all-gather) and the norm of the whole product (one all-reduce of the
    column sums of what each rank brought)."""
    ws = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_worker_stage, args=(ws, _free_port(), out), nprocs=ws, join=True)
    for rank in range(ws):
        got, want, bt_ok, a_ok = out[rank]
        assert bt_ok and a_ok
        assert abs(got - want) < 1e-12 * max(1.0, abs(want))
    assert out[0][0] == out[1][0]


def _block_worker(rank, ws, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank),
                      WORLD_SIZE=str(ws), LOCAL_RANK=str(rank))
    import importlib, sys
    import numpy as np
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    importlib.import_module("dqva-and-circuit-cutting_b200")
    from qdc_b200 import dist as qd, recombine as rc
    qd.init_from_env(backend="gloo")
    k = 3
    # the int8 path's result block: [3 sums | norm | slices | 2 bounds | candidate count | ... | k indices | k probs]
    h = np.zeros(16 + 2 * k)
    h[0:3] = [0.25 * (rank + 1), 1.0 * rank, 2.0]
    h[7:8].view(np.int64)[0] = 2
    h[16:16 + k].view(np.int64)[:] = [(1 << 40) + rank, 20 + rank, -1]
    h[16 + k:16 + 2 * k] = [0.5 - 0.1 * rank, 0.3, float("-inf")]
    block = torch.from_numpy(h)
    res = rc.ScoreResult(block[0:3], block[16:16 + k].view(torch.int64), block[16 + k:], block[7:8].view(torch.int64),
                         cand_cap=8, block=block)
    total, top = qd.finish_block(res, k)
    pending = qd.finish_block(res, k, wait=False)              # read-back deferred: same answer
    total2, top2 = pending.get()
    assert total2.tolist() == total.tolist() and top2 == top
    out[rank] = (total.tolist(), top)
    dist.destroy_process_group()


def test_finish_block_world_size_2():
    """dist.finish_block: the whole result block is all-gathered, 64-bit indices survive the trip."""
    ws = 2
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(_block_worker, args=(ws, _free_port(), out), nprocs=ws, join=True)
    for rank in range(ws):
        total, top = out[rank]
        assert total == [0.75, 1.0, 4.0]
        assert top == [(1 << 40, 0.5), ((1 << 40) + 1, 0.4), (20, 0.3)]
