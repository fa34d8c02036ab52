"""GPU parity: K5 (DMMA contraction, fused MIS epilogue) and K8 (standalone reduction, top-k)
against the numpy oracle (oracle/mis_oracle.py) on seeded inputs.  Tolerance: 1e-10 max-abs on
probabilities and expectations (north_star FP64 bar); bit-exact on top-k indices and masks."""
import networkx as nx
import numpy as np
import pytest
import torch

from oracle import mis_oracle as mo

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _problem(row_bits, col_bits, kdim, seed, n_anc=0, concentrate=False):
    rng = np.random.default_rng(seed)
    nb = row_bits + col_bits
    A = rng.random((1 << row_bits, kdim)) - 0.3
    Bt = rng.random((1 << col_bits, kdim)) - 0.3
    if concentrate:
        A[rng.integers(0, 1 << row_bits)] *= 50
        Bt[rng.integers(0, 1 << col_bits)] *= 50
    bitpos = [int(b) for b in rng.permutation(nb)]
    n = nb - n_anc
    if n >= 4 and n % 2 == 0:
        G = nx.random_regular_graph(3, n, seed=int(seed))
    else:
        G = nx.path_graph(max(n, 1))
    adj = [int(a) for a in mo.adjacency_masks(G, n)] + [0] * n_anc
    node_mask = (1 << n) - 1
    return A, Bt, bitpos, adj, node_mask, n, G


def _oracle(A, Bt, bitpos, adj, n, n_anc, scale, threshold, k):
    P = scale * (A @ Bt.T)                      # united [row, col]
    u = np.arange(P.size, dtype=np.uint64)
    x = np.zeros(P.size, dtype=np.uint64)
    for b, pos in enumerate(bitpos):
        x |= ((u >> np.uint64(b)) & np.uint64(1)) << np.uint64(pos)
    full = np.zeros(P.size)
    full[x.astype(np.int64)] = P.reshape(-1)
    full /= full.sum()
    adj_n = np.array(adj[:n], dtype=np.uint64)
    s_p, s_pwok = mo.score_array(full, n, n_anc, adj_n, masked=True)
    _, s_pw = mo.score_array(full, n, n_anc, adj_n, masked=False)
    idx = np.nonzero(full > threshold)[0]
    order = np.lexsort((idx, -full[idx]))[:k]
    top = [(int(idx[i]), float(full[idx[i]])) for i in order]
    return full, s_p, s_pwok, s_pw, top


@pytest.mark.parametrize("row_bits,col_bits,kdim,n_anc,store", [
    (3, 4, 16, 0, "circuit"), (7, 7, 64, 0, "circuit"), (8, 9, 256, 2, "united"),
    (1, 1, 16, 0, "circuit"), (0, 5, 16, 0, None), (9, 8, 256, 0, None), (10, 3, 16, 1, "circuit"),
    (5, 0, 16, 0, "circuit"),
])
def test_contract_and_score_matches_oracle(cuda, row_bits, col_bits, kdim, n_anc, store):
    from qdc_b200 import recombine as rc
    A, Bt, bitpos, adj, node_mask, n, G = _problem(row_bits, col_bits, kdim, 7 + row_bits, n_anc)
    scale, thr, k = 0.5 ** 3, 1e-5, 5
    full, s_p, s_pwok, s_pw, top = _oracle(A, Bt, bitpos, adj, n, n_anc, scale, thr, k)
    tabs = rc.build_tables(row_bits, col_bits, bitpos, adj, node_mask, device=cuda)
    dA, dB = torch.from_numpy(A).to(cuda), torch.from_numpy(Bt).to(cuda)
    res = rc.contract_and_score(dA, dB, tabs, scale=scale, threshold=thr, k=k, store=store)
    out = res.get()
    assert abs(out["sum_p"] - s_p) < TOL
    assert abs(out["sum_pw_indset"] - s_pwok) < TOL * 10
    assert abs(out["sum_pw"] - s_pw) < TOL * 10
    assert [x for x, _ in out["top"]] == [x for x, _ in top]
    assert np.allclose([p for _, p in out["top"]], [p for _, p in top], atol=TOL, rtol=0)
    if store == "circuit":
        assert np.abs(res.P.cpu().numpy() - full).max() < TOL
    elif store == "united":
        P = scale * (A @ Bt.T)
        P /= P.sum()
        assert np.abs(res.P.cpu().numpy() - P).max() < TOL


def test_row_sharding_adds_up(cuda):
    """Prefix sharding: per-slab sums add to the whole; top-k merges."""
    from qdc_b200 import recombine as rc
    A, Bt, bitpos, adj, node_mask, n, G = _problem(9, 8, 64, 3)
    tabs = rc.build_tables(9, 8, bitpos, adj, node_mask, device=cuda)
    dA, dB = torch.from_numpy(A).to(cuda), torch.from_numpy(Bt).to(cuda)
    whole = rc.contract_and_score(dA, dB, tabs, k=4).get()
    parts = [rc.contract_and_score(dA, dB, tabs, k=4, row_lo=lo, n_rows=128).get()
             for lo in range(0, 512, 128)]
    for key in ("sum_p", "sum_pw_indset", "sum_pw"):
        assert abs(sum(p[key] for p in parts) - whole[key]) < 1e-12
    merged = sorted((t for p in parts for t in p["top"]), key=lambda t: (-t[1], t[0]))[:4]
    assert [x for x, _ in merged] == [x for x, _ in whole["top"]]


@pytest.mark.parametrize("dtype", [torch.float64, torch.float32])
@pytest.mark.parametrize("row_bits,col_bits", [(6, 6), (0, 1), (11, 10), (3, 0)])
def test_reduce_matches_oracle(cuda, dtype, row_bits, col_bits):
    from qdc_b200 import recombine as rc
    A, Bt, bitpos, adj, node_mask, n, G = _problem(row_bits, col_bits, 16, 11, concentrate=True)
    nb = row_bits + col_bits
    full, *_ = _oracle(A, Bt, list(range(nb)), adj, n, 0, 1.0, 1e-5, 3)
    tabs = rc.build_tables(row_bits, col_bits, list(range(nb)), adj, node_mask, device=cuda)
    P = torch.from_numpy(full).to(cuda).to(dtype)
    ref = P.double().cpu().numpy()
    s_p, s_pwok = mo.score_array(ref, n, 0, np.array(adj, dtype=np.uint64), masked=True)
    out = rc.reduce_and_score(P, tabs, threshold=1e-5, k=3).get()
    assert abs(out["sum_p"] - s_p) < TOL and abs(out["sum_pw_indset"] - s_pwok) < TOL * 10
    idx = np.nonzero(ref > 1e-5)[0]
    order = np.lexsort((idx, -ref[idx]))[:3]
    assert [x for x, _ in out["top"]] == [int(idx[i]) for i in order]


def test_indset_mask_against_python_is_indset(cuda):
    """Bit-exact: the table-driven mask/weight equal the reference's string algorithms
    (utils/graph_funcs.py:107-118, utils/helper_funcs.py:18-19) for every 12-bit string."""
    from qdc_b200 import recombine as rc
    n = 12
    G = nx.random_regular_graph(3, n, seed=5)
    adj = [int(a) for a in mo.adjacency_masks(G, n)]
    tabs = rc.build_tables(7, 5, list(range(n)), adj, (1 << n) - 1, device=cuda)
    rx = tabs.row_x.cpu().numpy(); rf = tabs.row_forb.cpu().numpy()
    rw = tabs.row_wok.cpu().numpy(); cx = tabs.col_x.cpu().numpy(); cw = tabs.col_wok.cpu().numpy()
    for u in range(1 << n):
        r, c = u >> 5, u & 31
        ok = bool((rw[r] >> 8) & (cw[c] >> 8) & 1) and (int(cx[c]) & int(rf[r])) == 0
        w = (rw[r] & 255) + (cw[c] & 255)
        s = format(u, f"0{n}b")
        assert ok == mo.is_indset(s, G) and w == mo.hamming_weight(s)
        assert int(rx[r]) | int(cx[c]) == u


@pytest.mark.parametrize("rb,cb,kdim,slices", [(7, 6, 256, 6), (8, 7, 64, 6), (5, 9, 16, 5), (9, 6, 256, 4), (8, 8, 256, 5), (7, 7, 128, 3)])
def test_int8_sliced_contraction_matches_fp64(cuda, rb, cb, kdim, slices):
    """The int8 contraction (tcgen05 int8 tensor cores, exact radix-256 slicing, all-integer scoring
    epilogue) against the FP64 kernel: the three sums within the summed stated bound
    kdim (S + 1.1) 2^(-8S) 2^(e_r + f_c) scale, the same top-k, the same numbers on a second call.
    (Per-output check: test_int8_sliced_contraction_per_output_bound.)"""
    import torch
    from qdc_b200 import recombine as rc
    rng = np.random.default_rng(rb * 100 + cb)
    M, N = (1 << rb) - 3, (1 << cb) - 2                     # ragged: not multiples of the tile
    A = rng.standard_normal((M, kdim)) * rng.uniform(1e-3, 1.0, (M, 1))
    Bt = rng.standard_normal((N, kdim)) * rng.uniform(1e-3, 1.0, (N, 1))
    G = __import__("networkx").gnm_random_graph(rb + cb, 2 * (rb + cb), seed=1)
    adj = [sum(1 << int(v) for v in G.neighbors(i)) for i in range(rb + cb)]
    tabs_full = rc.build_tables(rb, cb, list(range(rb + cb)), adj, (1 << (rb + cb)) - 1, device=cuda)
    tabs = rc.MisTables(rb, cb, tabs_full.row_x[:M].contiguous(), tabs_full.row_forb[:M].contiguous(),
                        tabs_full.row_wok[:M].contiguous(), tabs_full.col_x[:N].contiguous(),
                        tabs_full.col_wok[:N].contiguous(), tabs_full.bitpos, tabs_full.node_mask)
    A_d, B_d = torch.as_tensor(A, device=cuda), torch.as_tensor(Bt, device=cuda)
    scale = 0.25
    want = scale * (A @ Bt.T)
    thr = float(np.sort(want.reshape(-1))[-20])            # about 20 candidates above the threshold
    got = rc.contract_and_score(A_d, B_d, tabs, scale=scale, threshold=thr, k=4, normalize=False,
                                store=None, method="i8", slices=slices)
    ea = np.ceil(np.log2(np.abs(A).max(axis=1)))[:, None]
    fb = np.ceil(np.log2(np.abs(Bt).max(axis=1)))[None, :]
    bound = 2.0 ** (ea + fb + 1) * scale * kdim * (slices + 1.1) * 2.0 ** (-8 * slices)
    res = rc.contract_and_score(A_d, B_d, tabs, scale=scale, threshold=thr, k=4, normalize=False,
                                store=None, method="i8", slices=slices)
    s_got = got.get()
    assert s_got == res.get()                               # run-to-run deterministic
    ref = rc.contract_and_score(rc.pad_kdim(A_d), rc.pad_kdim(B_d), tabs, scale=scale, threshold=thr, k=4,
                                normalize=False, store=None).get()
    tot_bound = float(bound.sum()) * (rb + cb)
    for key in ("sum_p", "sum_pw", "sum_pw_indset"):
        assert abs(s_got[key] - ref[key]) <= tot_bound + 1e-9 * abs(ref[key])
    if slices >= 5:
        assert [x for x, _ in s_got["top"]] == [x for x, _ in ref["top"]]
    assert abs(s_got["sum_p"] - want.sum()) <= tot_bound + 1e-9 * abs(want.sum())


@pytest.mark.parametrize("slices", [4, 5, 6])
def test_int8_sliced_contraction_per_output_bound(cuda, slices):
    """Every output of the int8 contraction (united store) against the float64 product: the stated bound
    kdim (S + 1.1) 2^(-8S) 2^(e_r + f_c) scale holds per output, on operands whose rows span twelve
    orders of magnitude and include zero rows, negative entries and exact powers of two."""
    import torch
    from qdc_b200 import recombine as rc
    rb, cb, kdim = 8, 7, 256
    rng = np.random.default_rng(slices)
    M, N = 1 << rb, 1 << cb
    A = rng.standard_normal((M, kdim)) * 10.0 ** rng.uniform(-12, 0, (M, 1))
    Bt = rng.standard_normal((N, kdim)) * 10.0 ** rng.uniform(-6, 0, (N, 1))
    A[3] = 0.0
    A[5] = 2.0 ** -7                                      # row maximum an exact power of two
    Bt[7, :] = np.where(np.arange(kdim) % 2 == 0, 1.0, -1.0)
    G = __import__("networkx").gnm_random_graph(rb + cb, 30, seed=2)
    adj = [sum(1 << int(v) for v in G.neighbors(i)) for i in range(rb + cb)]
    tabs = rc.build_tables(rb, cb, list(range(rb + cb)), adj, (1 << (rb + cb)) - 1, device=cuda)
    scale = 1.0 / 16
    res = rc.contract_and_score(torch.as_tensor(A, device=cuda), torch.as_tensor(Bt, device=cuda), tabs,
                                scale=scale, threshold=float("inf"), k=1, normalize=False,
                                store="united", method="i8", slices=slices)
    got = res.P.cpu().numpy()
    want = scale * (A @ Bt.T)
    with np.errstate(divide="ignore"):
        ea = np.where(np.abs(A).max(axis=1) > 0, np.floor(np.log2(np.abs(A).max(axis=1))) + 1, -2000.0)[:, None]
        fb = np.floor(np.log2(np.abs(Bt).max(axis=1))) + 1
    # a column more than 2^20 below the largest of its 32-column group is sliced at that floor
    fb = np.maximum(fb, np.repeat(fb.reshape(-1, 32).max(axis=1), 32) - 20)[None, :]
    # + 1: the exponent is bumped when the scaled maximum would exceed 127 / 128
    bound = 2.0 ** (ea + fb + 2) * scale * kdim * (slices + 1.1) * 2.0 ** (-8 * slices)
    fp64_noise = 4 * kdim * 2.0 ** -53 * scale * (np.abs(A) @ np.abs(Bt).T)
    assert np.all(np.abs(got - want) <= bound + fp64_noise), float(np.max(np.abs(got - want) / (bound + fp64_noise + 1e-300)))
    assert np.all(got[3] == 0.0)
    s = res.get()
    assert abs(s["sum_p"] - want.sum()) <= bound.sum() + 1e-12 * np.abs(want).sum()
    # float32 outputs (what CutQC's build stores): integer scoring + FP32 fold of the orders + vector stores; the
    # same bound plus float32 rounding of the result (5 roundings of the fold, 2 of the scaling), same sums
    res32 = rc.contract_and_score(torch.as_tensor(A, device=cuda), torch.as_tensor(Bt, device=cuda), tabs,
                                  scale=scale, threshold=float("inf"), k=1, normalize=False,
                                  store="united", out_dtype=torch.float32, method="i8", slices=slices)
    assert res32.P.dtype == torch.float32
    got32 = res32.P.cpu().numpy().astype(np.float64)
    ulp = 2.0 ** -23 * (2.0 ** (ea + fb) * scale * kdim)          # float32 rounding relative to the largest possible |output|
    assert np.all(np.abs(got32 - want) <= bound + fp64_noise + 8 * ulp + 1e-45), float(np.max(np.abs(got32 - want) / (bound + 8 * ulp + 1e-300)))
    assert np.all(got32[3] == 0.0)
    s32 = res32.get()
    for key in ("sum_p", "sum_pw", "sum_pw_indset"):
        assert abs(s32[key] - s[key]) <= 2 * bound.sum() * (rb + cb) + 1e-12 * abs(s[key])


@pytest.mark.parametrize("rb,cb,kdim,slices", [(9, 8, 64, 6), (11, 8, 256, 5)])     # second: slabs wide enough for the CTA-pair kernel
def test_int8_contraction_row_slabs_and_given_norm(cuda, rb, cb, kdim, slices):
    """The sharded layout of bench.py's e2e leg on one GPU: every 'rank' holds only its rows of the
    row operand and divides by a norm computed from all slabs (dist.sharded_norm, world size 1 per
    call here, so the column sums are added by hand); slab sums add up to the full evaluation and
    the union of the slab candidates is the full candidate set."""
    import torch
    from qdc_b200 import recombine as rc, dist as qdist
    rng = np.random.default_rng(11)
    M, N = 1 << rb, 1 << cb
    A = torch.as_tensor(rng.random((M, kdim)) * 10.0 ** rng.uniform(-4, 0, (M, 1)), device=cuda)
    Bt = torch.as_tensor(rng.random((N, kdim)), device=cuda)
    G = __import__("networkx").gnm_random_graph(rb + cb, 24, seed=3)
    adj = [sum(1 << int(v) for v in G.neighbors(i)) for i in range(rb + cb)]
    tabs = rc.build_tables(rb, cb, list(range(rb + cb)), adj, (1 << (rb + cb)) - 1, device=cuda)
    scale, thr = 0.125, 2e-5
    full = rc.contract_and_score(A, Bt, tabs, scale=scale, threshold=thr, k=8, method="i8", slices=slices).get()
    norm = scale * torch.dot(A.sum(0), Bt.sum(0)).reshape(1)
    assert abs(float(qdist.sharded_norm(A, Bt, scale)) - float(norm)) < 1e-12 * float(norm)
    tot = {"sum_p": 0.0, "sum_pw": 0.0, "sum_pw_indset": 0.0}
    cands = []
    for lo, cnt in [qdist.row_shard(M, r, 3) for r in range(3)]:
        slab_tabs = rc.MisTables(rb, cb, tabs.row_x[lo:lo + cnt].contiguous(), tabs.row_forb[lo:lo + cnt].contiguous(),
                                 tabs.row_wok[lo:lo + cnt].contiguous(), tabs.col_x, tabs.col_wok, tabs.bitpos,
                                 tabs.node_mask)
        part = rc.contract_and_score(A[lo:lo + cnt].contiguous(), Bt, slab_tabs, scale=scale, threshold=thr, k=8,
                                     method="i8", slices=slices, norm=norm).get()
        # the same slab addressed inside the full operand (row_lo / n_rows form) must agree exactly
        part2 = rc.contract_and_score(A, Bt, tabs, scale=scale, threshold=thr, k=8, method="i8", slices=slices,
                                      row_lo=lo, n_rows=cnt).get()
        # (the given norm and the library's own differ in the last bit)
        assert abs(part["sum_pw_indset"] - part2["sum_pw_indset"]) < 1e-13
        assert [x for x, _ in part["top"]] == [x for x, _ in part2["top"]]
        assert all(abs(p - q) < 1e-15 for (_, p), (_, q) in zip(part["top"], part2["top"]))
        for key in tot:
            tot[key] += part[key]
        cands += part["top"]
    for key in tot:
        assert abs(tot[key] - full[key]) < 1e-12, (key, tot[key], full[key])
    assert abs(tot["sum_p"] - 1.0) < 1e-12
    merged = qdist.merge_topk(cands, 8)
    assert [x for x, _ in merged] == [x for x, _ in full["top"]]
    assert all(abs(p - q) < 1e-15 for (_, p), (_, q) in zip(merged, full["top"]))


@pytest.mark.parametrize("M,N,kdim,slices", [
    (20000, 200, 32, 6),     # more row blocks than SMs: CTAs take several work items; all slices fit in TMEM
    (19500, 130, 64, 5),
    (300, 9000, 256, 6),     # few row blocks, many column blocks: the column walk is split into chunks
    (129, 66, 256, 2),       # one row and two columns past a tile edge (the FP64 kernel wants an even column count)
    (1000, 1000, 128, 3),
    (37, 5000, 16, 4),
])
def test_int8_contraction_shapes(cuda, M, N, kdim, slices):
    """Work-item scheduling, TMEM-resident slices and ragged edges of the int8 contraction on shapes that
    the power-of-two configs never produce, against the FP64 kernel (sums within the stated bound,
    same candidates)."""
    import torch
    from qdc_b200 import recombine as rc
    rng = np.random.default_rng(M + N + kdim)
    A = rng.random((M, kdim)) * 10.0 ** rng.uniform(-3, 0, (M, 1))
    Bt = rng.random((N, kdim)) * 10.0 ** rng.uniform(-3, 0, (N, 1))
    f64 = dict(dtype=torch.float64, device=cuda)
    # tables by hand: arbitrary 40-bit "bitstrings", a third of the rows / columns not independent
    row_x = torch.as_tensor(rng.integers(0, 1 << 20, M) << 20, device=cuda)
    col_x = torch.as_tensor(rng.integers(0, 1 << 20, N), device=cuda)
    row_forb = torch.as_tensor(rng.integers(0, 1 << 20, M) & rng.integers(0, 1 << 20, M), device=cuda)
    row_wok = torch.as_tensor((rng.integers(0, 20, M) | ((rng.random(M) > 0.33).astype(np.int64) << 8)).astype(np.int32), device=cuda)
    col_wok = torch.as_tensor((rng.integers(0, 20, N) | ((rng.random(N) > 0.33).astype(np.int64) << 8)).astype(np.int32), device=cuda)
    tabs = rc.MisTables(20, 20, row_x, row_forb, row_wok, col_x, col_wok, tuple(range(40)), (1 << 40) - 1)
    A_d, B_d = torch.as_tensor(A, **f64), torch.as_tensor(Bt, **f64)
    want = A @ Bt.T
    top13 = np.sort(np.partition(want.reshape(-1), -13)[-13:])
    thr = float(0.5 * (top13[0] + top13[1])) / want.sum()                   # 12 candidates, none on the edge
    ref = rc.contract_and_score(rc.pad_kdim(A_d), rc.pad_kdim(B_d), tabs, threshold=thr, k=6).get()
    got = rc.contract_and_score(A_d, B_d, tabs, threshold=thr, k=6, method="i8", slices=slices).get()
    ea = np.floor(np.log2(A.max(axis=1)))[:, None] + 2
    fb = np.floor(np.log2(Bt.max(axis=1)))[None, :] + 2
    bound = float((2.0 ** (ea + fb)).sum()) * kdim * (slices + 1.1) * 2.0 ** (-8 * slices) / want.sum() * 40
    for key in ("sum_p", "sum_pw", "sum_pw_indset"):
        assert abs(got[key] - ref[key]) <= bound + 1e-12, (key, got[key], ref[key], bound)
    if slices >= 5:
        assert got["n_candidates"] == ref["n_candidates"]
        assert [x for x, _ in got["top"]] == [x for x, _ in ref["top"]]


@pytest.mark.parametrize("tol,spread", [(1e-10, 0), (1e-6, 0), (1e-10, 8), (1e-14, 0)])
def test_int8_auto_slices_keeps_its_bounds(cuda, tol, spread):
    """slices="auto": the count is chosen ON THE DEVICE (i8_plan_kernel); its a-posteriori bounds
    equal the torch restatement recombine.i8_error_bounds, are below tol unless even 6 slices cannot
    do it, one slice fewer would violate them, and the measured differences from the FP64 kernel
    stay inside the bounds."""
    import torch
    from qdc_b200 import recombine as rc
    rb, cb, kdim = 9, 8, 64
    rng = np.random.default_rng(int(-np.log10(tol)) + spread)
    M, N = 1 << rb, 1 << cb
    A = rng.random((M, kdim)) * 10.0 ** rng.uniform(-spread, 0, (M, 1))
    Bt = rng.random((N, kdim)) * 10.0 ** rng.uniform(-spread, 0, (N, 1))
    G = __import__("networkx").gnm_random_graph(rb + cb, 20, seed=5)
    adj = [sum(1 << int(v) for v in G.neighbors(i)) for i in range(rb + cb)]
    tabs = rc.build_tables(rb, cb, list(range(rb + cb)), adj, (1 << (rb + cb)) - 1, device=cuda)
    A_d, B_d = torch.as_tensor(A, device=cuda), torch.as_tensor(Bt, device=cuda)
    res = rc.contract_and_score(A_d, B_d, tabs, scale=0.25, k=4, method="i8", slices="auto", tol=tol,
                                store="united")
    ref = rc.contract_and_score(A_d, B_d, tabs, scale=0.25, k=4, store="united")
    S, eb = res.slices, res.error_bound
    assert 3 <= S <= 6 and eb["slices"] == S
    all_bounds = rc.i8_error_bounds(A_d, B_d, tabs, 0.25, ref.norm)
    assert abs(all_bounds[S]["cost"] - eb["cost"]) <= 1e-12 * eb["cost"]
    assert abs(all_bounds[S]["per_output"] - eb["per_output"]) <= 1e-12 * eb["per_output"]
    # the unstored (all-integer epilogue) form picks the same count and lands on the same cost
    res0 = rc.contract_and_score(A_d, B_d, tabs, scale=0.25, k=4, method="i8", slices="auto", tol=tol)
    assert res0.slices == S
    assert abs(res0.get()["sum_pw_indset"] - ref.get()["sum_pw_indset"]) <= eb["cost"] + 1e-13
    if S < 6 or max(all_bounds[6].values()) <= tol:
        assert eb["per_output"] <= tol and eb["cost"] <= tol
    if S > 3:
        assert max(all_bounds[S - 1].values()) > tol
    got, want = res.get(), ref.get()
    fp64_noise = 1e-13
    assert abs(got["sum_pw_indset"] - want["sum_pw_indset"]) <= eb["cost"] + fp64_noise
    assert float((res.P - ref.P).abs().max()) <= eb["per_output"] + fp64_noise


def test_int8_prepared_operands_are_reusable(cuda):
    """prepare_i8 once, contract many times: the column operand of a fragment whose tensor does not
    change between parameter sets is prepared once and reused with different row operands; results
    equal the one-call form bit for bit, and a 6-slice buffer serves a 4-slice contraction."""
    import torch
    from qdc_b200 import recombine as rc
    rb, cb, kdim = 8, 9, 64
    rng = np.random.default_rng(21)
    M, N = 1 << rb, 1 << cb
    Bt = torch.as_tensor(rng.random((N, kdim)) * 10.0 ** rng.uniform(-3, 0, (N, 1)), device=cuda)
    G = __import__("networkx").gnm_random_graph(rb + cb, 30, seed=4)
    adj = [sum(1 << int(v) for v in G.neighbors(i)) for i in range(rb + cb)]
    tabs = rc.build_tables(rb, cb, list(range(rb + cb)), adj, (1 << (rb + cb)) - 1, device=cuda)
    pb = rc.prepare_i8(Bt, tabs.col_wok, tabs.col_x, slices=6, column_role=True)
    for seed in (1, 2):
        A = torch.as_tensor(np.random.default_rng(seed).random((M, kdim)), device=cuda)
        one = rc.contract_and_score(A, Bt, tabs, scale=0.5, threshold=1e-4, k=3, method="i8", slices=6).get()
        pa = rc.prepare_i8(A, tabs.row_wok, slices=6)
        two = rc.contract_and_score(pa, pb, tabs, scale=0.5, threshold=1e-4, k=3, method="i8", slices=6).get()
        assert one == two
        ref = rc.contract_and_score(A, Bt, tabs, scale=0.5, threshold=1e-4, k=3).get()
        assert abs(one["sum_pw_indset"] - ref["sum_pw_indset"]) < 1e-12
        assert [x for x, _ in one["top"]] == [x for x, _ in ref["top"]]
    # fewer slices than stored: truncation of the balanced digits, inside the 4-slice bound
    auto = rc.contract_and_score(pa, pb, tabs, scale=0.5, threshold=1e-4, k=3, method="i8", slices="auto", tol=1e-6)
    assert auto.slices in (3, 4) and auto.error_bound["cost"] <= 1e-6
    assert abs(auto.get()["sum_pw_indset"] - ref["sum_pw_indset"]) <= auto.error_bound["cost"] + 1e-13


def test_int8_candidates_of_a_peaked_distribution(cuda):
    """The all-integer epilogue flags candidate outputs on the leading digit order alone and
    re-evaluates them exactly from the FP64 operands: on a distribution with a few hundred outputs
    above the threshold (and many just below it) the candidate count, the top-k indices and their
    probabilities equal the FP64 kernel's; zero rows, zero 16-column groups and a threshold of the
    size of the largest output are covered."""
    import torch
    from qdc_b200 import recombine as rc
    rb, cb, kdim = 9, 9, 256
    rng = np.random.default_rng(5)
    M, N = 1 << rb, 1 << cb
    A = rng.random((M, kdim)) * 1e-4
    Bt = rng.random((N, kdim)) * 1e-4
    hot_r, hot_c = rng.choice(M, 12, replace=False), rng.choice(N, 20, replace=False)
    A[hot_r] = rng.random((12, kdim)) * rng.uniform(0.5, 2.0, (12, 1))
    Bt[hot_c] = rng.random((20, kdim)) * rng.uniform(0.5, 2.0, (20, 1))
    A[40:60] = 0.0                                         # zero rows
    Bt[64:96] = 0.0                                        # two all-zero 16-column groups
    G = __import__("networkx").gnm_random_graph(rb + cb, 20, seed=6)
    adj = [sum(1 << int(v) for v in G.neighbors(i)) for i in range(rb + cb)]
    tabs = rc.build_tables(rb, cb, list(range(rb + cb)), adj, (1 << (rb + cb)) - 1, device=cuda)
    A_d, B_d = torch.as_tensor(A, device=cuda), torch.as_tensor(Bt, device=cuda)
    want = A @ Bt.T
    want /= want.sum()
    hot = np.sort(want[np.ix_(hot_r, hot_c)].reshape(-1))
    for thr in (float(0.5 * (hot[40] + hot[41])), float(0.5 * (hot[-2] + hot[-1])), float(hot[-1] * 2)):
        n_want = int((want > thr).sum())
        ref = rc.contract_and_score(A_d, B_d, tabs, threshold=thr, k=16).get()
        for slices in (4, 5, 6, "auto"):
            got = rc.contract_and_score(A_d, B_d, tabs, threshold=thr, k=16, method="i8", slices=slices).get()
            assert got["n_candidates"] == ref["n_candidates"] == n_want
            assert [x for x, _ in got["top"]] == [x for x, _ in ref["top"]]
            assert all(abs(p - q) <= 1e-15 + 1e-13 * q for (_, p), (_, q) in zip(got["top"], ref["top"]))
            assert abs(got["sum_p"] - 1.0) < 1e-12


def test_colsum_and_norm_kernels(cuda):
    """qdc_colsum (full-grid streaming column sums over two row ranges) and qdc_contract_norm against
    numpy, on even / odd contraction lengths and a misaligned base pointer."""
    import ctypes as C
    import torch
    from qdc_b200 import _lib
    lib = _lib.load()
    rng = np.random.default_rng(9)
    for rows, kdim in ((5000, 256), (777, 48), (1000, 7), (33, 512)):
        X = rng.standard_normal((rows, kdim))
        X_d = torch.as_tensor(X, device=cuda)
        n_part = lib.qdc_colsum_partials()
        part = torch.empty(n_part, kdim, dtype=torch.float64, device=cuda)
        lo, hi = rows // 5, rows // 2
        _lib.check(lib.qdc_colsum(_lib.ptr(X_d), 0, lo, hi, rows, kdim, _lib.ptr(part), _lib.stream_ptr()))
        want = X[:lo].sum(0) + X[hi:].sum(0)
        assert np.abs(part.sum(0).cpu().numpy() - want).max() < 1e-11
        Y = rng.standard_normal((300, kdim))
        Y_d = torch.as_tensor(Y, device=cuda)
        ws = torch.empty(lib.qdc_norm_workspace_doubles(kdim), dtype=torch.float64, device=cuda)
        norm = torch.empty(1, dtype=torch.float64, device=cuda)
        _lib.check(lib.qdc_contract_norm(_lib.ptr(X_d), rows, _lib.ptr(Y_d), 300, kdim, 0.25, _lib.ptr(ws),
                                         _lib.ptr(norm), _lib.stream_ptr()))
        want_n = 0.25 * float(X.sum(0) @ Y.sum(0))
        assert abs(float(norm) - want_n) < 1e-9 * max(1.0, abs(want_n))


def test_int8_flagged_list_overflow_is_reported(cuda):
    """The scoring-only int8 kernels file every output their integer filter flags and a follow-up kernel
    evaluates those exactly (i8_reeval_kernel).  With threshold 0 every output is flagged: 2^23 of them do
    not fit the 2^22-entry list, the exact test then runs inline, and because true candidates found no room
    the result says so instead of silently dropping them; with a threshold that leaves a handful of candidates
    the same operands give the FP64 kernel's candidates, and an infinite threshold flags nothing."""
    import torch
    from qdc_b200 import recombine as rc
    from qdc_b200._lib import QdcError
    rb, cb, kdim = 12, 11, 32
    rng = np.random.default_rng(3)
    A = rng.random((1 << rb, kdim)); Bt = rng.random((1 << cb, kdim))
    A[17] *= 50.0; Bt[5] *= 50.0
    tabs = rc.build_tables(rb, cb, list(range(rb + cb)), [0] * (rb + cb), (1 << (rb + cb)) - 1, device=cuda)
    A_d, B_d = torch.as_tensor(A, device=cuda), torch.as_tensor(Bt, device=cuda)
    res = rc.contract_and_score(A_d, B_d, tabs, threshold=0.0, k=1, method="i8", slices=4)
    with pytest.raises(QdcError, match="flagged"):
        res.get()
    want = A @ Bt.T
    want /= want.sum()
    thr = float(0.5 * (np.sort(want.reshape(-1))[-6] + np.sort(want.reshape(-1))[-7]))
    got = rc.contract_and_score(A_d, B_d, tabs, threshold=thr, k=6, method="i8", slices=4).get()
    ref = rc.contract_and_score(rc.pad_kdim(A_d), rc.pad_kdim(B_d), tabs, threshold=thr, k=6).get()
    assert got["n_candidates"] == ref["n_candidates"] == 6
    assert [x for x, _ in got["top"]] == [x for x, _ in ref["top"]]
    none = rc.contract_and_score(A_d, B_d, tabs, threshold=float("inf"), k=1, method="i8", slices=4)
    assert none.get()["n_candidates"] == 0 and float(none.block[9:10].view(torch.int64).item()) == 0


@pytest.mark.parametrize("slices", [3, 4, 5])
def test_int8_cta_pair_equals_single_cta_kernel(cuda, slices, monkeypatch):
    """The CTA-pair form of the contraction (cta_group::2: 256-row units, the streamed tile split between two
    CTAs, column table staged in shared memory) against the single-CTA kernel on the same prepared operands:
    integer arithmetic in another tiling, so the masked sum, the candidates and the top-k are bit-identical.
    Rows are ragged (the last 256-row unit has only its first half, and that half is partial)."""
    import torch
    from qdc_b200 import recombine as rc
    rb, cb, kdim = 11, 9, 256
    rng = np.random.default_rng(40 + slices)
    M, N = (1 << rb) - 200, 1 << cb                        # 1848 rows: 14 full units + 56 rows of a 15th
    A = rng.random((M, kdim)) * 10.0 ** rng.uniform(-3, 0, (M, 1))
    Bt = rng.random((N, kdim)) * 10.0 ** rng.uniform(-3, 0, (N, 1))
    A[rng.choice(M, 5, replace=False)] *= 300.0
    Bt[rng.choice(N, 5, replace=False)] *= 300.0
    G = __import__("networkx").gnm_random_graph(rb + cb, 25, seed=9)
    adj = [sum(1 << int(v) for v in G.neighbors(i)) for i in range(rb + cb)]
    full = rc.build_tables(rb, cb, list(range(rb + cb)), adj, (1 << (rb + cb)) - 1, device=cuda)
    tabs = rc.MisTables(rb, cb, full.row_x[:M].contiguous(), full.row_forb[:M].contiguous(), full.row_wok[:M].contiguous(),
                        full.col_x, full.col_wok, full.bitpos, full.node_mask)
    A_d, B_d = torch.as_tensor(A, device=cuda), torch.as_tensor(Bt, device=cuda)
    want = A @ Bt.T
    want /= want.sum()
    thr = float(np.sort(want.reshape(-1))[-12])
    out = {}
    for pair in ("1", "0"):
        monkeypatch.setenv("QDC_I8_PAIR", pair)
        out[pair] = rc.contract_and_score(A_d, B_d, tabs, threshold=thr, k=8, method="i8", slices=slices).get()
    assert out["1"] == out["0"]
    ref = rc.contract_and_score(A_d, B_d, tabs, threshold=thr, k=8).get()
    assert out["1"]["n_candidates"] == ref["n_candidates"] == 11
    if slices >= 4:
        assert [x for x, _ in out["1"]["top"]] == [x for x, _ in ref["top"]]
