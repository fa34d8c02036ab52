"""GPU parity of `streamed.StreamedContraction` (host operands copied in pieces under the contraction, normalisation
speculated from a hint and verified) against the unstreamed `recombine.contract_and_score` on the same operands and
against the numpy product: the three sums to 1e-13 relative, the same top-k indices, probabilities to 1e-13 relative;
a hint that is too large, or none, must take the ordinary path and still give the same answer."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _setup(cuda, rb=11, cb=10, kdim=256, seed=3):
    import networkx as nx
    from qdc_b200 import recombine as rc
    rng = np.random.default_rng(seed)
    M, N = 1 << rb, 1 << cb
    A = rng.random((M, kdim)) * 10.0 ** rng.uniform(-3, 0, (M, 1))
    Bt = rng.random((N, kdim)) * 10.0 ** rng.uniform(-3, 0, (N, 1))
    A[rng.choice(M, 6, replace=False)] *= 400.0
    Bt[rng.choice(N, 6, replace=False)] *= 400.0
    G = nx.gnm_random_graph(rb + cb, 28, seed=seed)
    adj = [sum(1 << int(v) for v in G.neighbors(i)) for i in range(rb + cb)]
    bitpos = [int(b) for b in rng.permutation(rb + cb)]
    tabs = rc.build_tables(rb, cb, bitpos, adj, (1 << (rb + cb)) - 1, device=cuda)
    want = A @ Bt.T
    want /= want.sum()
    top = np.sort(want.reshape(-1))
    thr = float(0.5 * (top[-20] + top[-21]))                  # 20 entries above it, none on it
    return A, Bt, tabs, thr


@pytest.mark.parametrize("chunks", [(4, 4), (1, 1), ((1, 1, 2), 2)])
def test_streamed_equals_unstreamed(cuda, chunks):
    from qdc_b200 import recombine as rc
    from qdc_b200.streamed import StreamedContraction
    A, Bt, tabs, thr = _setup(cuda)
    scale, k = 0.25, 8
    ref = rc.contract_and_score(torch.as_tensor(A, device=cuda), torch.as_tensor(Bt, device=cuda), tabs, scale=scale,
                                threshold=thr, k=k, method="i8", slices="auto").get()
    norm = scale * float(A.sum(axis=0) @ Bt.sum(axis=0))
    A_pin, Bt_pin = torch.from_numpy(A).pin_memory(), torch.from_numpy(Bt).pin_memory()
    sc = StreamedContraction(tabs, A.shape[1], row_chunks=chunks[0], col_chunks=chunks[1], k=k)

    def same(out):
        for key in ("sum_p", "sum_pw_indset", "sum_pw"):
            assert abs(out[key] - ref[key]) <= 1e-13 * max(1.0, abs(ref[key])), key
        assert [x for x, _ in out["top"]] == [x for x, _ in ref["top"]]
        assert np.allclose([p for _, p in out["top"]], [p for _, p in ref["top"]], rtol=1e-13, atol=0)
        assert abs(out["norm"] - norm) <= 1e-12 * abs(norm)

    first = sc.run(A_pin, Bt_pin, scale=scale, threshold=thr)            # no hint: the ordinary path
    assert first["streamed"] is False
    same(first)
    for hint in (first["norm"], first["norm"] * 1.9, first["norm"] * 0.7):
        out = sc.run(A_pin, Bt_pin, scale=scale, threshold=thr, norm_hint=hint)
        assert out["streamed"] is True and len(out["slices"]) == sc.R * sc.C
        same(out)
        assert out["n_candidates"] >= ref["n_candidates"]
    exact = sc.run(A_pin, Bt_pin, scale=scale, threshold=thr, norm_hint=first["norm"])
    assert exact["n_candidates"] >= ref["n_candidates"] == 20
    before = sc.fallbacks
    out = sc.run(A_pin, Bt_pin, scale=scale, threshold=thr, norm_hint=first["norm"] * 2.5)   # norm < hint / 2
    assert out["streamed"] is False and sc.fallbacks == before + 1
    same(out)


def test_streamed_new_operands_every_step(cuda):
    """The staging buffers are reused: a second evaluation with other operands (and the previous norm as the hint)
    must not see anything of the first."""
    from qdc_b200 import recombine as rc
    from qdc_b200.streamed import StreamedContraction
    A, Bt, tabs, thr = _setup(cuda, seed=5)
    sc = StreamedContraction(tabs, A.shape[1], row_chunks=2, col_chunks=4, k=4)
    hint = None
    rng = np.random.default_rng(0)
    for step in range(3):
        A2 = A * rng.uniform(0.9, 1.1, A.shape)
        A_pin, Bt_pin = torch.from_numpy(A2).pin_memory(), torch.from_numpy(Bt).pin_memory()
        out = sc.run(A_pin, Bt_pin, scale=1.0, threshold=thr, norm_hint=hint)
        ref = rc.contract_and_score(torch.as_tensor(A2, device=cuda), torch.as_tensor(Bt, device=cuda), tabs,
                                    threshold=thr, k=4, method="i8", slices="auto").get()
        assert out["streamed"] is (step > 0)
        assert abs(out["sum_pw_indset"] - ref["sum_pw_indset"]) <= 1e-13 * max(1.0, abs(ref["sum_pw_indset"]))
        assert [x for x, _ in out["top"]] == [x for x, _ in ref["top"]]
        hint = out["norm"]
