"""BASELINE.json configs[3] (C4: 17 + 17 qubits, K = 4, 2^34 outputs) at its OWN size, on the bench's operands:
(1) sampled rows x all 2^17 columns against a numpy restatement (FP64 product, oracle/mis_oracle.py's independence
test and weights) for the headline int8 method and the FP64 kernel, with the normalisation of the full run;
(2) size-independent properties of the full 2^34-output evaluation: sum p = 1, row slabs add up to the whole, the
int8 result within its own a-posteriori bound of the FP64 kernel's.  Tolerances: 1e-12 absolute on sums that are
O(1e-3) (FP64 summation order), the stated bound for int8 vs FP64."""
import numpy as np
import pytest
import torch

from oracle import mis_oracle as mo

pytestmark = pytest.mark.gpu
RB = CB = 17


@pytest.fixture(scope="module")
def c4(cuda):
    import bench
    from qdc_b200 import recombine as rc
    n = RB + CB
    A, Bt = bench.make_operands(RB, CB)
    _, adj = bench.make_graph(n)
    bitpos = bench.make_bitpos(n, CB)
    tabs = rc.build_tables(RB, CB, bitpos, adj, (1 << n) - 1, device=cuda)
    scale = 0.5 ** bench.NUM_CUTS
    norm = scale * float(A.sum(axis=0) @ Bt.sum(axis=0))
    return dict(A=A, Bt=Bt, adj=adj, bitpos=bitpos, tabs=tabs, scale=scale, norm=norm, n=n,
                A_d=torch.from_numpy(A).to(cuda), Bt_d=torch.from_numpy(Bt).to(cuda))


def _wires(u, bitpos):
    x = np.zeros_like(u)
    for b, pos in enumerate(bitpos):
        x |= ((u >> np.uint64(b)) & np.uint64(1)) << np.uint64(pos)
    return x


@pytest.mark.parametrize("method", ["i8", "f64"])
def test_c4_sampled_rows_match_numpy(cuda, c4, method):
    from qdc_b200 import recombine as rc
    rng = np.random.default_rng(17)
    rows = np.sort(rng.choice(1 << RB, 40, replace=False)).astype(np.int64)
    A_s = np.ascontiguousarray(c4["A"][rows])
    # numpy: the normalised probabilities of those rows, their bitstrings, weights and independence
    P = (c4["scale"] / c4["norm"]) * (A_s @ c4["Bt"].T)
    u = (rows.astype(np.uint64)[:, None] << np.uint64(CB)) | np.arange(1 << CB, dtype=np.uint64)[None, :]
    x = _wires(u.reshape(-1), c4["bitpos"])
    w = mo.popcount64(x)
    ok = mo.indset_mask(x, np.array(c4["adj"], dtype=np.uint64))
    p = P.reshape(-1)
    want = (float(p.sum()), float((p * w * ok).sum()), float((p * w).sum()))
    t = c4["tabs"]
    idx = torch.as_tensor(rows, device=cuda)
    tabs_s = rc.MisTables(t.row_bits, t.col_bits, t.row_x[idx].contiguous(), t.row_forb[idx].contiguous(),
                          t.row_wok[idx].contiguous(), t.col_x, t.col_wok, t.bitpos, t.node_mask)
    norm_t = torch.tensor([c4["norm"]], dtype=torch.float64, device=cuda)
    kw = dict(method="i8", slices="auto", tol=1e-10) if method == "i8" else dict(method="f64")
    got = rc.contract_and_score(torch.from_numpy(A_s).to(cuda), c4["Bt_d"], tabs_s, scale=c4["scale"],
                                threshold=1e-5, k=2, norm=norm_t, **kw).get()
    assert abs(got["sum_p"] - want[0]) < 1e-13
    assert abs(got["sum_pw_indset"] - want[1]) < 1e-12
    assert abs(got["sum_pw"] - want[2]) < 1e-12
    assert got["top"] == [] and want[1] > 0.0


def test_c4_full_size_properties(cuda, c4):
    from qdc_b200 import recombine as rc
    args = dict(scale=c4["scale"], threshold=1e-5, k=2)
    full = rc.contract_and_score(c4["A_d"], c4["Bt_d"], c4["tabs"], method="i8", slices="auto", tol=1e-10, **args)
    whole = full.get()
    assert abs(whole["sum_p"] - 1.0) < 1e-12
    assert abs(float(full.norm.item()) - c4["norm"]) < 1e-12 * c4["norm"]
    bound = full.error_bound
    assert bound["slices"] == 5 and bound["cost"] < 1e-10
    # row slabs with the norm of the whole add up to the whole (prefix sharding), at three ragged cuts
    norm_t = full.norm.reshape(1).clone()
    cuts = [0, 1000, (1 << RB) // 3, 100_001, 1 << RB]
    parts = [rc.contract_and_score(c4["A_d"], c4["Bt_d"], c4["tabs"], method="i8", slices=5, norm=norm_t,
                                   row_lo=lo, n_rows=hi - lo, **args).get() for lo, hi in zip(cuts, cuts[1:])]
    for key in ("sum_p", "sum_pw_indset", "sum_pw"):
        assert abs(sum(p[key] for p in parts) - whole[key]) < 1e-12 * max(1.0, abs(whole[key])), key
    # the FP64 DMMA kernel on all 2^34 outputs: the int8 cost lies within its stated bound of it
    f64 = rc.contract_and_score(c4["A_d"], c4["Bt_d"], c4["tabs"], method="f64", **args).get()
    assert abs(f64["sum_pw_indset"] - whole["sum_pw_indset"]) <= bound["cost"] + 1e-13
    assert abs(f64["sum_p"] - 1.0) < 1e-12
