"""Host logic of qdc_b200.qsplit.qsplit_mlrecon_methods against the oracle restatement:
wire bookkeeping (bit-exact), the united-bit -> circuit-wire map that replaces
`united_axis_permutation`, the Pauli <-> Choi change of basis behind the dict views, and the
closed-form eigenvalue clipping of `maximum_likelihood_model` (run on CPU tensors here: these
functions do torch algebra on whatever device their input lives on)."""
import networkx as nx
import numpy as np
import pytest
import torch

from oracle import qsplit_oracle as qo
from qdc_b200.qsplit import qsplit_mlrecon_methods as qmm
from test_oracle_golden import _cut_dqva, _ghz3
from qdc_b200.qsplit import qsplit_circuit_cutter as qcc


def _scenarios():
    circ, cuts = _ghz3()
    yield qcc.cut_circuit(circ, cuts)
    G = nx.Graph()
    G.add_edges_from([(0, 1), (1, 2), (2, 0), (2, 3), (3, 4), (4, 5), (5, 3), (1, 4)])
    for max_cuts, seed in [(1, 0), (2, 1), (2, 5)]:
        _, _, frags, wpm = _cut_dqva(G, [[0, 1, 2], [3, 4, 5]], max_cuts, seed, True)
        yield frags, wpm
    G3 = nx.Graph()
    G3.add_edges_from([(0, 1), (1, 2), (2, 3), (3, 4), (4, 5), (5, 6), (6, 7), (7, 8), (0, 2),
                       (3, 5), (6, 8)])
    _, _, frags, wpm = _cut_dqva(G3, [[0, 1, 2], [3, 4, 5], [6, 7, 8]], 2, 3, True)
    yield frags, wpm


def test_wire_bookkeeping_matches_oracle():
    for frags, wpm in _scenarios():
        assert qmm.identify_stitches(wpm) == qo.identify_stitches(wpm)
        assert qmm.identify_frag_targets(wpm) == qo.identify_frag_targets(wpm)
        assert qmm.fragment_cuts(wpm) == qo.fragment_cuts(wpm)
        assert qmm.fragment_variants(wpm) == qo.fragment_variants(wpm)
        assert qmm.united_axis_permutation(wpm) == qo.united_axis_permutation(wpm)
        assert qmm.get_frag_num(wpm) == qo.get_frag_num(wpm) == len(frags)


def test_united_bit_wires_is_the_axis_permutation():
    """final char j = joined char perm[j]; joined char c <-> united bit N-1-c; final char j <->
    circuit bit N-1-j  =>  united bit u feeds circuit bit bitpos[u]."""
    for frags, wpm in _scenarios():
        bitpos, widths = qmm.united_bit_wires(wpm)
        perm = qo.united_axis_permutation(wpm)
        N = len(perm)
        assert sorted(bitpos) == list(range(N)) and sum(widths) == N
        for j in range(N):
            assert bitpos[N - 1 - perm[j]] == N - 1 - j


def test_label_matrices_match_oracle():
    for lab in ["I", "X", "Y", "Z", "Zp", "Zm", "Xp", "Xm", "Yp", "Ym", "S0", "S1", "S2", "S3"]:
        assert np.abs(qmm.label_to_matrix(lab) - qo.label_to_matrix(lab)).max() < 1e-15
    t = (("S1", "S3"), ("Xm", "Zp", "Yp"))
    assert np.abs(qmm.target_labels_to_matrix(t) - qo.target_labels_to_matrix(t)).max() < 1e-15


def test_choi_view_of_pauli_tensor_matches_oracle_lstsq():
    """The dict view of a FragmentModel (Pauli tensor -> Choi blocks) equals the reference's
    per-bitstring least-squares fit on the same exact data."""
    for frags, wpm in list(_scenarios())[:3]:
        ft = qo.identify_frag_targets(wpm)
        data = qo.collect_fragment_data(frags, wpm)
        want = qo.direct_fragment_model(data)
        for i, frag in enumerate(frags):
            R = qo.fragment_pauli_tensor(frag, ft[i]["prep"], ft[i]["meas"])
            p, m = len(ft[i]["prep"]), len(ft[i]["meas"])
            T = torch.as_tensor(R.reshape(R.shape[0], -1)).unsqueeze(0)
            model = qmm.FragmentModel(T, p, m, None)
            assert set(model.keys()) == set(want[i].keys())
            for bits, block in want[i].items():
                assert np.abs(model[bits] - block).max() < 1e-10
            # and back
            blocks = model.choi_blocks()
            T2 = qmm._tensor_from_choi(blocks, p, m)
            assert (T2 - T[0]).abs().max() < 1e-12


@pytest.mark.parametrize("noise", [0.0, 0.05])
def test_maximum_likelihood_matches_oracle(noise):
    """Closed-form clipping == the reference's sequential loop, including on non-PSD (noisy)
    blocks; compared through the Choi dict view."""
    rng = np.random.default_rng(4)
    frags, wpm = list(_scenarios())[2]
    ft = qo.identify_frag_targets(wpm)
    for i, frag in enumerate(frags):
        R = qo.fragment_pauli_tensor(frag, ft[i]["prep"], ft[i]["meas"])
        p, m = len(ft[i]["prep"]), len(ft[i]["meas"])
        T = R.reshape(R.shape[0], -1)
        T = T + noise * rng.standard_normal(T.shape) * (np.abs(T).max(axis=1, keepdims=True) > 0)
        model = qmm.FragmentModel(torch.as_tensor(T).unsqueeze(0), p, m, {"prep": [], "meas": []})
        direct = {k: v for k, v in model.items()}
        want = qo.maximum_likelihood_model(direct)
        got = qmm.maximum_likelihood_model(model)
        assert set(got.keys()) == set(want.keys())
        for bits in want:
            assert np.abs(got[bits] - want[bits]).max() < 1e-10


def test_contraction_spec_parsing(monkeypatch):
    """recombine.resolve_contraction: 'f64' | 'i8' | 'i8:S', environment default, loud errors."""
    from qdc_b200 import recombine as rc
    monkeypatch.delenv("QDC_CONTRACTION", raising=False)
    assert rc.resolve_contraction(None) == ("f64", 6)
    assert rc.resolve_contraction("i8") == ("i8", 6)
    assert rc.resolve_contraction("i8:5") == ("i8", 5)
    monkeypatch.setenv("QDC_CONTRACTION", "i8:4")
    assert rc.resolve_contraction(None) == ("i8", 4)
    for bad in ("i8:7", "i8:1", "i8x", "tf32", ""):
        with pytest.raises(ValueError):
            rc.resolve_contraction(bad)


def test_i8_error_bounds_formula():
    """recombine.i8_error_bounds on the CPU (pure torch mirror of the device plan kernel): the
    per-output bound is 2^(e_r + f_c) scale / norm kdim (S + 1.1) 2^(-8S) at the largest rows, the
    cost bound only counts rows / columns that are independent sets themselves (plus the rounding
    of the epilogue's 16-column sums), both shrink 256-fold per slice."""
    from qdc_b200 import recombine as rc
    A = torch.tensor([[0.3, -0.7, 0.0, 0.1], [0.01, 0.02, 0.0, 0.0], [0.0, 0.0, 0.0, 0.0]], dtype=torch.float64)
    Bt = torch.tensor([[1.5, 0.0, 0.0, 0.2], [0.2, 0.1, 0.0, 0.0]], dtype=torch.float64)
    row_wok = torch.tensor([256 | 2, 1, 256], dtype=torch.int32)      # rows 0 and 2 independent
    col_wok = torch.tensor([256 | 1, 256], dtype=torch.int32)
    tabs = rc.MisTables(2, 1, None, None, row_wok, None, col_wok, (0, 1, 2), 0b111)
    norm = torch.tensor([2.0], dtype=torch.float64)
    b = rc.i8_error_bounds(A, Bt, tabs, 0.5, norm)
    g5 = 4 * 6.1 * 2.0 ** -40
    # exponents the slicing applies (|x| 2^(7 - e) <= 127): 0.7 -> 2^0, 0.02 -> 2^-5, 1.5 -> 2^1, 0.2 -> 2^-2
    assert abs(b[5]["per_output"] - 1 * 2 * 0.25 * g5) < 1e-25
    fold = 0.5 * 1 * 2.0 ** -26                                       # one 16-column group
    assert abs(b[5]["cost"] - 3 * 0.25 * ((1 + 0) * (2 + 0.25) * g5 + 1 * 2 * fold * 2.0 ** -40)) < 1e-25
    assert abs(b[4]["per_output"] / b[5]["per_output"] - 256 * 5.1 / 6.1) < 1e-9
    # a column more than 2^20 below the largest of its 16-column group is sliced at that floor
    Bt2 = torch.tensor([[1.5, 0.0, 0.0, 0.2], [2e-9, 1e-9, 0.0, 0.0]], dtype=torch.float64)
    b2 = rc.i8_error_bounds(A, Bt2, tabs, 0.5, norm)
    assert abs(b2[5]["cost"] - 3 * 0.25 * (1 * (2 + 2.0 ** -19) * g5 + 1 * 2 * fold * 2.0 ** -40)) < 1e-25
    assert set(b) == {2, 3, 4, 5, 6}
