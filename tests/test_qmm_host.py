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


def test_maximum_likelihood_is_gpu_only():
    """No CPU fallback: a model on the CPU is refused loudly."""
    from qdc_b200 import _lib
    model = qmm.FragmentModel(torch.zeros(1, 2, 4, dtype=torch.float64), 1, 0, {"prep": [], "meas": []})
    with pytest.raises(_lib.QdcError, match="GPU only"):
        qmm.maximum_likelihood_model(model)


@pytest.mark.parametrize("p,m", [(0, 0), (1, 0), (0, 1), (1, 1), (2, 1), (2, 2), (4, 0), (0, 4)])
def test_ml_basis_tables_reproduce_the_change_of_basis(p, m):
    """qmm._ml_tables (what qdc_ml_project is given): the sparse forward table rebuilds every Choi
    block entry, the sparse inverse table every tensor component, exactly as the dense basis does."""
    rng = np.random.default_rng(10 * p + m)
    D = 2 ** (p + m)
    T = torch.as_tensor(rng.standard_normal((5, D * D)))
    fwd_k, fwd_c, inv_ij, inv_c = [x.numpy() for x in qmm._ml_tables(p, m, torch.device("cpu"))]
    want = qmm._choi_from_tensor(T, p, m).numpy()
    fc = fwd_c[..., 0] + 1j * fwd_c[..., 1]
    got = (T.numpy()[:, fwd_k] * fc[None]).sum(axis=2).reshape(5, D, D)
    assert np.abs(got - want).max() < 1e-14
    L = rng.standard_normal((5, D, D)) + 1j * rng.standard_normal((5, D, D))
    L = L + np.conj(np.swapaxes(L, 1, 2))
    want_T = qmm._tensor_from_choi(torch.as_tensor(L), p, m).numpy()
    ic = inv_c[..., 0] + 1j * inv_c[..., 1]
    got_T = (ic[None] * L[:, inv_ij & 0xff, inv_ij >> 8]).sum(axis=2).real
    assert np.abs(got_T - want_T).max() < 1e-13


def test_soft_threshold_is_the_reference_clipping_walk():
    """The reference's sequential walk over the sorted eigenvalues (:333-343, restated line by line in
    the oracle) equals lambda' = max(lambda - t0, 0) with t0 reached by Newton steps from 0 on
    h(t) = sum max(lambda - t, 0) - sum lambda -- the form the device kernel uses."""
    rng = np.random.default_rng(8)
    for trial in range(200):
        n = int(rng.integers(2, 80))
        eigs = rng.standard_normal(n) * rng.uniform(0.01, 1.0) + rng.uniform(0.0, 0.6)
        if trial % 7 == 0:
            eigs = np.abs(eigs)                                  # nothing to clip
        fake = {format(i, "b"): np.diag([v]).astype(complex) for i, v in enumerate(eigs)}
        want = np.array([qo.maximum_likelihood_model(fake).get(format(i, "b"), np.zeros((1, 1)))[0, 0].real
                         for i in range(n)])
        t, total, steps = 0.0, eigs.sum(), 0
        for steps in range(64):
            pos = eigs > t
            h = (eigs[pos] - t).sum() - total
            if not h > 1e-16 * np.abs(eigs).sum() or pos.sum() == 0:
                break
            t += h / pos.sum()
        got = np.maximum(eigs - t, 0.0)
        assert np.abs(got - want).max() < 1e-12 and steps < 12


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
    # a column more than 2^20 below the largest of its 32-column group is sliced at that floor
    Bt2 = torch.tensor([[1.5, 0.0, 0.0, 0.2], [2e-9, 1e-9, 0.0, 0.0]], dtype=torch.float64)
    b2 = rc.i8_error_bounds(A, Bt2, tabs, 0.5, norm)
    assert abs(b2[5]["cost"] - 3 * 0.25 * (1 * (2 + 2.0 ** -19) * g5 + 1 * 2 * fold * 2.0 ** -40)) < 1e-25
    assert set(b) == {2, 3, 4, 5, 6}


def test_ckt_summation_terms_numbering_matches_restatement():
    """Host bookkeeping of `cutqc.ckt.generate_summation_terms` (digit-tuple lookup) == the oracle's
    label-string restatement of wire_cutting_post_processing.py:225-337, dict for dict, including
    the first-appearance numbering of entries and instances; a wrong num_cuts is refused."""
    import cutqc_cases as cc
    from oracle import cutqc_oracle as co
    from qdc_b200.cutqc import ckt
    for name in cc.CASES:
        circ, frags, wpm = cc.CASES[name]()
        cpm = co.complete_path_map_from_wire_path_map(wpm)
        K = sum(len(p) - 1 for p in cpm.values())
        got = ckt.generate_summation_terms(frags, cpm, K)
        want = co.ckt_generate_summation_terms(frags, cpm, K)
        assert got == want
        for s in want[2]:
            assert list(got[2][s].items()) == list(want[2][s].items())
        with pytest.raises(ValueError):
            ckt.generate_summation_terms(frags, cpm, K + 1)


def test_get_dirname_spelling():
    from qdc_b200.cutqc.helper_fun import get_dirname
    assert get_dirname("bv", 10, None, None, None, None, "cutter") == "./cutqc_data/bv/cc_10"
    assert get_dirname("bv", 10, None, "sv", None, None, "evaluator") == "./cutqc_data/bv/cc_10/sv"
    assert get_dirname("bv", 10, 1, "sv", None, None, "vertical_collapse") == "./cutqc_data/bv/cc_10/sv/et_1"
    assert get_dirname("bv", 10, 1, "sv", 4, 20, "build") == "./cutqc_data/bv/cc_10/sv_1_20_4"
    assert get_dirname("bv", 10, None, None, 4, 20, "runtime") == "./runtime/bv/cc_10/q_20_4"
    with pytest.raises(Exception):
        get_dirname("bv", 10, None, None, None, None, "nope")


def test_top_strings_merges_ancilla_partners_before_threshold_and_cutoff():
    """mis.py:229-242 applies threshold and cutoff AFTER strip_ancillas (which sums entries that
    differ only in ancilla bits); the fused device list ranks full-width strings, so equal node
    strings are merged first - checked against the reference order of operations on a distribution
    with weight on ancilla = 1."""
    import networkx as nx
    from qdc_b200 import mis
    from oracle import mis_oracle as mo
    G = nx.path_graph(3)                                  # independent sets: 000 001 010 100 101
    probs = {"0101": 0.30, "1101": 0.25, "0010": 0.35, "1010": 0.05, "0100": 0.05}   # ancilla = first character

    class Score:
        def __init__(self, top, n_cand):
            self.top_x = [x for x, _ in top]
            self._r = {"top": top, "n_candidates": n_cand}

        def get(self):
            return self._r

    class Dist(dict):
        n_wires = 4
        P = object()

    full = sorted(((int(k, 2), v) for k, v in probs.items()), key=lambda t: -t[1])
    want = mo.better_strs(mo.top_probs(mo.strip_ancillas(probs, 1), threshold=1e-5, cutoff=1), G, "000")
    assert want == [("101", 2)]                           # 0.30 + 0.25 beats 0.35 + 0.05 only after merging
    d = Dist(probs); d.score = Score(full, len(full))     # every candidate returned: merge is complete
    assert mis.top_strings(d, G, cutoff=1) == want
    d.score = Score(full[:2], 5)                          # truncated list (0010, 0101): the gap does not cover the
    assert mis.top_strings(d, G, cutoff=1) == want        # unseen partners -> falls back to the dict path
    d.score = Score(full[:1], 5)
    d.P = None                                            # nothing stored: best effort from the fused list
    assert mis.top_strings(d, G, cutoff=1) == [("010", 1)]


def test_streamed_combine_blocks_rescales_and_refilters():
    """Host side of streamed.StreamedContraction: pair blocks computed with a norm HINT are rescaled to the true
    norm (hint x total of the pairs' sum p), candidates re-tested against the exact threshold, and a hint more
    than twice the norm (or a pair out of candidate room) refuses the speculation."""
    import numpy as np
    from qdc_b200.streamed import combine_blocks, SLACK
    k, thr, hint, true = 2, 0.1, 2.0, 1.6                        # true norm = 0.8 x hint
    def block(sum_p, cost, top, n_cand=2, lost=0, slices=5):
        b = np.zeros(16 + 2 * k)
        b[0:3] = [sum_p, cost, 2 * cost]
        b[4] = slices
        b[7:8].view(np.int64)[0] = n_cand
        b[10:11].view(np.int64)[0] = lost
        xs = [x for x, _ in top] + [-1] * (k - len(top))
        ps = [p for _, p in top] + [float("-inf")] * (k - len(top))
        b[16:16 + k].view(np.int64)[:] = xs
        b[16 + k:] = ps
        return b
    # probabilities as the kernels report them (divided by the hint): exact value = reported x hint / true
    blocks = np.stack([block(0.5, 0.05, [(7, 0.30), (3, 0.09)]), block(0.3, 0.01, [(9, 0.07)], n_cand=1, slices=4)])
    out = combine_blocks(blocks, [8, 8], hint, thr, k)
    f = hint / true
    assert out["streamed"] and abs(out["norm"] - true) < 1e-15
    assert abs(out["sum_p"] - 1.0) < 1e-15 and abs(out["sum_pw_indset"] - 0.06 * f) < 1e-15
    assert [x for x, _ in out["top"]] == [7, 3]                  # 0.09 f = 0.1125 > thr stays, 0.07 f = 0.0875 goes
    assert abs(out["top"][0][1] - 0.30 * f) < 1e-15 and out["slices"] == [5, 4] and out["n_candidates"] == 3
    low = np.stack([block(0.25, 0.02, []), block(0.15, 0.01, [])])                       # pairs add up to 0.4 = norm / hint
    assert combine_blocks(low, [8, 8], hint, thr, k) == {"streamed": False, "norm": 0.8}   # norm < SLACK x hint
    assert SLACK == 0.5
    assert combine_blocks(blocks, [1, 8], hint, thr, k)["streamed"] is False            # candidate room exceeded
    lost = np.stack([block(0.8, 0.0, [], n_cand=0, lost=3)])
    assert combine_blocks(lost, [8], hint, thr, k) == {"streamed": False, "norm": 1.6}


def test_piece_boundaries():
    from qdc_b200.streamed import _splits
    assert _splits(64, (1, 1, 2, 4)) == [(0, 8), (8, 8), (16, 16), (32, 32)]
    assert _splits(64, 4) == [(0, 16), (16, 16), (32, 16), (48, 16)]
    assert _splits(10, 2) == [(0, 5), (5, 5)]
    assert [n % 8 for _, n in _splits(1 << 17, (1, 1, 2, 4), align=8)] == [0] * 4
    for n, parts in [(100, (1, 1, 2)), (7, 4), (1 << 17, (1, 1, 2, 4, 8))]:
        sp = _splits(n, parts)
        assert sp[0][0] == 0 and sum(c for _, c in sp) == n and all(a + c == b for (a, c), (b, _) in zip(sp, sp[1:]))


def test_bind_host_near_gpu_is_a_no_op_without_a_gpu(monkeypatch):
    from qdc_b200 import dist as qd
    import torch
    if not torch.cuda.is_available():
        assert qd.bind_host_near_gpu(0) == []
    monkeypatch.setenv("QDC_BIND_NUMA", "0")
    assert qd.bind_host_near_gpu(0) == []
