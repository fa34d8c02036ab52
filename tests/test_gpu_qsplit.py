"""GPU parity of the qsplit path end to end (collect -> model -> recombine -> MIS score) through
the reference-named API, against the oracle restatement and the uncut statevector.  1e-10 max-abs
on probabilities / costs, bit-exact keys and top-k bitstrings."""
import networkx as nx
import numpy as np
import pytest
import torch

from oracle import mis_oracle as mo
from oracle import qsplit_oracle as qo
from oracle import statevector as sv
from test_oracle_golden import _cut_dqva, _ghz3

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _be(cuda, **kw):
    from qdc_b200.engine import B200Backend
    return B200Backend(device=cuda, **kw)


@pytest.mark.parametrize("mode", ["direct", "likely"])
def test_reference_test_cutting_known_answer(cuda, mode):
    """test_cutting.py:7-40 of the reference: {000: 0.5, 111: 0.5}."""
    from qdc_b200.qsplit import qsplit_circuit_cutter as qcc
    from qdc_b200.utils.cutting_funcs import sim_with_cutting
    circ, cuts = _ghz3()
    frags, wpm = qcc.cut_circuit(circ, cuts)
    dist = sim_with_cutting(frags, wpm, 1000, _be(cuda), mode=mode)
    assert abs(dist["000"] - 0.5) < TOL and abs(dist["111"] - 0.5) < TOL
    assert all(abs(v) < TOL for k, v in dist.items() if k not in ("000", "111"))
    with pytest.raises(Exception):
        sim_with_cutting(frags, wpm, 1000, _be(cuda), mode="bogus")


def test_stage_dict_views_match_oracle(cuda):
    """collect_fragment_data / direct_fragment_model / maximum_likelihood_model as Mappings have
    the reference's layouts and the oracle's numbers (counts = shots * probability)."""
    from qdc_b200.qsplit import qsplit_mlrecon_methods as qmm
    G = nx.Graph()
    G.add_edges_from([(0, 1), (1, 2), (2, 0), (2, 3), (3, 4), (4, 5), (5, 3), (1, 4)])
    _, _, frags, wpm = _cut_dqva(G, [[0, 1, 2], [3, 4, 5]], 2, 1, False)     # gate-faithful
    shots = 4096
    data = qmm.collect_fragment_data(frags, wpm, shots=shots, tomography_backend=_be(cuda, tile_bits=4))
    want = qo.collect_fragment_data(frags, wpm, drop_below=0.0)
    for got_f, want_f in zip(data, want):
        # outcomes whose exact probability is 0 up to rounding may or may not have a key
        for bits, row in want_f.items():
            if max(row.values()) < 1e-14:
                continue
            assert bits in got_f
            for lab, v in row.items():
                assert abs(got_f[bits][lab] / shots - v) < TOL
    direct = qmm.direct_fragment_model(data)
    want_direct = qo.direct_fragment_model(want)
    for got_f, want_f in zip(direct, want_direct):
        for bits, block in want_f.items():
            if np.abs(block).max() < 1e-14:
                continue
            assert np.abs(got_f[bits] / shots - block).max() < TOL
    # counts-dict entry point (one GEMM with the folded pseudo-inverse) == kernel path
    from_counts = qmm.direct_fragment_model({k: dict(v) for k, v in data[0].items()})
    for bits, block in direct[0].items():
        assert np.abs(from_counts[bits] - block).max() < 1e-8 * shots
    likely = qmm.maximum_likelihood_model(direct)
    want_likely = qo.maximum_likelihood_model(want_direct)
    for got_f, want_f in zip(likely, want_likely):
        for bits, block in want_f.items():
            if np.abs(block).max() < 1e-12:
                continue
            assert np.abs(got_f[bits] / shots - block).max() < 1e-9
    # models of exactly simulated fragments are Gram matrices: they carry `exact` and skip the Cholesky test
    # (the oracle comparison above ran that way); running the test anyway, or the full Jacobi path, agrees
    for model, fast in zip(direct, likely):
        assert model.exact and not qmm.ml_info_host(fast)[0]
        tested = qmm.maximum_likelihood_model(model, assume_psd=False)
        assert not qmm.ml_info_host(tested)[0] and torch.equal(tested.T, fast.T)
        full = qmm.maximum_likelihood_model(model, force_full=True)
        assert float((full.T - fast.T).abs().max()) < 1e-9 * shots


@pytest.mark.parametrize("fused", [False, True])
@pytest.mark.parametrize("max_cuts,seed", [(1, 0), (2, 1), (2, 5)])
def test_cut_dqva_matches_uncut_and_oracle(cuda, fused, max_cuts, seed):
    from qdc_b200.utils.cutting_funcs import sim_with_cutting
    G = nx.Graph()
    G.add_edges_from([(0, 1), (1, 2), (2, 0), (2, 3), (3, 4), (4, 5), (5, 3), (1, 4)])
    circ, cuts, frags, wpm = _cut_dqva(G, [[0, 1, 2], [3, 4, 5]], max_cuts, seed, fused)
    N = circ.num_qubits
    full = sv.probabilities(sv.simulate(N, circ.to_instructions()))
    be = _be(cuda, graph=G, tile_bits=3)
    dist = sim_with_cutting(frags, wpm, 1, be, mode="likely", topk=4)
    assert np.abs(dist.P.cpu().numpy() - full).max() < TOL
    want = qo.sim_with_cutting(frags, wpm, mode="likely")
    for k, v in want.items():
        assert abs(dist.get(k, 0.0) - v) < TOL
    n = G.number_of_nodes()
    s_p, s_pwok = mo.score_array(full, n, 2, mo.adjacency_masks(G, n))
    out = dist.score.get()
    assert abs(out["sum_p"] - 1) < TOL and abs(out["sum_pw_indset"] - s_pwok) < TOL
    cost = mo.avg_cost(mo.strip_ancillas(dict(dist), 2), G)            # the reference's own loop
    assert abs(-cost - out["sum_pw_indset"]) < TOL
    idx = np.nonzero(full > 1e-5)[0]
    order = idx[np.lexsort((idx, -full[idx]))][:4]
    assert [x for x, _ in out["top"]] == [int(i) for i in order]


def test_three_fragments(cuda):
    from qdc_b200.utils.cutting_funcs import sim_with_cutting
    G = nx.Graph()
    G.add_edges_from([(0, 1), (1, 2), (2, 3), (3, 4), (4, 5), (5, 6), (6, 7), (7, 8), (0, 2),
                      (3, 5), (6, 8)])
    circ, cuts, frags, wpm = _cut_dqva(G, [[0, 1, 2], [3, 4, 5], [6, 7, 8]], 2, 3, True)
    N = circ.num_qubits
    full = sv.probabilities(sv.simulate(N, circ.to_instructions()))
    dist = sim_with_cutting(frags, wpm, 1, _be(cuda, graph=G), mode="direct")
    assert np.abs(dist.P.cpu().numpy() - full).max() < TOL


def test_parameter_batch(cuda):
    """Unbound fragments + a table of parameter vectors == one bound evaluation per row."""
    import random
    from qdc_b200.ansatz import dqv_cut_ansatz
    from qdc_b200.circuit import Parameter
    from qdc_b200.qsplit import qsplit_circuit_cutter as qcc
    from qdc_b200.utils import graph_funcs
    from qdc_b200.utils.cutting_funcs import simple_choose_nodes, sim_with_cutting
    G = nx.Graph()
    G.add_edges_from([(0, 1), (1, 2), (2, 0), (2, 3), (3, 4), (4, 5), (5, 3), (1, 4), (5, 6), (6, 7), (7, 4)])
    partition = [[0, 1, 2], [3, 4, 5, 6, 7]]
    random.seed(2); np.random.seed(2)
    n = G.number_of_nodes()
    subgraphs, cut_edges = graph_funcs.get_subgraphs(G, partition)
    cut_nodes, hot_nodes = simple_choose_nodes(G, partition, cut_edges, 2)
    order = [v for part in partition for v in part]
    syms = [Parameter(f"var_{i}") for i in range(n + 1)]
    circ, cuts = dqv_cut_ansatz.gen_dqva(G, partition, cut_nodes, hot_nodes, order, params=syms,
                                         init_state="0" * n, fused=True)
    frags, wpm = qcc.cut_circuit(circ, cuts)
    used = circ.parameters
    table = np.random.uniform(0, 2 * np.pi, size=(5, len(used)))
    be = _be(cuda, graph=G, tile_bits=4)
    dists = sim_with_cutting(frags, wpm, 1, be, mode="direct", params=table, param_order=used)
    assert len(dists) == 5
    N = circ.num_qubits
    for row, dist in zip(table, dists):
        bound = circ.bind_parameters(dict(zip(used, row)))
        full = sv.probabilities(sv.simulate(N, bound.to_instructions()))
        assert np.abs(dist.P.cpu().numpy() - full).max() < TOL


def test_support_compaction_is_exact(cuda):
    """recombine(..., compact=True) drops the all-zero tensor rows (fragment outcomes that never
    occur): same distribution, same fused scores and top-k as the full contraction."""
    from qdc_b200 import recombine as rc
    from qdc_b200.qsplit import qsplit_mlrecon_methods as qmm
    G = nx.Graph()
    G.add_edges_from([(0, 1), (1, 2), (2, 0), (2, 3), (3, 4), (4, 5), (5, 3), (1, 4), (5, 6), (6, 7), (7, 4)])
    _, _, frags, wpm = _cut_dqva(G, [[0, 1, 2, 3], [4, 5, 6, 7]], 2, 3, True)
    be = _be(cuda)
    models = qmm.direct_fragment_model(qmm.collect_fragment_data(frags, wpm, 1, be))
    full = qmm.recombine_fragment_models(models, wpm, graph=G, topk=3, store=True, compact=False)
    comp = qmm.recombine_fragment_models(models, wpm, graph=G, topk=3, store=True, compact=True)
    assert torch.equal(full.P, comp.P)
    a, b = full.score.get(), comp.score.get()
    assert a["top"] == b["top"]
    for key in ("sum_p", "sum_pw", "sum_pw_indset"):
        assert abs(a[key] - b[key]) < 1e-13
    # the support really is a strict subset here (non-independent sets of a fragment have amplitude 0)
    T = models[0].T[0]
    assert int((T.abs().amax(dim=1) > 0).sum()) < T.shape[0]


def test_finite_shot_mode_converges_to_exact(cuda):
    """sample=True: counts drawn per variant circuit, least-squares model from the
    counts table, ML projection, recombination.  Statistical test: the error shrinks like
    1/sqrt(shots), the draws are reproducible, the dict view has the reference's layout."""
    from qdc_b200.qsplit import qsplit_mlrecon_methods as qmm
    from qdc_b200.utils.cutting_funcs import sim_with_cutting
    G = nx.Graph()
    G.add_edges_from([(0, 1), (1, 2), (2, 0), (2, 3), (3, 4), (4, 5), (5, 3), (1, 4)])
    circ, _, frags, wpm = _cut_dqva(G, [[0, 1, 2], [3, 4, 5]], 2, 5, True)
    be = _be(cuda)
    exact = sim_with_cutting(frags, wpm, 1, be, mode="direct", store=True).P.cpu().numpy()
    errs = []
    for shots in (2000, 200000):
        a = sim_with_cutting(frags, wpm, shots, be, mode="likely", store=True, sample=True, seed=7)
        b = sim_with_cutting(frags, wpm, shots, be, mode="likely", store=True, sample=True, seed=7)
        assert torch.equal(a.P, b.P)
        got = a.P.cpu().numpy()
        assert abs(got.sum() - 1) < 1e-9
        errs.append(np.abs(got - exact).max())
    assert errs[1] < 0.01 and errs[1] < errs[0]
    data = qmm.collect_fragment_data(frags, wpm, shots=500, tomography_backend=be, sample=True, seed=1)
    for d in data:
        p, m = len(d.prep_qubits), len(d.meas_qubits)
        assert abs(float(d.counts.sum()) - 500 * 4 ** p * 3 ** m) < 1e-6      # every draw lands in one cell
        row = next(iter(d.values()))
        assert len(row) == 4 ** p * 6 ** m


def _scenario_models(noise, seed):
    """Fragment Pauli tensors of small cut circuits from the ORACLE (optionally with noise, so that the
    Choi blocks are not positive semidefinite), as (tensor [n_b, 4^q], n_prep, n_meas)."""
    import cutqc_cases as cc
    rng = np.random.default_rng(seed)
    out = []
    for name in ("ghz3", "ckt5", "there_and_back", "chain3", "dqva6"):
        circ, frags, wpm = cc.CASES[name]()
        ft = qo.identify_frag_targets(wpm)
        for i, frag in enumerate(frags):
            R = qo.fragment_pauli_tensor(frag, ft[i]["prep"], ft[i]["meas"])
            T = R.reshape(R.shape[0], -1)
            T = T + noise * rng.standard_normal(T.shape) * (np.abs(T).max(axis=1, keepdims=True) > 0)
            out.append((name, i, T, len(ft[i]["prep"]), len(ft[i]["meas"])))
    return out


@pytest.mark.parametrize("noise,force_full", [(0.0, False), (0.0, True), (0.05, False), (1e-3, False)])
def test_maximum_likelihood_kernel_matches_oracle(cuda, noise, force_full):
    """qdc_ml_project against the oracle's line-faithful maximum_likelihood_model
    (qsplit_mlrecon_methods.py:320-356: scipy eigh per block, sequential clipping walk, rebuild through
    to_projector) on exact tensors (early-out: every block is PSD), on the same with the early-out
    switched off (the Jacobi path must reproduce the conjugate), and on noisy, non-PSD tensors; block
    sizes 2 x 2 to 16 x 16 occur.  Keys (which outcomes survive) and blocks to 1e-10."""
    from qdc_b200.qsplit import qsplit_mlrecon_methods as qmm
    seen_dims, seen_full = set(), set()
    for name, i, T, p, m in _scenario_models(noise, 3):
        model = qmm.FragmentModel(torch.as_tensor(T, device=cuda).unsqueeze(0), p, m, {"prep": [], "meas": []})
        direct = {k: v for k, v in model.items()}
        want = qo.maximum_likelihood_model(direct)
        got_model = qmm.maximum_likelihood_model(model, force_full=force_full)
        need_full, steps, t0 = qmm.ml_info_host(got_model)
        got = dict(got_model.items())
        seen_dims.add(2 ** (p + m)); seen_full.add(need_full)
        scale = max(np.abs(b).max() for b in direct.values())
        for bits, block in want.items():
            if np.abs(block).max() < 1e-12 * scale:
                continue
            assert bits in got, (name, i, bits)
            assert np.abs(got[bits] - block).max() < 1e-10 * max(scale, 1.0), (name, i, bits)
        for bits, block in got.items():
            if bits not in want:
                assert np.abs(block).max() < 1e-10 * max(scale, 1.0), (name, i, bits)
        if noise == 0.0 and not force_full:
            assert not need_full and t0 == 0.0
        if noise >= 0.05:
            assert need_full and t0 > 0.0 and 1 <= steps <= 12
    assert {2, 4} <= seen_dims and max(seen_dims) >= 4
    assert (True in seen_full) == (force_full or noise > 0)


@pytest.mark.parametrize("p,m", [(1, 0), (1, 1), (2, 1), (2, 2), (4, 0), (0, 4)])
def test_maximum_likelihood_kernel_random_blocks(cuda, p, m):
    """Random Hermitian blocks of every supported size (2 .. 16), many of them indefinite, some of
    them exactly rank one or zero: Jacobi eigen-decomposition + global soft threshold + conjugated
    rebuild against numpy (eigh per block, the oracle's clipping walk over all eigenvalues)."""
    from qdc_b200.qsplit import qsplit_mlrecon_methods as qmm
    rng = np.random.default_rng(100 * p + m)
    D, n_b = 2 ** (p + m), 300
    L = rng.standard_normal((n_b, D, D)) + 1j * rng.standard_normal((n_b, D, D))
    L = L @ np.conj(np.swapaxes(L, 1, 2)) / D + 0.3 * (L + np.conj(np.swapaxes(L, 1, 2)))
    v = rng.standard_normal((n_b, D)) + 1j * rng.standard_normal((n_b, D))
    L[::7] = v[::7, :, None] * np.conj(v[::7, None, :])            # rank one
    L[5] = 0.0                                                      # an outcome that never occurs
    T = qmm._tensor_from_choi(torch.as_tensor(L), p, m)
    model = qmm.FragmentModel(T.to(cuda).unsqueeze(0), p, m, {"prep": [], "meas": []})
    got_model = qmm.maximum_likelihood_model(model)
    need_full, steps, t0 = qmm.ml_info_host(got_model)
    assert need_full and t0 > 0
    got = got_model.choi_blocks().cpu().numpy()
    want_dict = qo.maximum_likelihood_model({format(b, "012b"): L[b] for b in range(n_b) if np.abs(L[b]).max() > 0})
    want = np.zeros_like(L)
    for bits, block in want_dict.items():
        want[int(bits, 2)] = block
    assert np.abs(got - want).max() < 1e-10 * np.abs(L).max()
    assert np.abs(got[5]).max() == 0.0
    # positive semidefinite and trace preserving (the point of the projection)
    assert np.linalg.eigvalsh(got).min() > -1e-10 * np.abs(L).max()
    assert abs(np.trace(got, axis1=1, axis2=2).sum().real - np.trace(L, axis1=1, axis2=2).sum().real) < 1e-9 * n_b
