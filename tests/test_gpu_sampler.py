"""The device shot sampler (qdc_sv_sample_counts, csrc/shot_sampler.cu) through the C-ABI and
through collect_fragment_data(sample=True): what the reference gets from qasm_simulator jobs +
organize_tomography_data (qsplit_mlrecon_methods.py:20-42, 169-232).  A sampler has no bit-exact
oracle; the checks are the ones the domain offers: every draw lands in exactly one cell, states
with one outcome are sampled exactly, counts agree with shots * |amp|^2 within binomial bounds
cell by cell (6 sigma + chi-square), same arguments -> same table, other seed / stream -> other
table, and the least-squares model built from the counts converges on the exact model of every
parameter set."""
import ctypes as C

import networkx as nx
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _sample(cuda, states, n, n_init, n_sets, flip, shots, seed, stream_id, meas_mask, cols, ld=None):
    from qdc_b200 import _lib
    lib = _lib.load()
    m = bin(meas_mask).count("1")
    ld = ld or int(cols.max()) + 1
    counts = torch.zeros(n_sets, 1 << (n - m), ld, dtype=torch.float64, device=cuda)
    ws = torch.empty(int(lib.qdc_sv_sample_workspace_doubles(n, n_init * n_sets)), dtype=torch.float64, device=cuda)
    d_cols = torch.as_tensor(cols.astype(np.int32), device=cuda)
    with torch.cuda.device(cuda):
        _lib.check(lib.qdc_sv_sample_counts(_lib.ptr(states), n, n_init, n_sets, C.c_uint64(flip), shots,
                                            C.c_uint64(seed), C.c_uint64(stream_id), C.c_uint64(meas_mask),
                                            _lib.ptr(d_cols), _lib.ptr(counts), ld, _lib.ptr(ws), None),
                   "qdc_sv_sample_counts")
    torch.cuda.synchronize()
    return counts


def _pext(x, mask):
    out = j = 0
    q = 0
    while mask >> q:
        if (mask >> q) & 1:
            out |= ((x >> q) & 1) << j
            j += 1
        q += 1
    return out


@pytest.mark.parametrize("n,meas_mask,flip", [(3, 0b010, 0b101), (10, 0b1000010010, 0b0110011010), (13, 0, 0),
                                              (17, 0b11 << 15, 1 << 16)])
def test_sampler_counts_match_probabilities(cuda, n, meas_mask, flip):
    rng = np.random.default_rng(n)
    n_init, n_sets, shots = 4, 3, 40000
    m = bin(meas_mask).count("1")
    psi = rng.normal(size=(n_init * n_sets, 1 << n)) + 1j * rng.normal(size=(n_init * n_sets, 1 << n))
    psi *= rng.random((n_init * n_sets, 1 << n)) < 0.3                  # many exact zeros
    psi[:, : (1 << n) // 2] *= 3.0                                      # uneven mass over the tiles
    psi[1] *= 0.25                                                      # an unnormalised state is sampled by its own mass
    psi[2] = 0.0                                                        # a state without norm draws nothing
    states = torch.as_tensor(psi, device=cuda)
    cols = np.arange(n_init * (1 << m)).reshape(n_init, 1 << m)         # one column per (init, measured value)
    counts = _sample(cuda, states, n, n_init, n_sets, flip, shots, 1234, 5, meas_mask, cols).cpu().numpy()
    assert np.array_equal(counts, np.round(counts)) and counts.min() >= 0
    keep = ((1 << n) - 1) & ~meas_mask
    xs = np.arange(1 << n)
    b = np.array([_pext(int(x), keep) for x in xs]) if n <= 13 else None
    if b is None:                                                       # measured qubits on top: b = low bits
        b, v = xs & keep, xs >> 15
    else:
        v = np.array([_pext(int(x), meas_mask) for x in xs])
    chi2 = dof = 0
    for init in range(n_init):
        for s in range(n_sets):
            stored = np.abs(psi[init * n_sets + s]) ** 2
            p = np.empty_like(stored)
            p[xs] = stored[xs ^ flip]                                   # logical x lives at stored index x ^ flip
            hist = counts[s][b, init * (1 << m) + v]                    # histogram of this state over x
            if p.sum() == 0.0:
                assert hist.sum() == 0
                continue
            p /= p.sum()
            assert hist.sum() == shots
            assert np.all(hist[p == 0] == 0)
            sigma = np.sqrt(shots * p * (1 - p))
            # 6 sigma, plus room for the Poisson tail of the ~10^6 cells whose expectation is << 1
            assert np.all(np.abs(hist - shots * p) <= 6 * sigma + 6)
            big = shots * p >= 5
            chi2 += (((hist - shots * p) ** 2)[big] / (shots * p)[big]).sum()
            dof += int(big.sum())
    assert dof == 0 or abs(chi2 - dof) < 6 * np.sqrt(2 * dof) + 10
    again = _sample(cuda, states, n, n_init, n_sets, flip, shots, 1234, 5, meas_mask, cols).cpu().numpy()
    assert np.array_equal(again, counts)                                # a function of its arguments
    other_seed = _sample(cuda, states, n, n_init, n_sets, flip, shots, 1235, 5, meas_mask, cols).cpu().numpy()
    other_stream = _sample(cuda, states, n, n_init, n_sets, flip, shots, 1234, 6, meas_mask, cols).cpu().numpy()
    assert not np.array_equal(other_seed, counts) and not np.array_equal(other_stream, counts)


@pytest.mark.parametrize("n", [1, 5, 8, 16, 21])
def test_sampler_single_outcome_states_are_exact(cuda, n):
    """A basis state has one outcome: all shots land there, wherever it sits in the three-level walk
    (first / last amplitude, tile and super-tile boundaries), with and without the flip frame;
    shots = 0 draws nothing."""
    N = 1 << n
    spots = sorted({0, N - 1, N // 2, min(N - 1, 255), min(N - 1, 256), min(N - 1, 65535), min(N - 1, 65536),
                    (N - 1) ^ (N >> 1)})
    flip = (N - 1) & 0b1011011
    psi = np.zeros((len(spots), N), dtype=np.complex128)
    for i, x in enumerate(spots):
        psi[i, x] = 0.6 - 0.8j
    states = torch.as_tensor(psi, device=cuda)
    cols = np.zeros((len(spots), 1), dtype=np.int32)
    shots = 777
    for f in (0, flip):
        counts = _sample(cuda, states, n, len(spots), 1, f, shots, 9, 0, 0, cols + np.arange(len(spots))[:, None])
        counts = counts.cpu().numpy()[0]
        for i, x in enumerate(spots):
            want = np.zeros(N)
            want[x ^ f] = shots
            assert np.array_equal(counts[:, i], want)
    none = _sample(cuda, states, n, len(spots), 1, 0, 0, 9, 0, 0, cols)
    assert float(none.sum()) == 0.0


def test_sampler_argument_checks(cuda):
    from qdc_b200 import _lib
    lib = _lib.load()
    st = torch.zeros(2, 8, dtype=torch.complex128, device=cuda)
    buf = torch.zeros(64, dtype=torch.float64, device=cuda)
    col = torch.zeros(8, dtype=torch.int32, device=cuda)
    bad = [dict(n=3, flip=8), dict(n=3, meas=8), dict(n=0), dict(n=3, shots=-1), dict(n=3, ld=0)]
    for kw in bad:
        rc = lib.qdc_sv_sample_counts(_lib.ptr(st), kw.get("n", 3), 2, 1, C.c_uint64(kw.get("flip", 0)),
                                      kw.get("shots", 10), C.c_uint64(1), C.c_uint64(0), C.c_uint64(kw.get("meas", 0)),
                                      _lib.ptr(col), _lib.ptr(buf), kw.get("ld", 1), _lib.ptr(buf), None)
        assert rc != 0
    assert lib.qdc_sv_sample_workspace_doubles(0, 1) < 0
    assert lib.qdc_sv_sample_workspace_doubles(20, 3) == 3 * ((1 << 12) + 16)


def test_sampled_models_of_a_parameter_batch_converge(cuda):
    """collect_fragment_data(sample=True) over a BATCH of parameter vectors: one counts table per set
    (every variant circuit x every set sampled by one kernel call per basis choice), and the
    least-squares model of each set converges on that set's exact model like 1/sqrt(shots)."""
    from qdc_b200.engine import B200Backend
    from qdc_b200.qsplit import qsplit_mlrecon_methods as qmm
    import random
    from qdc_b200.ansatz import dqv_cut_ansatz
    from qdc_b200.circuit import Parameter
    from qdc_b200.qsplit import qsplit_circuit_cutter as qcc
    from qdc_b200.utils import graph_funcs
    from qdc_b200.utils.cutting_funcs import simple_choose_nodes
    G = nx.Graph()
    G.add_edges_from([(0, 1), (1, 2), (2, 0), (2, 3), (3, 4), (4, 5), (5, 3), (1, 4)])
    partition = [[0, 1, 2], [3, 4, 5]]
    random.seed(5); np.random.seed(5)
    _, cut_edges = graph_funcs.get_subgraphs(G, partition)
    cut_nodes, hot_nodes = simple_choose_nodes(G, partition, cut_edges, 2)
    syms = [Parameter(f"var_{i}") for i in range(7)]
    circ, cuts = dqv_cut_ansatz.gen_dqva(G, partition, cut_nodes, hot_nodes, [0, 1, 2, 3, 4, 5], params=syms,
                                         init_state="0" * 6, fused=True)
    frags, wpm = qcc.cut_circuit(circ, cuts)
    be = B200Backend(device=cuda)
    order = circ.parameters
    table = np.random.default_rng(3).uniform(0, 2 * np.pi, (3, len(order)))
    exact = qmm.direct_fragment_model(qmm.collect_fragment_data(frags, wpm, tomography_backend=be, params=table,
                                                                param_order=order))
    errs = []
    for shots in (1000, 100000):
        data = qmm.collect_fragment_data(frags, wpm, shots=shots, tomography_backend=be, params=table,
                                         param_order=order, sample=True, seed=11)
        models = qmm.direct_fragment_model(data)
        worst = 0.0
        for d, got, want in zip(data, models, exact):
            p, m = len(d.prep_qubits), len(d.meas_qubits)
            assert d.counts.shape[0] == 3
            per_set = d.counts.sum(dim=(1, 2)).cpu().numpy()
            assert np.array_equal(per_set, np.full(3, shots * 4 ** p * 3 ** m))
            assert got.T.shape == want.T.shape
            worst = max(worst, float((got.T - want.T).abs().max()))
        errs.append(worst)
    assert errs[1] < 0.05 and errs[1] < errs[0] / 3
    # the three sets really are different experiments
    assert float((models[0].T[0] - models[0].T[1]).abs().max()) > 1e-3
