"""GPU parity of the statevector engine (csrc/statevec.cu) against the numpy oracle
(oracle/statevector.py, oracle/qsplit_oracle.py) on seeded circuits.  Tolerance 1e-10 max-abs on
amplitudes / probabilities / tensor entries (north_star FP64 bar); the index bookkeeping (flip
frame, tile <-> full index maps, batch order) is exact by construction of the comparison."""
import itertools

import numpy as np
import pytest
import torch

from oracle import qsplit_oracle as qo
from oracle import statevector as sv
from test_sv_compiler import _faithful_mixer, _random_circuit

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _backend(cuda, **kw):
    from qdc_b200.engine import B200Backend
    return B200Backend(device=cuda, **kw)


def _unflip(states, flip):
    idx = np.arange(states.shape[-1]) ^ flip
    return states[..., idx]


@pytest.mark.parametrize("n,tile_bits,coalesce,seed", [
    (3, 0, -1, 0), (6, 3, 1, 1), (8, 4, 2, 2), (10, 0, -1, 3), (12, 12, 4, 4), (13, 13, 4, 5),
    (14, 12, 4, 6), (14, 7, 3, 7), (16, 12, 5, 8), (11, 5, 0, 9), (1, 0, -1, 10),
])
def test_random_circuits_match_oracle(cuda, n, tile_bits, coalesce, seed):
    rng = np.random.default_rng(seed)
    insts = _random_circuit(rng, n, 80)
    be = _backend(cuda, tile_bits=tile_bits, coalesce_bits=coalesce)
    prog, auto = be.compile_instructions(n, insts)
    states = be.run(prog, auto)
    want = sv.simulate(n, insts)
    got = _unflip(states.cpu().numpy(), prog.flip_mask)[0]
    assert np.abs(got - want).max() < TOL
    probs = be.probabilities(states, prog).cpu().numpy()[0]
    assert np.abs(probs - sv.probabilities(want)).max() < TOL


def test_parameter_sets_and_prep_variants(cuda):
    """batch = prep variant x parameter set; symbolic angles come from the parameter table."""
    from qdc_b200.circuit import Parameter
    from qdc_b200.engine import SIC_KETS
    n, prep = 9, [2, 7]
    rng = np.random.default_rng(3)
    pars = [Parameter(f"var_{i}") for i in range(4)]
    insts = [("x", (0,), (), None)]
    insts.append(("zcrx", (0, 1, 3), (2 * pars[0],), None))
    insts.append(("rx", (2,), (pars[1],), None))
    insts.append(("h", (5,), (), None))
    insts.append(("crx", (2, 7), (2 * pars[2],), None))
    insts.append(("zcrx", (7, 8, 6), (2 * pars[3],), None))
    insts.append(("rz", (7,), (pars[1] / 2,), None))
    insts.append(("rz", (2,), (-1 * pars[1] / 2,), None))
    insts.append(("cx", (5, 4), (), None))
    insts.append(("ry", (4,), (0.3,), None))
    be = _backend(cuda, tile_bits=5, coalesce_bits=2)
    prog, auto = be.compile_instructions(n, insts, param_order=pars, zero_mask=((1 << n) - 1) & ~0b10000100)
    table = rng.uniform(0, 2 * np.pi, size=(3, 4))
    states = be.run(prog, auto, params=table, prep_qubits=prep)
    assert states.shape == (16 * 3, 1 << n)
    got = _unflip(states.cpu().numpy(), prog.flip_mask)
    for v, (c0, c1) in enumerate(itertools.product(range(4), repeat=2)):
        init = np.zeros(1 << n, dtype=complex)
        for b0 in range(2):
            for b1 in range(2):
                init[(b0 << prep[0]) | (b1 << prep[1])] = SIC_KETS[c0][b0] * SIC_KETS[c1][b1]
        for s in range(3):
            bound = [(nm, qs, tuple(p.bind(table[s][pars.index(p.param)]) if hasattr(p, "bind") else p
                                    for p in ps), mt) for nm, qs, ps, mt in insts]
            want = sv.simulate(n, bound, init=init)
            assert np.abs(got[v * 3 + s] - want).max() < TOL, (v, s)


def test_gate_faithful_mixers_equal_fused(cuda):
    rng = np.random.default_rng(5)
    n, anc = 12, 11
    faithful, fused = [("x", (3,), (), None)], [("x", (3,), (), None)]
    for q in range(11):
        nb = [int(v) for v in rng.choice([k for k in range(11) if k != q], size=3, replace=False)]
        a = float(rng.uniform(0, 2 * np.pi))
        _faithful_mixer(faithful, nb, anc, q, a)
        fused.append(("zcrx", tuple(nb) + (q,), (2 * a,), None))
    for be in (_backend(cuda, tile_bits=6, fuse_ancilla=True), _backend(cuda, tile_bits=6, fuse_ancilla=False)):
        prog, auto = be.compile_instructions(n, faithful)
        got = _unflip(be.run(prog, auto).cpu().numpy(), prog.flip_mask)[0]
        assert np.abs(got - sv.simulate(n, fused)).max() < TOL


@pytest.mark.parametrize("n,prep,meas,tile_bits,seed", [
    (6, [1], [4], 0, 0), (8, [], [2, 6], 4, 1), (8, [0, 5], [], 4, 2), (9, [3, 8], [1, 6], 5, 3),
    (10, [2], [0, 9, 4], 12, 4), (7, [], [], 3, 5), (13, [], [12, 3, 7, 9], 13, 6),
    (11, [10, 0, 4, 6], [], 6, 7),
])
def test_pauli_tensor_matches_oracle(cuda, n, prep, meas, tile_bits, seed):
    """K3+K4 fused: tensor straight from amplitudes == oracle's 3^m measured variants combined
    with signs and the SIC frame inverse (oracle.fragment_pauli_tensor)."""
    rng = np.random.default_rng(seed)
    insts = _random_circuit(rng, n, 50)
    be = _backend(cuda, tile_bits=tile_bits)
    zero_mask = ((1 << n) - 1) & ~sum(1 << q for q in prep)
    prog, auto = be.compile_instructions(n, insts, zero_mask=zero_mask)
    states = be.run(prog, auto, prep_qubits=prep)
    T = be.pauli_tensor(states, prog, 1, len(prep), meas).cpu().numpy()[0]
    R = qo.fragment_pauli_tensor((n, insts), prep, meas)
    assert T.shape == (1 << (n - len(meas)), 4 ** (len(prep) + len(meas)))
    assert np.abs(T - R.reshape(T.shape)).max() < TOL
    if prep and len(prep) + len(meas) <= 4:
        # linearity: the 2^p basis-input evolutions give the same tensor as the 4^p SIC variants
        from qdc_b200.engine import BASIS_KETS
        basis = be.run(prog, auto, prep_qubits=prep, kets=BASIS_KETS)
        assert basis.shape[0] == 2 ** len(prep)
        Tb = be.pauli_tensor(basis, prog, 1, len(prep), meas, basis_inputs=True).cpu().numpy()[0]
        assert np.abs(Tb - R.reshape(T.shape)).max() < TOL


def test_pauli_tensor_strides_and_sets(cuda):
    rng = np.random.default_rng(9)
    n, prep, meas = 7, [5], [2, 0]
    insts = [(nm, qs, ps, mt) for nm, qs, ps, mt in _random_circuit(rng, n, 40)]
    be = _backend(cuda, tile_bits=4)
    prog, auto = be.compile_instructions(n, insts, zero_mask=((1 << n) - 1) & ~(1 << 5))
    states = be.run(prog, auto, params=np.zeros((2, 0)), prep_qubits=prep)
    strides = [1, 16, 4]                                   # axis order (prep0, meas0, meas1) -> custom layout
    T = be.pauli_tensor(states, prog, 2, 1, meas, axis_strides=strides, ld=64).cpu().numpy()
    R = qo.fragment_pauli_tensor((n, insts), prep, meas)   # [b, a_prep, a_m0, a_m1]
    want = np.zeros((1 << (n - 2), 64))
    for a0, a1, a2 in itertools.product(range(4), repeat=3):
        want[:, a0 * 1 + a1 * 16 + a2 * 4] = R[:, a0, a1, a2]
    assert np.abs(T[0] - want).max() < TOL and np.abs(T[1] - want).max() < TOL
    from qdc_b200.engine import BASIS_KETS
    basis = be.run(prog, auto, params=np.zeros((2, 0)), prep_qubits=prep, kets=BASIS_KETS)
    Tb = be.pauli_tensor(basis, prog, 2, 1, meas, axis_strides=strides, ld=64, basis_inputs=True).cpu().numpy()
    assert np.abs(Tb[0] - want).max() < TOL and np.abs(Tb[1] - want).max() < TOL


def test_amplitude_scoring_matches_oracle(cuda):
    """Uncut cost f(params) (mis.py:639-665): sum p * popcount straight from amplitudes, with the
    flip frame folded into the scoring tables."""
    import networkx as nx
    from oracle import mis_oracle as mo
    from qdc_b200 import recombine as rc
    n = 12
    rng = np.random.default_rng(2)
    insts = _random_circuit(rng, n, 70)
    be = _backend(cuda, tile_bits=7)
    prog, auto = be.compile_instructions(n, insts)
    states = be.run(prog, auto)
    G = nx.random_regular_graph(3, n - 1, seed=1) if (n - 1) % 2 == 0 else nx.path_graph(n - 1)
    adj = [int(a) for a in mo.adjacency_masks(G, n - 1)] + [0]      # top wire = ancilla
    tabs = rc.build_tables(6, 6, list(range(n)), adj, (1 << (n - 1)) - 1, device=cuda,
                           index_flip=prog.flip_mask)
    res = rc.reduce_and_score(torch.view_as_real(states[0]).reshape(-1), tabs, k=3, amplitudes=True).get()
    p = sv.probabilities(sv.simulate(n, insts))
    s_p, s_pwok = mo.score_array(p, n - 1, 1, np.array(adj[:n - 1], dtype=np.uint64), masked=True)
    _, s_pw = mo.score_array(p, n - 1, 1, np.array(adj[:n - 1], dtype=np.uint64), masked=False)
    assert abs(res["sum_p"] - s_p) < TOL and abs(res["sum_pw_indset"] - s_pwok) < TOL
    assert abs(res["sum_pw"] - s_pw) < TOL
    order = np.lexsort((np.arange(p.size), -p))[:3]
    assert [x for x, _ in res["top"]] == [int(i) for i in order if p[i] > 1e-5]


def test_uncut_cost_batch_matches_oracle(cuda):
    """mis.uncut_cost == the reference's f(params) (mis.py:639-665) per parameter vector."""
    import networkx as nx
    from oracle import mis_oracle as mo
    from qdc_b200 import mis
    from qdc_b200.ansatz import dqv_ansatz
    from qdc_b200.circuit import Parameter
    n = 10
    G = nx.random_regular_graph(3, n, seed=5)
    syms = [Parameter(f"var_{i}") for i in range(n + 1)]
    init = "0001000000"
    circ = dqv_ansatz.gen_dqva(G, P=1, params=syms, init_state=init, mixer_order=list(range(n)))
    used = circ.parameters
    table = np.random.default_rng(2).uniform(0, 2 * np.pi, (5, len(used)))
    be = _backend(cuda, tile_bits=6)
    got = mis.uncut_cost(circ, be, G, params=table, param_order=used, max_states=2)
    got_masked = mis.uncut_cost(circ, be, G, params=table, param_order=used, masked=True)
    adj = mo.adjacency_masks(G, n)
    for s in range(5):
        bound = circ.bind_parameters(dict(zip(used, table[s])))
        p = sv.probabilities(sv.simulate(n + 1, bound.to_instructions()))
        _, s_pw = mo.score_array(p, n, 1, adj, masked=False)
        _, s_pwok = mo.score_array(p, n, 1, adj, masked=True)
        assert abs(got[s] + s_pw) < TOL and abs(got_masked[s] + s_pwok) < TOL


@pytest.mark.parametrize("n,tile_bits", [(10, 6), (10, 12), (13, 5), (7, 3)])
def test_run_scored_fuses_the_mis_reduction_into_the_last_group(cuda, n, tile_bits):
    """engine.run_scored (qdc_sv_run_score): the three sums reduced inside the last gate group equal the
    separate pass over the stored state (qdc_mis_reduce on the amplitudes) and the oracle, for several tile
    sizes (one group / many groups, local bits scattered over the index); with keep_state the state written
    is bit-identical to backend.run's, and the sums do not depend on whether it is written."""
    import networkx as nx
    import torch
    from oracle import mis_oracle as mo
    from qdc_b200 import recombine as rc
    from qdc_b200.ansatz import dqv_ansatz
    from qdc_b200.circuit import Parameter
    G = nx.gnm_random_graph(n, 2 * n, seed=n)
    syms = [Parameter(f"var_{i}") for i in range(n + 1)]
    init = "".join("1" if i == 2 else "0" for i in range(n))
    circ = dqv_ansatz.gen_dqva(G, P=1, params=syms, init_state=init, mixer_order=list(range(n)))
    used = circ.parameters
    table = np.random.default_rng(n).uniform(0, 2 * np.pi, (3, len(used)))
    be = _backend(cuda, tile_bits=tile_bits)
    prog, auto = be.compile(circ, param_order=used)
    nq = prog.n_qubits
    adj = [int(a) for a in mo.adjacency_masks(G, n)] + [0] * (nq - n)
    node_mask = (1 << n) - 1
    states = be.run(prog, auto, table)
    scratch = torch.zeros_like(states)
    sums_kept = be.run_scored(prog, auto, table, adj, node_mask, scratch=scratch, keep_state=True)
    assert torch.equal(scratch, states)
    sums = be.run_scored(prog, auto, table, adj, node_mask).cpu().numpy()
    assert np.array_equal(sums, sums_kept.cpu().numpy())
    cb = min(nq, 14)
    tabs = rc.build_tables(nq - cb, cb, list(range(nq)), adj, node_mask, device=cuda, index_flip=prog.flip_mask)
    mo_adj = mo.adjacency_masks(G, n)
    for s in range(3):
        ref = rc.reduce_and_score(torch.view_as_real(states[s]).reshape(-1), tabs, threshold=2.0, k=1, cand_cap=1,
                                  amplitudes=True).sums.cpu().numpy()
        assert np.abs(sums[s] - ref).max() < 1e-12
        bound = circ.bind_parameters(dict(zip(used, table[s])))
        p = sv.probabilities(sv.simulate(nq, bound.to_instructions()))
        _, s_pw = mo.score_array(p, n, nq - n, mo_adj, masked=False)
        _, s_pwok = mo.score_array(p, n, nq - n, mo_adj, masked=True)
        assert abs(sums[s][0] - 1.0) < TOL and abs(sums[s][2] - s_pw) < TOL and abs(sums[s][1] - s_pwok) < TOL


def test_uncut_cost_on_a_tiny_graph(cuda):
    """Graphs of 3-4 nodes give fewer than 64 table columns, where qdc_mis_reduce takes its scalar
    path: complex128 amplitudes must be scored as |amp|^2 there as well (they were once read as
    float32)."""
    import networkx as nx
    from oracle import mis_oracle as mo
    from qdc_b200 import mis
    from qdc_b200.ansatz import dqv_ansatz
    from qdc_b200.circuit import Parameter
    for n, edges in ((3, [(0, 1), (1, 2)]), (4, [(0, 1), (1, 2), (2, 3), (3, 0)])):
        G = nx.Graph()
        G.add_nodes_from(range(n))
        G.add_edges_from(edges)
        syms = [Parameter(f"var_{i}") for i in range(n + 1)]
        circ = dqv_ansatz.gen_dqva(G, P=1, params=syms, init_state="0" * n, mixer_order=list(range(n)))
        used = circ.parameters
        table = np.random.default_rng(n).uniform(0, 2 * np.pi, (3, len(used)))
        be = _backend(cuda)
        got = mis.uncut_cost(circ, be, G, params=table, param_order=used)
        got_masked = mis.uncut_cost(circ, be, G, params=table, param_order=used, masked=True)
        adj = mo.adjacency_masks(G, n)
        for s in range(3):
            bound = circ.bind_parameters(dict(zip(used, table[s])))
            p = sv.probabilities(sv.simulate(n + 1, bound.to_instructions()))
            _, s_pw = mo.score_array(p, n, 1, adj, masked=False)
            _, s_pwok = mo.score_array(p, n, 1, adj, masked=True)
            assert abs(got[s] + s_pw) < TOL and abs(got_masked[s] + s_pwok) < TOL


def test_qaoa_plus_circuit_on_engine(cuda):
    import networkx as nx
    from qdc_b200.ansatz import qaoa_plus
    G = nx.random_regular_graph(3, 12, seed=1)
    circ = qaoa_plus.construct_qaoa_plus(2, G, [0.4, 0.9, 1.3, 0.25])
    be = _backend(cuda, tile_bits=7)
    prog, auto = be.compile(circ)
    got = be.probabilities(be.run(prog, auto), prog).cpu().numpy()[0]
    want = sv.probabilities(sv.simulate(12, circ.to_instructions()))
    assert np.abs(got - want).max() < TOL


@pytest.mark.parametrize("n_bits,dtype", [(5, "f64"), (13, "f64"), (16, "f32"), (15, "amp")])
def test_penalty_reduce_matches_direct_sums(cuda, n_bits, dtype):
    """qdc_penalty_reduce: [sum p, sum p #edges, sum p |x|] with a flip frame and a non-node wire."""
    import networkx as nx
    import torch
    from qdc_b200 import recombine as rc
    rng = np.random.default_rng(n_bits)
    n_nodes = n_bits - 1                                    # top wire is an ancilla
    G = nx.gnm_random_graph(n_nodes, 2 * n_nodes, seed=3)
    adj = [0] * n_bits
    for a, b in G.edges():
        adj[a] |= 1 << b
        adj[b] |= 1 << a
    flip = int(rng.integers(0, 1 << n_bits))
    N = 1 << n_bits
    if dtype == "amp":
        amps = rng.normal(size=N) + 1j * rng.normal(size=N)
        p = np.abs(amps) ** 2
        dev = torch.view_as_real(torch.as_tensor(amps, device=cuda)).reshape(-1)
    else:
        p = rng.random(N).astype(np.float32 if dtype == "f32" else np.float64)
        dev = torch.as_tensor(p, device=cuda)
        p = p.astype(np.float64)
    got = rc.penalty_reduce(dev, n_bits, adj, (1 << n_nodes) - 1, index_flip=flip,
                            amplitudes=dtype == "amp").cpu().numpy()
    x = (np.arange(N) ^ flip) & ((1 << n_nodes) - 1)
    w = np.array([bin(v).count("1") for v in x])
    e = np.zeros(N)
    for a, b in G.edges():
        e += ((x >> a) & (x >> b) & 1)
    want = np.array([p.sum(), (p * e).sum(), (p * w).sum()])
    assert np.abs(got - want).max() < 1e-9 * max(1.0, want.max())


def test_qaoa_plus_cost_batch_matches_oracle(cuda):
    """mis.qaoa_plus_cost == -expectation_value(counts, G, Lambda) (ansatz/qaoa_plus.py:40-53)."""
    import networkx as nx
    from oracle import mis_oracle as mo
    from qdc_b200 import mis
    from qdc_b200.ansatz import qaoa_plus
    from qdc_b200.circuit import Parameter
    n, P, Lambda = 10, 2, 1.5
    G = nx.random_regular_graph(3, n, seed=7)
    syms = [Parameter(f"var_{i}") for i in range(2 * P)]
    circ = qaoa_plus.construct_qaoa_plus(P, G, syms)
    table = np.random.default_rng(4).uniform(0, 2 * np.pi, (4, 2 * P))
    got = mis.qaoa_plus_cost(circ, _backend(cuda, tile_bits=6), G, Lambda, params=table,
                             param_order=syms, max_states=3)
    for s in range(4):
        bound = circ.bind_parameters(dict(zip(syms, table[s])))
        p = sv.probabilities(sv.simulate(n, bound.to_instructions()))
        counts = {format(x, f"0{n}b"): p[x] for x in range(1 << n)}
        assert abs(got[s] + mo.qaoa_plus_expectation(counts, G, Lambda)) < TOL


@pytest.mark.parametrize("n,tile_bits,coalesce", [(9, 3, 1), (10, 5, 2), (11, 7, 3), (12, 8, 2), (13, 12, 5)])
def test_controlled_diagonal_runs_match_oracle(cuda, n, tile_bits, coalesce):
    """Diagonal gates with one control join the table pass of a diagonal run whenever control and target are not
    both tile-local (csrc/statevec.cu: the control bit then selects diag(1, d) / diag(d, 1) in the low or high
    table, a control outside the tile is a per-tile condition).  QAOA+-style circuits (cx . rz . cx fused to
    controlled diagonals, with rz and u1 in the middle) mixed with cz, X-frame flips and non-diagonal gates in between, amplitudes
    against the oracle for tiles small enough that every placement of control and target occurs."""
    import networkx as nx
    rng = np.random.default_rng(100 * n + tile_bits)
    G = nx.gnm_random_graph(n, 3 * n, seed=n)
    insts = [("h", (q,), (), None) for q in range(n)]
    for k, (a, b) in enumerate(G.edges()):
        g = float(rng.uniform(0, 2 * np.pi))
        if k % 5 == 0:
            insts.append(("x", (int(rng.integers(n)),), (), None))
        if k % 4 == 0:
            insts.append(("cz", (a, b), (), None))
        elif k % 4 == 1:
            insts += [("cx", (b, a), (), None), ("u1", (a,), (g,), None), ("cx", (b, a), (), None)]
        else:
            insts += [("rz", (a,), (g / 2,), None), ("rz", (b,), (g / 2,), None), ("cx", (a, b), (), None),
                      ("rz", (b,), (-g / 2,), None), ("cx", (a, b), (), None)]
        if k % 7 == 6:
            insts.append(("ry", (int(rng.integers(n)),), (float(rng.uniform(0, 3)),), None))      # breaks the diagonal block
    insts += [("rx", (q,), (float(rng.uniform(0, 3)),), None) for q in range(n)]
    be = _backend(cuda, tile_bits=tile_bits, coalesce_bits=coalesce)
    prog, auto = be.compile_instructions(n, insts)
    got = be.run(prog, auto, None)[0].cpu().numpy()
    want = sv.simulate(n, insts)
    idx = np.arange(1 << n) ^ prog.flip_mask
    assert np.abs(got[idx] - want).max() < TOL
