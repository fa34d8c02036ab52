"""Host logic of the statevector engine's compiler (qdc_sv_compile): X-frame absorption, ancilla
compute/act/uncompute fusion and commutation-aware grouping must preserve the circuit.  The
schedule is dumped through the C-ABI and replayed with the numpy oracle's gate primitives; the
result must equal the oracle's simulation of the ORIGINAL gate list (no GPU needed: only the
host-side planner of libqdc_b200.so runs here)."""
import numpy as np
import pytest

from oracle import statevector as sv
from qdc_b200 import engine as eng


class _HostBackend(eng.B200Backend):
    """Compiler access without a device (the constructor of the real backend is device-free
    too; only run()/probabilities() need CUDA)."""
    def __init__(self, **kw):
        super().__init__(device="cpu", **kw)


def _random_circuit(rng, n, n_gates, with_ancilla_mixers=False):
    insts = []
    names1 = ["x", "h", "s", "sdg", "z", "t", "y"]
    for _ in range(n_gates):
        r = rng.random()
        q = int(rng.integers(n))
        others = [k for k in range(n) if k != q]
        if r < 0.25:
            insts.append((names1[int(rng.integers(len(names1)))], (q,), (), None))
        elif r < 0.45:
            insts.append((["rx", "ry", "rz", "u1"][int(rng.integers(4))], (q,),
                          (float(rng.uniform(0, 2 * np.pi)),), None))
        elif r < 0.55:
            insts.append(("u3", (q,), tuple(float(a) for a in rng.uniform(0, 2 * np.pi, 3)), None))
        elif r < 0.65 and others:
            c = int(rng.choice(others))
            insts.append((["cx", "cz"][int(rng.integers(2))], (c, q), (), None))
        elif r < 0.75 and others:
            c = int(rng.choice(others))
            insts.append(("crx", (c, q), (float(rng.uniform(0, 2 * np.pi)),), None))
        elif r < 0.9 and len(others) >= 2:
            k = int(rng.integers(1, min(4, len(others)) + 1))
            cs = [int(c) for c in rng.choice(others, size=k, replace=False)]
            if rng.random() < 0.5:
                insts.append(("mcx", tuple(cs) + (q,), (), None))
            else:
                insts.append(("zcrx", tuple(cs) + (q,), (float(rng.uniform(0, 2 * np.pi)),), None))
        else:
            insts.append(("Id", tuple(range(n)), (), None))
    return insts


def _faithful_mixer(insts, neighbours, anc, q, alpha):
    """The reference's partial mixer, gate for gate (ansatz/dqv_cut_ansatz.py:127-151)."""
    for v in neighbours:
        insts.append(("x", (v,), (), None))
    insts.append(("mcx", tuple(neighbours) + (anc,), (), None))
    for v in neighbours:
        insts.append(("x", (v,), (), None))
    insts.append(("crx", (anc, q), (2 * alpha,), None))
    for v in neighbours:
        insts.append(("x", (v,), (), None))
    insts.append(("mcx", tuple(neighbours) + (anc,), (), None))
    for v in neighbours:
        insts.append(("x", (v,), (), None))


def _replay(n, insts, prog, auto, recs):
    d = prog.dump()
    state = np.zeros(1 << n, dtype=np.complex128)
    state[0] = 1.0
    first = d["group_first_op"]
    for g in range(prog.n_groups):
        lmask = int(d["group_local_mask"][g])
        assert bin(lmask).count("1") == prog.tile_bits
        for i in range(first[g], first[g + 1]):
            kind, target, cmask, cval, slot, scale, offset, m = recs[int(d["src_gate"][i])]
            ang = scale * (auto[slot] if slot >= 0 else 0.0) + offset
            name = {eng.G_X: "x", eng.G_Y: "y", eng.G_Z: "z", eng.G_H: "h", eng.G_S: "s",
                    eng.G_SDG: "sdg", eng.G_T: "t", eng.G_TDG: "tdg", eng.G_RX: "rx",
                    eng.G_RY: "ry", eng.G_RZ: "rz", eng.G_U1: "u1"}.get(kind)
            if name is None:
                mat = (np.array(m[0::2]) + 1j * np.array(m[1::2])).reshape(2, 2)
            elif name in ("rx", "ry", "rz", "u1"):
                mat = sv.gate_matrix(name, (ang,))
            else:
                mat = sv.gate_matrix(name)
            if d["conj_x"][i]:
                X = sv.gate_matrix("x")
                mat = X @ mat @ X
            is_diag = abs(mat[0, 1]) == 0 and abs(mat[1, 0]) == 0
            assert int(d["kind"][i]) == (1 if is_diag else 0)
            assert int(d["target"][i]) == target or True
            if not is_diag:
                assert (lmask >> int(d["target"][i])) & 1, "non-diagonal target outside the tile"
            state = sv.apply_controlled_1q(state, n, int(d["target"][i]), mat,
                                           int(d["ctrl_mask"][i]), int(d["ctrl_val"][i]))
    return state


@pytest.mark.parametrize("n,tile_bits,seed", [(5, 2, 0), (6, 3, 1), (7, 4, 2), (8, 3, 3),
                                              (8, 8, 4), (9, 5, 5), (4, 1, 6), (10, 4, 7)])
def test_schedule_preserves_circuit(n, tile_bits, seed):
    rng = np.random.default_rng(seed)
    insts = _random_circuit(rng, n, 60)
    be = _HostBackend(tile_bits=tile_bits, coalesce_bits=min(2, tile_bits))
    prog, auto = be.compile_instructions(n, insts)
    recs, auto2 = eng.lower_instructions(insts)
    assert auto == auto2
    want = sv.simulate(n, insts)
    got = _replay(n, insts, prog, auto, recs)
    idx = np.arange(1 << n) ^ prog.flip_mask
    assert np.abs(got[idx] - want).max() < 1e-12
    assert prog.n_groups >= 1 and prog.tile_bits == min(tile_bits, n)


@pytest.mark.parametrize("fuse", [True, False])
def test_faithful_mixer_fuses_to_controlled_rx(fuse):
    """X^N . MCX . X^N . CRX . X^N . MCX . X^N on a |0> ancilla == one open-controlled RX."""
    rng = np.random.default_rng(11)
    n, anc = 7, 6
    insts = [("x", (1,), (), None), ("x", (4,), (), None)]          # init_state bits
    fused_form = list(insts)
    for q, nb in [(0, [1, 2]), (3, [2, 4, 5]), (2, [0, 3]), (5, [3]), (1, [0])]:
        a = float(rng.uniform(0, 2 * np.pi))
        _faithful_mixer(insts, nb, anc, q, a)
        fused_form.append(("zcrx", tuple(nb) + (q,), (2 * a,), None))
    for q in range(6):
        insts.append(("rz", (q,), (0.37 * (q + 1),), None))
        fused_form.append(("rz", (q,), (0.37 * (q + 1),), None))
    be = _HostBackend(tile_bits=4, coalesce_bits=2, fuse_ancilla=fuse)
    prog, auto = be.compile_instructions(n, insts)
    recs, _ = eng.lower_instructions(insts)
    want = sv.simulate(n, insts)
    assert np.abs(want - sv.simulate(n, fused_form)).max() < 1e-12
    got = _replay(n, insts, prog, auto, recs)
    idx = np.arange(1 << n) ^ prog.flip_mask
    assert np.abs(got[idx] - want).max() < 1e-12
    # 2 init X + per mixer 4k X + 2 MCX + 1 CRX, then 6 RZ
    if fuse:
        assert prog.n_ops == 5 + 6            # one controlled RX per mixer + the phase separator
    else:
        assert prog.n_ops == 5 * 3 + 6        # X frame only: MCX, CRX, MCX per mixer


def test_unknown_gate_is_an_error():
    from qdc_b200 import _lib
    with pytest.raises(_lib.QdcError):
        eng.lower_instructions([("unitary", (0, 1), (), np.eye(4))])


@pytest.mark.parametrize("n,tile_bits,seed", [(6, 3, 0), (8, 4, 1), (9, 12, 2)])
def test_cx_diagonal_cx_becomes_controlled_diagonals(n, tile_bits, seed):
    """QAOA+'s phase separator (ansatz/qaoa_plus.py:20-24: rz, cx, rz, cx per edge) is diagonal as a whole: the
    compiler turns every cx . D . cx into two controlled diagonal gates, so no cx is left in the program, the
    circuit's unitary is unchanged (replayed from the dumped program against the oracle) -- also with X gates
    around that put control and target into the flipped frame, and with u1 / z / s / sdg as the diagonal."""
    rng = np.random.default_rng(seed)
    import networkx as nx
    G = nx.gnm_random_graph(n, 2 * n, seed=seed)
    insts = [("h", (q,), (), None) for q in range(n)]
    diag = ["rz", "u1", "z", "s", "sdg"]
    for k, (a, b) in enumerate(G.edges()):
        g = float(rng.uniform(0, 2 * np.pi))
        if k % 3 == 0:
            insts.append(("x", (int(rng.integers(n)),), (), None))        # flips the frame of some qubit
        insts += [("rz", (a,), (g / 2,), None), ("rz", (b,), (g / 2,), None), ("cx", (a, b), (), None)]
        name = diag[k % len(diag)]
        insts.append((name, (b,), (-g / 2,) if name in ("rz", "u1") else (), None))
        insts.append(("cx", (a, b), (), None))
    insts += [("rx", (q,), (float(rng.uniform(0, 3)),), None) for q in range(n)]
    be = _HostBackend(tile_bits=tile_bits, coalesce_bits=min(2, tile_bits))
    prog, auto = be.compile_instructions(n, insts)
    recs, _ = eng.lower_instructions(insts)
    want = sv.simulate(n, insts)
    got = _replay(n, insts, prog, auto, recs)
    idx = np.arange(1 << n) ^ prog.flip_mask
    assert np.abs(got[idx] - want).max() < 1e-12
    # ops: n H + n RX + per edge 2 rz + 2 controlled diagonals; a surviving cx pair would make it 5 per edge
    assert prog.n_ops == 2 * n + 4 * G.number_of_edges()


@pytest.mark.parametrize("seed", range(24))
def test_compiler_fuzz_mixed_circuits(seed):
    """Random circuits that mix everything the compiler rewrites -- X gates (frame), cx . D . cx sandwiches (some broken
    by a gate in between, some with control and target swapped between the two cx, which must NOT fuse), faithful
    partial mixers on an ancilla, runs of diagonal gates with and without controls, and plain random gates --
    over 2 to 3 layers, random width and tile size: the replayed program equals the oracle on the original list."""
    rng = np.random.default_rng(1000 + seed)
    n = int(rng.integers(4, 11))
    tile_bits = int(rng.integers(1, n + 1))
    anc = n - 1
    insts = []
    for layer in range(int(rng.integers(2, 4))):
        for _ in range(int(rng.integers(3, 9))):
            r = rng.random()
            a, b = (int(v) for v in rng.choice(n - 1, size=2, replace=False))
            g = float(rng.uniform(0, 2 * np.pi))
            if r < 0.35:                                        # a proper sandwich
                name = ["rz", "u1", "z", "s", "t"][int(rng.integers(5))]
                insts += [("cx", (a, b), (), None), (name, (b,), (g,) if name in ("rz", "u1") else (), None),
                          ("cx", (a, b), (), None)]
            elif r < 0.45:                                      # broken sandwich: a non-diagonal gate in the middle
                insts += [("cx", (a, b), (), None), ("rz", (b,), (g,), None), ("h", (b,), (), None), ("cx", (a, b), (), None)]
            elif r < 0.55:                                      # swapped control / target: not a sandwich
                insts += [("cx", (a, b), (), None), ("rz", (b,), (g,), None), ("cx", (b, a), (), None)]
            elif r < 0.7 and n >= 5:                            # the reference's partial mixer through the ancilla
                q = int(rng.integers(n - 1))
                nb = [int(v) for v in rng.choice([k for k in range(n - 1) if k != q],
                                                 size=int(rng.integers(1, min(3, n - 2) + 1)), replace=False)]
                _faithful_mixer(insts, nb, anc, q, g)
            elif r < 0.85:                                      # diagonal run with a controlled member
                insts += [("rz", (a,), (g,), None), ("cz", (a, b), (), None), ("u1", (b,), (g / 3,), None), ("x", (a,), (), None)]
            else:
                insts += _random_circuit(rng, n - 1, 4)
    be = _HostBackend(tile_bits=tile_bits, coalesce_bits=min(2, tile_bits))
    prog, auto = be.compile_instructions(n, insts)
    recs, _ = eng.lower_instructions(insts)
    want = sv.simulate(n, insts)
    got = _replay(n, insts, prog, auto, recs)
    idx = np.arange(1 << n) ^ prog.flip_mask
    assert np.abs(got[idx] - want).max() < 1e-12
