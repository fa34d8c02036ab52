"""GPU parity of the CutQC-style stages.  (1) File-protocol shims (`qdc_b200.cutqc.stages`, the
argv/files of cutqc/main.py:298,383,400,425) against the reference's OWN C programs compiled into
oracle/_ref (float32 + `%e` text: tolerance 2e-6 relative to the vector's scale); (2) the
in-memory kernels against float64 numpy restatements of the same formulas (1e-10)."""
import os
import shutil

import numpy as np
import pytest
import torch

from oracle import ref_cutqc

pytestmark = pytest.mark.gpu


def _close(a, b, tol=2e-6):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    assert a.shape == b.shape
    assert np.abs(a - b).max() <= tol * max(1.0, np.abs(b).max())


def _read_vec(path, header=None):
    with open(path) as f:
        if header is not None:
            first = f.readline().split()
            assert first[0] == header
            return int(first[1]), np.array(f.read().split(), dtype=np.float64)
        return np.array(f.read().split(), dtype=np.float64)


def _write_raw(folder, sub, inst, d, eff, init, meas, probs):
    with open(os.path.join(folder, f"raw_{sub}_{inst}.txt"), "w") as f:
        f.write(f"d={d} effective={eff}\n")
        f.write(" ".join(init) + "\n" + " ".join(meas) + "\n")
        f.write(" ".join("%e" % p for p in probs) + " ")


needs_ref = pytest.mark.skipif(not ref_cutqc.available(), reason="oracle/_ref binaries not built")


@needs_ref
def test_measure_files_match_reference(cuda, tmp_path):
    from qdc_b200.cutqc import stages
    rng = np.random.default_rng(0)
    d = 7
    comp = [0, 2, 3, 6]                                   # effective qubits
    ref_dir, our_dir = tmp_path / "ref", tmp_path / "our"
    for folder in (ref_dir, our_dir):
        os.makedirs(folder)
    insts = list(range(5))
    for i in insts:
        meas = ["comp" if q in comp else ["I", "X", "Y", "Z"][int(rng.integers(4))] for q in range(d)]
        init = [["zero", "one", "plus", "plusI"][int(rng.integers(4))] for _ in range(d)]
        probs = rng.random(1 << d).astype(np.float32)
        for folder in (ref_dir, our_dir):
            _write_raw(folder, 1, i, d, len(comp), init, meas, probs)
    ref_cutqc.run_measure(str(ref_dir), 12, 1, insts, rank=3)
    stages.measure(3, str(our_dir), 12, 1, insts, device=cuda)
    for i in insts:
        assert not os.path.exists(our_dir / f"raw_1_{i}.txt")
        _close(_read_vec(our_dir / f"measured_1_{i}.txt"), _read_vec(ref_dir / f"measured_1_{i}.txt"))
    assert open(our_dir / "rank_3_summary.txt").read().strip().split("\n")[-1] == "measure DONE"


@needs_ref
@pytest.mark.parametrize("early", [0, 1])
def test_vertical_collapse_files_match_reference(cuda, tmp_path, early):
    from qdc_b200.cutqc import stages
    rng = np.random.default_rng(1)
    n_eff = 5
    ev_ref, ev_our, vc_ref, vc_our = (tmp_path / n for n in ("ev_ref", "ev_our", "vc_ref", "vc_our"))
    for folder in (ev_ref, ev_our, vc_ref, vc_our):
        os.makedirs(folder)
    for sub, n_inst in ((0, 6), (1, 4)):
        for i in range(n_inst):
            v = rng.random(1 << n_eff).astype(np.float32)
            for folder in (ev_ref, ev_our):
                with open(folder / f"measured_{sub}_{i}.txt", "w") as f:
                    f.write(" ".join("%e" % x for x in v) + " ")
    terms = {0: [(3, [(1, 0), (1, 1)]), (7, [(2, 2), (-1, 0), (-1, 1)]), (8, [(1, 4), (-1, 4)]),
                 (9, [(1, 5)])],
             1: [(0, [(1, 0), (-1, 3)]), (2, [(2, 1), (-1, 2), (-1, 3)])]}
    for folder in (ev_ref, ev_our):
        with open(folder / "subcircuit_kron_terms_0.txt", "w") as f:
            f.write(f"{len(terms)} subcircuits\n")
            for sub, tl in terms.items():
                f.write(f"subcircuit {sub} kron_terms {len(tl)} num_effective {n_eff}\n")
                for kidx, pairs in tl:
                    f.write(f"subcircuit_kron_index={kidx} kron_term_len={len(pairs)}\n")
                    f.write("".join(f"{c},{i} " for c, i in pairs) + "\n")
    ref_cutqc.run_vertical_collapse(10, str(ev_ref / "subcircuit_kron_terms_0.txt"), str(ev_ref),
                                    str(vc_ref), early_termination=early, rank=0)
    stages.vertical_collapse(10, str(ev_our / "subcircuit_kron_terms_0.txt"), str(ev_our), str(vc_our),
                             early, 0, device=cuda)
    assert sorted(os.listdir(vc_ref)) == sorted(os.listdir(vc_our))
    assert ("kron_0_8.txt" in os.listdir(vc_our)) == (early == 0)       # the all-zero term
    for name in os.listdir(vc_ref):
        n1, v1 = _read_vec(vc_ref / name, "num_effective")
        n2, v2 = _read_vec(vc_our / name, "num_effective")
        assert n1 == n2 == n_eff
        _close(v2, v1)
    assert open(ev_our / "rank_0_summary.txt").read().strip().split("\n")[-1] == "vertical_collapse DONE"


@needs_ref
def test_merge_files_match_reference(cuda, tmp_path):
    from qdc_b200.cutqc import stages
    rng = np.random.default_rng(2)
    data, d_ref, d_our = tmp_path / "data", tmp_path / "ref", tmp_path / "our"
    for folder in (data, d_ref, d_our):
        os.makedirs(folder)
    jobs = [(0, 4, [-1, -2, 1, -1, -2, -1]), (0, 5, [-1, -2, 1, -1, -2, -1]),
            (1, 0, [-2, -2, -1, 0, -1]), (2, 9, [-1, -1, -1]), (3, 1, [-2, -2, 0, 1])]
    with open(tmp_path / "merge_0.txt", "w") as f:
        f.write(f"num_files_to_merge={len(jobs)}\n")
        for sub, k, states in jobs:
            v = rng.random(1 << len(states)).astype(np.float32)
            ref_cutqc.write_kron(str(data), sub, k, v, header="num_effective")
            active = sum(1 for s in states if s == -1)
            f.write(f"subcircuit_idx={sub} subcircuit_kron_index={k} num_effective={len(states)} "
                    f"num_active={active}\n")
            f.write(" ".join(str(s) for s in states) + " \n")
    ref_cutqc.run_merge(str(tmp_path / "merge_0.txt"), str(data), str(d_ref))
    stages.merge(str(tmp_path / "merge_0.txt"), str(data), str(d_our), 0, 0, device=cuda)
    for sub, k, states in jobs:
        n1, v1 = _read_vec(d_ref / f"kron_{sub}_{k}.txt", "num_active")
        n2, v2 = _read_vec(d_our / f"kron_{sub}_{k}.txt", "num_active")
        assert n1 == n2 == sum(1 for s in states if s == -1)
        _close(v2, v1)
    assert open(d_our / "rank_0_summary.txt").read().split("\n")[-1] == "DONE"


@needs_ref
@pytest.mark.parametrize("actives,missing", [([3, 4], None), ([2, 3, 4], None), ([3, 2, 2], (1, 2)),
                                             ([5], None), ([2, 2, 1, 3], (3, 0))])
def test_build_files_match_reference(cuda, tmp_path, actives, missing):
    from qdc_b200.cutqc import stages
    rng = np.random.default_rng(3)
    data, d_ref, d_our = tmp_path / "data", tmp_path / "ref", tmp_path / "our"
    for folder in (data, d_ref, d_our):
        os.makedirs(folder)
    n_cuts, n_kron, n_terms = 2, 5, 16
    for s, a in enumerate(actives):
        for k in range(n_kron):
            if missing == (s, k):
                continue
            ref_cutqc.write_kron(str(data), s, k, (rng.random(1 << a) - 0.3).astype(np.float32))
    terms = [[(s, int(rng.integers(n_kron))) for s in range(len(actives))] for _ in range(n_terms)]
    bf = tmp_path / "build_0.txt"
    ref_cutqc.write_build_file(str(bf), sum(actives), len(actives), terms, n_cuts)
    import subprocess
    subprocess.run([os.path.join(ref_cutqc.REF_DIR, "build"), str(bf), str(data), str(d_ref), "0", "0"],
                   check=True)
    stages.build(str(bf), str(data), str(d_our), 0, 0, device=cuda)
    _close(_read_vec(d_our / "reconstructed_prob_0.txt"), _read_vec(d_ref / "reconstructed_prob_0.txt"))
    lines = open(d_our / "rank_0_summary.txt").read().split("\n")
    assert lines[-1] == "DONE" and lines[-2].startswith("Total build time = ")


def test_kernels_match_float64_formulas(cuda):
    """In-memory kernels vs float64 numpy (CKT's numpy twin of the same math,
    circuit_knitting_toolbox/.../wire_cutting_post_processing.py:340-443)."""
    from qdc_b200.cutqc import kernels
    rng = np.random.default_rng(4)
    # measure
    d, keep = 9, 0b101100110
    raw = rng.random((3, 1 << d))
    signs = [0b010010001, 0, 0b000001001]
    got = kernels.measure(torch.as_tensor(raw, device=cuda), keep, signs).cpu().numpy()
    x = np.arange(1 << d)
    e = np.zeros_like(x)
    for j, q in enumerate([q for q in range(d) if (keep >> q) & 1]):
        e |= ((x >> q) & 1) << j
    for i in range(3):
        sgn = 1 - 2 * (np.array([bin(v & signs[i]).count("1") for v in x]) & 1)
        want = np.bincount(e, weights=raw[i] * sgn, minlength=1 << bin(keep).count("1"))
        assert np.abs(got[i] - want).max() < 1e-10
    # collapse
    meas = rng.random((6, 64))
    terms = [[(1, 0), (1, 1)], [(2, 2), (-1, 0), (-1, 1)], [(1, 3), (-1, 3)], []]
    out, flags = kernels.vertical_collapse(torch.as_tensor(meas, device=cuda), terms)
    for t, term in enumerate(terms):
        want = sum((c * meas[i] for c, i in term), np.zeros(64))
        assert np.abs(out[t].cpu().numpy() - want).max() < 1e-12
    assert flags.cpu().tolist() == [1, 1, 0, 0]
    # merge (MSB-first states)
    states = [-1, -2, 1, -1, -2, -2, -1, 0]
    vec = rng.random((2, 256))
    got = kernels.merge(torch.as_tensor(vec, device=cuda), states).cpu().numpy()
    want = np.zeros((2, 8))
    for s in range(256):
        bits = [(s >> (7 - q)) & 1 for q in range(8)]
        if bits[2] != 1 or bits[7] != 0:
            continue
        a = bits[0] * 4 + bits[3] * 2 + bits[6]
        want[:, a] += vec[:, s]
    assert np.abs(got - want).max() < 1e-12
    # build, 3 subcircuits, one missing kron
    kron = [rng.random((4, 8)), rng.random((3, 16)), rng.random((5, 4))]
    index = [list(rng.integers(0, 4, 20)), list(rng.integers(0, 3, 20)), list(rng.integers(0, 5, 20))]
    index[1][7] = -1
    got = kernels.build([torch.as_tensor(k, device=cuda) for k in kron], index, 3).cpu().numpy()
    want = np.zeros(8 * 16 * 4)
    for t in range(20):
        if index[1][t] < 0:
            continue
        want += np.kron(np.kron(kron[0][index[0][t]], kron[1][index[1][t]]), kron[2][index[2][t]])
    assert np.abs(got - want / 8).max() < 1e-10


@pytest.mark.parametrize("out_dtype", ["f64", "f32"])
def test_build_int8_contraction_matches_fp64(cuda, out_dtype):
    """kernels.build through the int8-sliced contraction (contraction="i8:6") against the FP64
    DMMA kernel: three subcircuits (Khatri-Rao fold), a missing kron file, float64 and float32 out."""
    import torch
    from qdc_b200.cutqc import kernels as ck
    rng = np.random.default_rng(8)
    K, n_terms = 3, 64
    actives = (5, 4, 6)
    kron = [torch.as_tensor(rng.standard_normal((n_terms, 1 << a)) * 10.0 ** rng.uniform(-3, 0, (n_terms, 1)),
                            device=cuda) for a in actives]
    idx = [list(range(n_terms)) for _ in actives]
    idx[1][7] = -1                                         # that summation term is skipped everywhere
    dt = torch.float64 if out_dtype == "f64" else torch.float32
    ref = ck.build(kron, idx, K, out_dtype=dt, contraction="f64").cpu().numpy().astype(np.float64)
    got = ck.build(kron, idx, K, out_dtype=dt, contraction="i8:6").cpu().numpy().astype(np.float64)
    scale = np.abs(ref).max()
    tol = 1e-10 * scale if out_dtype == "f64" else 2e-7 * scale
    assert got.shape == ref.shape == (1 << sum(actives),)
    assert np.abs(got - ref).max() <= tol, np.abs(got - ref).max() / scale


@pytest.mark.parametrize("dtype,d,summed", [
    ("float32", 14, (0, 1)),            # both summed bits inside one 16-byte load of floats
    ("float32", 15, (0, 13)),           # one trailing, one above the tile
    ("float32", 14, (12, 13)),          # high summed bits only: nothing added inside a load
    ("float64", 14, (0,)),              # doubles: one trailing bit per load
    ("float64", 15, (0, 12, 14)),
    ("float32", 13, (1,)),              # bit 1 summed but bit 0 kept: NOT the streaming form (tiled kernel)
])
def test_measure_streaming_form_equals_tiled_form(cuda, monkeypatch, dtype, d, summed):
    """measure / merge when no summed bit lies between the bits a 16-byte load covers and bit 12 go through
    `marginal_direct_kernel` (no shared-memory tile): against float64 numpy, and bit for bit against the tiled
    kernel (QDC_MARG_DIRECT=0) -- same arithmetic in the same order."""
    from qdc_b200.cutqc import kernels
    rng = np.random.default_rng(d + len(summed))
    n_inst = 5
    raw = (rng.random((n_inst, 1 << d)) - 0.3).astype(dtype)
    keep = ((1 << d) - 1) & ~sum(1 << b for b in summed)
    signs = [0] + [int(rng.integers(0, 1 << d)) & ~keep | (int(rng.integers(0, 1 << d)) & keep & 0x5) for _ in range(n_inst - 1)]
    dev = torch.as_tensor(raw, device=cuda)
    out = {}
    for knob in ("1", "0"):
        monkeypatch.setenv("QDC_MARG_DIRECT", knob)
        out[knob] = kernels.measure(dev, keep, signs).cpu().numpy()
    assert np.array_equal(out["1"], out["0"])
    x = np.arange(1 << d)
    e = np.zeros_like(x)
    for j, q in enumerate([q for q in range(d) if (keep >> q) & 1]):
        e |= ((x >> q) & 1) << j
    for i in range(n_inst):
        sgn = 1 - 2 * (np.array([bin(int(v) & signs[i]).count("1") for v in x]) & 1)
        want = np.bincount(e, weights=raw[i].astype(np.float64) * sgn, minlength=1 << bin(keep).count("1"))
        assert np.abs(out["1"][i] - want).max() < 1e-10


def test_merge_streaming_form(cuda, monkeypatch):
    """merge with the merged (summed) qubits at the top and the bottom of the index and no zoomed qubit."""
    from qdc_b200.cutqc import kernels
    rng = np.random.default_rng(9)
    n_eff = 15
    states = [-2, -1, -2] + [-1] * 11 + [-2]               # MSB first: bits 14, 12 and 0 are merged
    vec = rng.random((3, 1 << n_eff))
    dev = torch.as_tensor(vec, device=cuda)
    out = {}
    for knob in ("1", "0"):
        monkeypatch.setenv("QDC_MARG_DIRECT", knob)
        out[knob] = kernels.merge(dev, states).cpu().numpy()
    assert np.array_equal(out["1"], out["0"])
    want = vec.reshape(3, 2, 2, 2, 1 << 11, 2).sum(axis=(1, 3, 5)).reshape(3, -1)
    assert np.abs(out["1"] - want).max() < 1e-12
