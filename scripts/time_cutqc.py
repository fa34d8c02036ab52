"""Device timing of the CutQC-style stage kernels (csrc/cutqc.cu: measure / vertical_collapse /
merge / build) at C3-like sizes, with the algorithmic bytes each moves and the fraction of the
measured HBM copy peak (MEASURED_PEAKS.json).  CUDA events, 1 warm-up + best of 5."""
import importlib, json, os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
importlib.import_module("dqva-and-circuit-cutting_b200")
from qdc_b200.cutqc import kernels as ck

dev = torch.device("cuda:0")
try:
    HBM = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
except (OSError, KeyError, ValueError):
    HBM = 6500.0


def best(fn, reps=5):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e-3)
    return min(ts)


def report(name, t, nbytes, note=""):
    gbs = nbytes / t / 1e9
    print(f"{name:34s} {t * 1e3:9.3f} ms  {nbytes / 1e6:10.1f} MB  {gbs:8.1f} GB/s  {100 * gbs / HBM:5.1f} % of copy peak  {note}")


g = torch.Generator(device="cpu").manual_seed(0)
# measure: one subcircuit of d qubits, rho = 2 cut inputs, O = 2 cut outputs -> 4^2 * 3^2 = 144 instances
for d, n_inst in [(16, 144), (20, 144), (22, 36)]:
    raw = torch.rand(n_inst, 1 << d, dtype=torch.float32, generator=g).to(dev)
    keep = (1 << d) - 1 - 0b11                       # two measured cut qubits are summed out with signs
    signs = [i % 4 for i in range(n_inst)]
    t = best(lambda: ck.measure(raw, keep, signs))
    report(f"measure d={d} x{n_inst} (f32 in)", t, n_inst * ((1 << d) * 4 + (1 << (d - 2)) * 8))

# vertical_collapse: 4^K kron terms from the measured instances (each term <= 2^K instances)
for eff, n_inst, n_terms in [(14, 144, 256), (18, 144, 256), (20, 36, 64)]:
    meas = torch.rand(n_inst, 1 << eff, dtype=torch.float64, generator=g).to(dev)
    rng = np.random.default_rng(1)
    terms = [[(int(rng.choice([1, -1, 2])), int(rng.integers(n_inst))) for _ in range(4)] for _ in range(n_terms)]
    t = best(lambda: ck.vertical_collapse(meas, terms))
    report(f"collapse eff={eff} {n_terms} terms x4", t, n_terms * (1 << eff) * 8 * (4 + 1))

# merge: sum out merged qubits, keep active ones
for eff, n_vec in [(18, 256), (22, 64)]:
    vecs = torch.rand(n_vec, 1 << eff, dtype=torch.float64, generator=g).to(dev)
    states = [-1 if q % 2 == 0 else -2 for q in range(eff)]          # half active, half merged
    t = best(lambda: ck.merge(vecs, states))
    report(f"merge eff={eff} x{n_vec} (half merged)", t, n_vec * ((1 << eff) + (1 << (eff // 2))) * 8)

# build: 0.5^K sum_t kron(v0[t], v1[t]) -> 2^(a0 + a1) outputs, float32 out as build.c writes
for a0, a1, K in [(12, 12, 4), (14, 14, 4), (15, 15, 3)]:
    n_terms = 4 ** K
    k0 = torch.rand(n_terms, 1 << a0, dtype=torch.float64, generator=g).to(dev)
    k1 = torch.rand(n_terms, 1 << a1, dtype=torch.float64, generator=g).to(dev)
    idx = [list(range(n_terms)), list(range(n_terms))]
    t = best(lambda: ck.build([k0, k1], idx, K, out_dtype=torch.float32), reps=3)
    outs = float(1 << (a0 + a1))
    report(f"build {a0}+{a1} active, {n_terms} terms", t, outs * 4,
           f"{outs / t / 1e9:.1f} Gprobs/s stored f32, {2 * outs * n_terms / t / 1e12:.1f} TFLOP/s FP64")
    for S in (6, 5, 4):
        t = best(lambda: ck.build([k0, k1], idx, K, out_dtype=torch.float32, contraction=f"i8:{S}"), reps=3)
        report(f"build {a0}+{a1} (int8 x {S} slices)", t, outs * 4, f"{outs / t / 1e9:.1f} Gprobs/s stored f32")
