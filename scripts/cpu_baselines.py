"""CPU baselines of SURVEY.md 8(d) on the host cores of the box it runs on (no GPU work):
 (i)   the reference's own C build path (oracle/_ref/build) at 2^20 and 2^24 outputs, K = 4;
 (ii)  the line-faithful Python restatement of recombine_fragment_models (per-output loop);
 (iii) best-effort numpy: A @ B.T in row slabs + vectorised popcount / independent-set scoring;
 (iv)  numpy statevector oracle, fragment-variant simulations per second (stand-in for Aer,
       which is not installable here).
Test infrastructure only (uses oracle/); prints one line per measurement."""
import importlib, os, sys, tempfile, time
import networkx as nx, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
importlib.import_module("dqva-and-circuit-cutting_b200")
from oracle import mis_oracle as mo, qsplit_oracle as qo, ref_cutqc, statevector as sv

cores = os.cpu_count()
print(f"host cores: {cores}", flush=True)

# (i) reference C
if ref_cutqc.available():
    for bits in (20, 24):
        with tempfile.TemporaryDirectory() as d:
            rate, secs, _, _ = ref_cutqc.time_build_sample(bits // 2, bits - bits // 2, 4, cores, d)
        print(f"(i) reference cutqc/build.c, {cores} ranks, K=4, 2^{bits} outputs: {secs:.1f} s -> {rate:.3e} probs/s "
              f"(gcc -O2 + plain-C BLAS shim, not MKL; text-file protocol included)", flush=True)

# (ii) line-faithful Python recombination
from test_oracle_golden import _cut_dqva
G = nx.Graph(); G.add_edges_from([(0, 1), (1, 2), (2, 0), (2, 3), (3, 4), (4, 5), (5, 3), (1, 4), (5, 6), (6, 7), (7, 4)])
_, _, frags, wpm = _cut_dqva(G, [[0, 1, 2, 3], [4, 5, 6, 7]], 2, 3, True)
data = qo.collect_fragment_data(frags, wpm)
models = qo.maximum_likelihood_model(qo.direct_fragment_model(data))
n_out = 1
for m in models: n_out *= len(m)
reps, t0 = 0, time.time()
while time.time() - t0 < 2.0:
    dist = qo.recombine_fragment_models(models, wpm); reps += 1
t = (time.time() - t0) / reps
print(f"(ii) line-faithful Python recombine_fragment_models (network method), {n_out} outputs visited "
      f"(2^{len(next(iter(dist)))} bitstrings): {t:.2f} s -> {n_out / t:.3e} probs/s on 1 core", flush=True)

# (iii) numpy best effort
rng = np.random.default_rng(0)
rb, cb, kdim = 9, 17, 256
A = rng.random((1 << rb, kdim)); Bt = rng.random((1 << cb, kdim))
Gs = nx.random_regular_graph(3, rb + cb, seed=1)
adj = mo.adjacency_masks(Gs, rb + cb)
t0 = time.time(); tot = 0.0
for r0 in range(0, 1 << rb, 256):
    P = (A[r0:r0 + 256] @ Bt.T) * (0.5 ** 4)
    s_p, s_pw = mo.score_array(P.reshape(-1), rb + cb, 0, adj, lo=r0 << cb)
    tot += s_pw
t = time.time() - t0
print(f"(iii) numpy/OpenBLAS A @ B.T + vectorised scoring, 2^{rb + cb} outputs: {t:.1f} s -> "
      f"{2.0 ** (rb + cb) / t:.3e} probs/s on {cores} cores", flush=True)

# (iv) numpy statevector oracle
for n_nodes in (9, 13, 16, 20):
    Gf = nx.random_regular_graph(3, n_nodes + (n_nodes % 2), seed=7)
    nodes = list(Gf.nodes)[:n_nodes]
    insts = []
    for q in rng.permutation(n_nodes):
        nb = [int(v) for v in Gf.neighbors(int(q)) if v < n_nodes]
        insts.append(("zcrx", tuple(nb) + (int(q),), (float(rng.uniform(0, 6.28)),), None))
    insts += [("rz", (q,), (0.3,), None) for q in range(n_nodes)]
    n = n_nodes + 1
    reps = max(1, int(2e6 / (1 << n)))
    t0 = time.time()
    for _ in range(reps):
        sv.probabilities(sv.simulate(n, insts))
    t = (time.time() - t0) / reps
    print(f"(iv) numpy statevector oracle, {n}-qubit DQVA fragment ({len(insts)} fused gates): {t * 1e3:.2f} ms per "
          f"variant -> {1 / t:.3e} sims/s on 1 core (qiskit-Aer itself: unavailable on this box)", flush=True)
