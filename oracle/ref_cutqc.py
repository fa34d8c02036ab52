"""ORACLE (test infrastructure): drives the reference's OWN cutqc C programs, compiled from
/root/reference/cutqc/*.c into oracle/_ref/ by oracle/Makefile (gcc + the mkl.h shim), through
the argv + text-file protocol that cutqc/main.py uses:

  build  : cutqc/main.py:418-457  ->  ./build build_file data_folder dest_folder rank layer
  merge  : cutqc/main.py:391-416  ->  ./merge merge_file data_folder dest_folder rank layer
  vertical_collapse : main.py:351-389 -> ./vertical_collapse size terms_file eval vc early rank
  measure: cutqc/main.py:270-301  ->  ./measure rank eval_folder size subcircuit n idx...

Nothing here is product code; only tests/ and bench.py's reference / cpu_baseline legs use it.
"""
from __future__ import annotations

import os
import subprocess
import time

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(_HERE, "_ref")


def available() -> bool:
    return all(os.path.exists(os.path.join(REF_DIR, b))
               for b in ("build", "merge", "vertical_collapse", "measure"))


def _fmt(vec) -> str:
    return " ".join("%e" % float(v) for v in vec) + " "


def write_kron(folder, subcircuit_idx, kron_idx, vec, header="num_active"):
    """kron_%d_%d.txt: `num_active %d` (after merge) or `num_effective %d` (after collapse)."""
    n = int(np.log2(len(vec)))
    with open(os.path.join(folder, f"kron_{subcircuit_idx}_{kron_idx}.txt"), "w") as f:
        f.write(f"{header} {n}\n")
        f.write(_fmt(vec))


def read_kron(folder, subcircuit_idx, kron_idx):
    path = os.path.join(folder, f"kron_{subcircuit_idx}_{kron_idx}.txt")
    if not os.path.exists(path):
        return None
    with open(path) as f:
        f.readline()
        return np.array(f.readline().split(), dtype=np.float64)


def write_build_file(path, total_active, num_subcircuits, terms, num_cuts):
    """terms: list of summation terms, each a list of (subcircuit_idx, kron_idx) in smart_order
    (cutqc/distributor.py:175-187)."""
    with open(path, "w") as f:
        f.write(f"total_active_qubit={total_active} num_subcircuits={num_subcircuits} "
                f"num_summation_terms={len(terms)} num_cuts={num_cuts}\n")
        for term in terms:
            f.write(" ".join(f"{s},{k}" for s, k in term) + " \n")


def run_build(data_folder, dest_folder, total_active, num_subcircuits, terms, num_cuts,
              num_ranks=1):
    """Stripe `terms` over `num_ranks` processes like cutqc/distributor.py:145-187, run them
    concurrently, sum the per-rank vectors like cutqc/main.py:439-451.  Returns
    (reconstructed float64 vector, wall seconds of the rank processes, wall seconds total)."""
    os.makedirs(dest_folder, exist_ok=True)
    t0 = time.time()
    procs = []
    for rank in range(num_ranks):
        mine = terms[rank::num_ranks]
        bf = os.path.join(dest_folder, f"build_{rank}.txt")
        write_build_file(bf, total_active, num_subcircuits, mine, num_cuts)
        procs.append(subprocess.Popen([os.path.join(REF_DIR, "build"), bf, data_folder,
                                       dest_folder, str(rank), "0"]))
    t1 = time.time()
    for p in procs:
        p.wait()
    t2 = time.time()
    total = None
    for rank in range(num_ranks):
        with open(os.path.join(dest_folder, f"rank_{rank}_summary.txt")) as f:
            lines = f.read().strip().split("\n")
        assert lines[-1] == "DONE"                       # cutqc/main.py:436
        with open(os.path.join(dest_folder, f"reconstructed_prob_{rank}.txt")) as f:
            vec = np.array(f.readline().split(), dtype=np.float64)
        total = vec if total is None else total + vec
    t3 = time.time()
    return total, t2 - t1, t3 - t0


def run_merge(merge_file, data_folder, dest_folder, rank=0):
    os.makedirs(dest_folder, exist_ok=True)
    subprocess.run([os.path.join(REF_DIR, "merge"), merge_file, data_folder, dest_folder,
                    str(rank), "0"], check=True)


def run_vertical_collapse(full_circ_size, terms_file, eval_folder, vc_folder,
                          early_termination=0, rank=0):
    os.makedirs(vc_folder, exist_ok=True)
    subprocess.run([os.path.join(REF_DIR, "vertical_collapse"), str(full_circ_size), terms_file,
                    eval_folder, vc_folder, str(early_termination), str(rank)], check=True)


def run_measure(eval_folder, full_circ_size, subcircuit_idx, instance_indices, rank=0):
    subprocess.run([os.path.join(REF_DIR, "measure"), str(rank), eval_folder,
                    str(full_circ_size), str(subcircuit_idx), str(len(instance_indices))]
                   + [str(i) for i in instance_indices], check=True)


def time_build_sample(active0, active1, num_cuts, num_ranks, workdir, seed=0):
    """Bounded sample of the bench workload for the reference C path: two subcircuits with
    `active0` / `active1` active qubits, 4^num_cuts summation terms (term t = kron t of each
    subcircuit).  Returns (probs_per_second, seconds, reconstructed vector, (V0, V1))."""
    rng = np.random.default_rng(seed)
    nt = 4 ** num_cuts
    data = os.path.join(workdir, "data")
    dest = os.path.join(workdir, "dest")
    os.makedirs(data, exist_ok=True)
    V0 = rng.random((nt, 1 << active0)).astype(np.float32)
    V1 = rng.random((nt, 1 << active1)).astype(np.float32)
    for t in range(nt):
        write_kron(data, 0, t, V0[t])
        write_kron(data, 1, t, V1[t])
    terms = [[(0, t), (1, t)] for t in range(nt)]
    vec, t_ranks, t_total = run_build(data, dest, active0 + active1, 2, terms, num_cuts,
                                      num_ranks)
    return float(1 << (active0 + active1)) / t_total, t_total, vec, (V0, V1)


# ---------------------------------------------------------------------------------------------
# the whole post-simulation pipeline through the reference's C programs
# ---------------------------------------------------------------------------------------------
def write_raw(eval_folder, subcircuit_idx, index, counter, inits, meas, probs):
    """raw_%d_%d.txt as cutqc/evaluator.py:134-153 writes it."""
    with open(os.path.join(eval_folder, f"raw_{subcircuit_idx}_{index}.txt"), "w") as f:
        f.write("d=%d effective=%d\n" % (counter[subcircuit_idx]["d"], counter[subcircuit_idx]["effective"]))
        f.write("".join("%s " % x for x in inits) + "\n")
        f.write("".join("%s " % x for x in meas) + "\n")
        f.write("".join("%e " % x for x in probs))


def write_kron_terms_file(path, kronecker_terms, counter):
    """subcircuit_kron_terms_%d.txt for one rank holding every term (cutqc/main.py:335-349)."""
    with open(path, "w") as f:
        f.write("%d subcircuits\n" % len(kronecker_terms))
        for s in kronecker_terms:
            f.write("subcircuit %d kron_terms %d num_effective %d\n"
                    % (s, len(kronecker_terms[s]), counter[s]["effective"]))
            for term, k in kronecker_terms[s].items():
                f.write("subcircuit_kron_index=%d kron_term_len=%d\n" % (k, len(term)))
                f.write("".join("%d,%d " % (c, i) for c, i in term) + "\n")


def run_pipeline(workdir, full_circ_size, counter, raw_instances, kronecker_terms, write_layer_files,
                 early_termination=0):
    """raw_instances[s] = {mutated index: (inits, meas, probabilities)}.  Runs measure ->
    vertical_collapse -> merge -> build exactly as cutqc/main.py:270-457 launches them (one
    rank) and returns the reconstructed vector.  `write_layer_files(vc_folder, dd_folder)` must
    write merge_0.txt / build_0.txt (cutqc/distributor.py:143-187)."""
    ev = os.path.join(workdir, "eval")
    vc = os.path.join(workdir, "vc")
    dd = os.path.join(workdir, "dd")
    for d in (ev, vc, dd):
        os.makedirs(d, exist_ok=True)
    for s, insts in raw_instances.items():
        for idx, (inits, meas, probs) in insts.items():
            write_raw(ev, s, idx, counter, inits, meas, probs)
        run_measure(ev, full_circ_size, s, sorted(insts))
    terms_file = os.path.join(ev, "subcircuit_kron_terms_0.txt")
    write_kron_terms_file(terms_file, kronecker_terms, counter)
    run_vertical_collapse(full_circ_size, terms_file, ev, vc, early_termination)
    write_layer_files(vc, dd)
    run_merge(os.path.join(dd, "merge_0.txt"), vc, dd)
    with open(os.path.join(dd, "rank_0_summary.txt")) as f:
        lines = f.readlines()
    assert lines[-2].split(" = ")[0] == "Total merge time" and lines[-1] == "DONE"   # main.py:410
    subprocess.run([os.path.join(REF_DIR, "build"), os.path.join(dd, "build_0.txt"), dd, dd, "0", "0"],
                   check=True)
    with open(os.path.join(dd, "rank_0_summary.txt")) as f:
        lines = f.readlines()
    assert lines[-2].split(" = ")[0] == "Total build time" and lines[-1] == "DONE"   # main.py:436
    with open(os.path.join(dd, "reconstructed_prob_0.txt")) as f:
        return np.array(f.readline().split(), dtype=np.float64)
