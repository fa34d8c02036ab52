"""File-protocol compatible stages: the same argv, input files, output files and summary lines as
the reference's `./cutqc/measure`, `./cutqc/vertical_collapse`, `./cutqc/merge`, `./cutqc/build`
(cutqc/main.py:298, 383, 400, 425), computed on the GPU by `cutqc.kernels`.  The executables in
`cutqc/bin/` forward their argv here, so `CutQC._measure/_vertical_collapse/_merge/_build` can
Popen them unchanged.  Text parsing / `%e` formatting is host work (it is the protocol)."""
from __future__ import annotations

import os
import time
from typing import List

import numpy as np
import torch

from . import kernels

__all__ = ["measure", "vertical_collapse", "merge", "build", "main"]


def _fmt(vec) -> str:
    return " ".join(["%e" % v for v in np.asarray(vec, dtype=np.float64)]) + " "


def _read_raw(path):
    with open(path) as f:
        tokens = f.read().split()
    d = int(tokens[0].split("=")[1])
    eff = int(tokens[1].split("=")[1])
    init = tokens[2:2 + d]
    meas = tokens[2 + d:2 + 2 * d]
    probs = np.array(tokens[2 + 2 * d:2 + 2 * d + (1 << d)], dtype=np.float32)
    return d, eff, init, meas, probs


def measure(rank: int, eval_folder: str, full_circ_size: int, subcircuit_idx: int,
            eval_files: List[int], device=None):
    """measure.c:44-121: raw_%d_%d.txt -> measured_%d_%d.txt (raw file removed), then
    rank_%d_summary.txt ('Total measure time = %e', 'measure DONE')."""
    t0 = time.time()
    raws, signs, keep, d = [], [], 0, 0
    for i in eval_files:
        d, eff, _, meas, probs = _read_raw(os.path.join(eval_folder, f"raw_{subcircuit_idx}_{i}.txt"))
        keep = sum(1 << q for q in range(d) if meas[q] == "comp")
        signs.append(sum(1 << q for q in range(d) if meas[q] not in ("I", "comp")))
        raws.append(probs)
    if raws:
        dev = kernels._dev(device)
        out = kernels.measure(torch.as_tensor(np.stack(raws), device=dev), keep, signs).cpu().numpy()
        for row, i in zip(out, eval_files):
            os.remove(os.path.join(eval_folder, f"raw_{subcircuit_idx}_{i}.txt"))
            with open(os.path.join(eval_folder, f"measured_{subcircuit_idx}_{i}.txt"), "w") as f:
                f.write(_fmt(row))
    with open(os.path.join(eval_folder, f"rank_{rank}_summary.txt"), "w") as f:
        f.write("Total measure time = %e\n" % (time.time() - t0))
        f.write("measure DONE\n")


def vertical_collapse(full_circ_size: int, kron_terms_file: str, eval_folder: str,
                      vc_folder: str, early_termination: int, rank: int, device=None):
    """vertical_collapse.c:27-104."""
    t0 = time.time()
    dev = kernels._dev(device)
    with open(kron_terms_file) as f:
        lines = f.read().split("\n")
    pos = 0
    n_sub = int(lines[pos].split()[0]); pos += 1
    for _ in range(n_sub):
        _, sub, _, n_terms, _, n_eff = lines[pos].split(); pos += 1
        sub, n_terms, n_eff = int(sub), int(n_terms), int(n_eff)
        terms, kron_idx, inst_rows, rows = [], [], {}, []
        for _t in range(n_terms):
            head = dict(kv.split("=") for kv in lines[pos].split()); pos += 1
            kron_idx.append(int(head["subcircuit_kron_index"]))
            pairs = [tuple(map(int, tok.split(","))) for tok in lines[pos].split()]; pos += 1
            assert len(pairs) == int(head["kron_term_len"])
            term = []
            for coeff, inst in pairs:
                if inst not in inst_rows:
                    inst_rows[inst] = len(rows)
                    with open(os.path.join(eval_folder, f"measured_{sub}_{inst}.txt")) as mf:
                        rows.append(np.array(mf.read().split()[:1 << n_eff], dtype=np.float32))
                term.append((coeff, inst_rows[inst]))
            terms.append(term)
        if not terms:
            continue
        measured = torch.as_tensor(np.stack(rows).astype(np.float64), device=dev)
        out, flags = kernels.vertical_collapse(measured, terms)
        out, flags = out.cpu().numpy(), flags.cpu().numpy()
        for row, flag, k in zip(out, flags, kron_idx):
            if early_termination == 1 and not flag:
                continue                                              # :76-78
            with open(os.path.join(vc_folder, f"kron_{sub}_{k}.txt"), "w") as f:
                f.write(f"num_effective {n_eff}\n")
                f.write(_fmt(row))
    with open(os.path.join(eval_folder, f"rank_{rank}_summary.txt"), "a") as f:
        f.write("Total vertical_collapse time = %e\n" % (time.time() - t0))
        f.write("vertical_collapse DONE\n")


def _read_kron(path, header):
    with open(path) as f:
        first = f.readline().split()
        assert first[0] == header, (path, first)
        return int(first[1]), np.array(f.read().split(), dtype=np.float32).astype(np.float64)


def merge(merge_file: str, data_folder: str, dest_folder: str, rank: int, recursion_layer: int = 0,
          device=None):
    """merge.c:28-122."""
    t0 = time.time()
    dev = kernels._dev(device)
    with open(merge_file) as f:
        lines = [ln for ln in f.read().split("\n") if ln.strip()]
    n_files = int(lines[0].split("=")[1])
    groups = {}
    for k in range(n_files):
        head = dict(kv.split("=") for kv in lines[1 + 2 * k].split())
        states = tuple(int(x) for x in lines[2 + 2 * k].split())
        assert len(states) == int(head["num_effective"])
        groups.setdefault((int(head["subcircuit_idx"]), states), []).append(
            int(head["subcircuit_kron_index"]))
    for (sub, states), kron_ids in groups.items():
        vecs = []
        for k in kron_ids:
            n_eff, v = _read_kron(os.path.join(data_folder, f"kron_{sub}_{k}.txt"), "num_effective")
            assert n_eff == len(states)
            vecs.append(v)
        out = kernels.merge(torch.as_tensor(np.stack(vecs), device=dev), states).cpu().numpy()
        n_active = sum(1 for s in states if s == -1)
        for row, k in zip(out, kron_ids):
            with open(os.path.join(dest_folder, f"kron_{sub}_{k}.txt"), "w") as f:
                f.write(f"num_active {n_active}\n")
                f.write(_fmt(row))
    with open(os.path.join(dest_folder, f"rank_{rank}_summary.txt"), "w") as f:
        f.write("Total merge time = %e\n" % (time.time() - t0))
        f.write("DONE")


def build(build_file: str, data_folder: str, dest_folder: str, rank: int, recursion_layer: int = 0,
          device=None):
    """build.c:69-196: reconstructed_prob_%d.txt + appended summary."""
    t0 = time.time()
    dev = kernels._dev(device)
    with open(build_file) as f:
        lines = [ln for ln in f.read().split("\n") if ln.strip()]
    head = dict(kv.split("=") for kv in lines[0].split())
    total_active, n_sub = int(head["total_active_qubit"]), int(head["num_subcircuits"])
    n_terms, n_cuts = int(head["num_summation_terms"]), int(head["num_cuts"])
    terms = [[tuple(map(int, tok.split(","))) for tok in ln.split()] for ln in lines[1:1 + n_terms]]
    if n_terms == 0:
        vec = np.zeros(1 << total_active)
    else:
        order = [s for s, _ in terms[0]]                   # subcircuit order of the file (smart_order)
        kron, index = [], []
        for pos, sub in enumerate(order):
            rows, row_of, idx = [], {}, []
            for term in terms:
                s, k = term[pos]
                assert s == sub
                if k not in row_of:
                    path = os.path.join(data_folder, f"kron_{sub}_{k}.txt")
                    if os.path.exists(path):                # a missing file zeroes the term (:35-67)
                        row_of[k] = len(rows)
                        rows.append(_read_kron(path, "num_active")[1])
                    else:
                        row_of[k] = -1
                idx.append(row_of[k])
            if not rows:
                rows = [np.zeros(1)]
            kron.append(torch.as_tensor(np.stack(rows), device=dev))
            index.append(idx)
        vec = kernels.build(kron, index, n_cuts).cpu().numpy()
        if vec.size != 1 << total_active:                  # every term zero somewhere
            vec = np.zeros(1 << total_active)
    with open(os.path.join(dest_folder, f"reconstructed_prob_{rank}.txt"), "w") as f:
        f.write(_fmt(vec))
    with open(os.path.join(dest_folder, f"rank_{rank}_summary.txt"), "a") as f:
        f.write("\nTotal build time = %e\n" % (time.time() - t0))
        f.write("DONE")


def main(stage: str, argv: List[str]):
    """argv exactly as the reference's C programs take it."""
    if stage == "measure":
        n = int(argv[4])
        measure(int(argv[0]), argv[1], int(argv[2]), int(argv[3]), [int(a) for a in argv[5:5 + n]])
    elif stage == "vertical_collapse":
        vertical_collapse(int(argv[0]), argv[1], argv[2], argv[3], int(argv[4]), int(argv[5]))
    elif stage == "merge":
        merge(argv[0], argv[1], argv[2], int(argv[3]), int(argv[4]))
    elif stage == "build":
        build(argv[0], argv[1], argv[2], int(argv[3]), int(argv[4]))
    else:
        raise ValueError(stage)
