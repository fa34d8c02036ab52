"""Names and readers of the reference's on-disk CutQC state (cutqc/helper_fun.py:12-26; read_dict is
qiskit_helper_functions.non_ibmq_functions.read_dict, which the reference imports everywhere).

This package keeps the recursion state in memory; `CutQC.post_process(dest_folder=...)` mirrors it
to disk in the reference's layout so that tools written against `./cutqc_data/...` keep working:

    <dest_folder>/meta_data.pckl                         {'counter', 'dynamic_definition_schedule'}
    <dest_folder>/dynamic_definition_<layer>/build_output.pckl
                                                         {'zoomed_ctr', 'max_states', 'reconstructed_prob'}
(cutqc/distributor.py:121-141, 212-215; cutqc/main.py:452-457).
"""
from __future__ import annotations

import os
import pickle

__all__ = ["get_dirname", "read_dict"]

_FIELDS = {
    "cutter": ("cutqc_data", "{name}/cc_{msq}"),
    "evaluator": ("cutqc_data", "{name}/cc_{msq}/{mode}"),
    "vertical_collapse": ("cutqc_data", "{name}/cc_{msq}/{mode}/et_{et}"),
    "build": ("cutqc_data", "{name}/cc_{msq}/{mode}_{et}_{limit}_{threads}"),
    "slurm": ("slurm", "{name}/cc_{msq}"),
    "runtime": ("runtime", "{name}/cc_{msq}/q_{limit}_{threads}"),
}


def get_dirname(circuit_name, max_subcircuit_qubit, early_termination, eval_mode, num_threads,
                qubit_limit, field):
    """Directory of one pipeline stage, spelled as the reference spells it."""
    if field not in _FIELDS:
        raise Exception("Illegal field = %s" % field)
    root, tail = _FIELDS[field]
    ints = {"msq": max_subcircuit_qubit, "et": early_termination, "limit": qubit_limit, "threads": num_threads}
    used = {k: int(v) for k, v in ints.items() if "{" + k + "}" in tail}
    return "./%s/%s" % (root, tail.format(name=circuit_name, mode=eval_mode, **used))


def read_dict(filename):
    """All pickled dicts appended to `filename`, merged (the reference appends to summary.pckl);
    {} when the file does not exist."""
    out = {}
    if not os.path.isfile(filename):
        return out
    with open(filename, "rb") as f:
        while True:
            try:
                out.update(pickle.load(f))
            except EOFError:
                break
    return out
