"""solve_mis_cut_dqva end to end on the device path: the returned
bitstring is an independent set at least as heavy as the start, history records have the
reference's keys (mis.py:245-253), and the batched SPSA optimiser uses population evaluations."""
import random

import networkx as nx
import numpy as np
import pytest

from oracle import mis_oracle as mo

pytestmark = pytest.mark.gpu

KEYS = {"mixer_round", "inner_round", "cost", "function_evals", "init_state", "mixer_order",
        "num_params", "frag_shots", "frag_qubits", "cuts", "hot_nodes", "better_strs",
        "cur_params", "cur_args"}


def _graph():
    G = nx.Graph()
    G.add_edges_from([(0, 1), (1, 2), (2, 3), (3, 0), (0, 2), (4, 5), (5, 6), (6, 7), (7, 4), (5, 7), (3, 4)])
    return G


@pytest.mark.parametrize("optimizer,options", [("COBYLA", {"maxiter": 40}),
                                               ("SPSA", {"iterations": 6, "pairs": 8, "seed": 0})])
def test_solver_returns_valid_improving_independent_set(cuda, optimizer, options):
    from qdc_b200.engine import B200Backend
    from qdc_b200.solvers import solve_mis_cut_dqva
    G = _graph()
    random.seed(3)
    np.random.seed(3)
    out = solve_mis_cut_dqva("0" * 8, G, m=2, max_cuts=2, cutoff=2, optimizer=optimizer,
                             optimizer_options=options, backend=B200Backend(device=cuda),
                             partition=[[0, 1, 2, 3], [4, 5, 6, 7]])
    best, params, init_state, perm, partition, cut_nodes, history = out
    assert len(best) == 8 and mo.is_indset(best, G)
    assert mo.hamming_weight(best) >= 1                      # something better than the empty set
    assert sorted(perm) == list(range(8)) and len(history) == 2
    for rounds in history:
        for rec in rounds:
            assert set(rec) == KEYS
            assert rec["cost"] <= 1e-9                       # -sum p |x| [indset] is never positive
            for s, w in rec["better_strs"]:
                assert mo.is_indset(s, G) and mo.hamming_weight(s) == w
    if optimizer == "SPSA":
        assert history[0][0]["function_evals"] == 6 * 17 + 1
    with pytest.raises(ValueError):
        solve_mis_cut_dqva("0" * 8, G, max_cuts=0, num_frags=2, backend=B200Backend(device=cuda))
