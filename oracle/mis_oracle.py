"""ORACLE (test infrastructure, not product code): CPU restatement of the MIS scoring that
`solve_mis_cut_dqva` applies to the recombined distribution.

Follows, line by line:
  strip_ancillas      utils/helper_funcs.py:3-16
  hamming_weight      utils/helper_funcs.py:18-19
  is_indset           utils/graph_funcs.py:107-118
  avg_cost            mis.py:117-134   (sum of prob*weight over independent sets)
  top_probs/better    mis.py:232-242
  uncut cost f        mis.py:657-661   (no independent-set mask)
  QAOA+ expectation   ansatz/qaoa_plus.py:40-53
  brute force         utils/helper_funcs.py:46-62

Pinned by the reference's own golden files benchmark_graphs/brute_force_outputs/*/G*_brute_force.out
(copies under tests/golden/brute_force/), which fix `is_indset`, `hamming_weight` and the
bitstring endianness on real graphs.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/reference legs may import this.
"""
from __future__ import annotations

import numpy as np


# ---- dict / string forms (line-faithful) -------------------------------------------------

def strip_ancillas(counts, num_anc):                   # helper_funcs.py:3-16
    if not num_anc:
        raise Exception('A valid qiskit circuit or number of ancillas must be provided')
    new_counts = {}
    for key in counts:
        new_key = key[num_anc:]
        if new_key in new_counts.keys():
            new_counts[new_key] += counts[key]
        else:
            new_counts[new_key] = counts[key]
    return new_counts


def hamming_weight(bitstr):                            # helper_funcs.py:18-19
    return sum([1 for bit in bitstr if bit == '1'])


def is_indset(bitstr, G):                              # graph_funcs.py:107-118
    ind_set = []
    for idx, bit in enumerate(reversed(bitstr)):
        if bit == '1':
            cur_neighbors = list(G.neighbors(idx))
            for node in ind_set:
                if node in cur_neighbors:
                    return False
            else:
                ind_set.append(idx)
    return True


def avg_cost(probs, graph):                            # mis.py:125-134
    avg_weight = sum([prob * hamming_weight(bitstr) for bitstr, prob
                      in probs.items() if is_indset(bitstr, graph)])
    return -avg_weight


def top_probs(probs, threshold=1e-5, cutoff=1):        # mis.py:232-233
    return sorted([(key, val) for key, val in probs.items() if val > threshold],
                  key=lambda tup: tup[1], reverse=True)[:cutoff]


def better_strs(top, graph, best_indset):              # mis.py:236-242
    best_hamming_weight = hamming_weight(best_indset)
    out = []
    for bitstr, prob in top:
        this_hamming = hamming_weight(bitstr)
        if is_indset(bitstr, graph) and this_hamming > best_hamming_weight:
            out.append((bitstr, this_hamming))
    return sorted(out, key=lambda t: t[1], reverse=True)


def uncut_cost(probs):                                 # mis.py:657-661
    avg = 0
    for sample in probs.keys():
        x = [int(bit) for bit in list(sample)]
        avg += probs[sample] * sum(x)
    return -avg


def qaoa_plus_expectation(counts, G, Lambda):          # qaoa_plus.py:40-53
    total_shots = sum(counts.values())
    energy = 0
    for bitstr, count in counts.items():
        temp_energy = hamming_weight(bitstr)
        for edge in G.edges():
            q_i, q_j = edge
            rev_bitstr = list(reversed(bitstr))
            if rev_bitstr[q_i] == '1' and rev_bitstr[q_j] == '1':
                temp_energy += -1 * Lambda
        energy += count * temp_energy / total_shots
    return energy


def brute_force_search_memory_efficient(G):            # helper_funcs.py:46-62
    num_nodes = len((list(G.nodes)))
    best_hamming_weight = 0
    best_strs = []
    for counter in range(int(2 ** num_nodes)):
        bitstr = f'{counter:0{num_nodes}b}'
        if hamming_weight(bitstr) >= best_hamming_weight and is_indset(bitstr, G):
            best_hamming_weight = hamming_weight(bitstr)
            if len(best_strs) > 0 and best_hamming_weight > hamming_weight(best_strs[0]):
                best_strs = [bitstr]
            else:
                best_strs.append(bitstr)
    return best_strs, best_hamming_weight


# ---- array forms (same semantics on integer indices; bit i <-> node i) --------------------

def adjacency_masks(G, n):
    """adj[i] = OR of 1<<j over neighbours j of node i (uint64)."""
    adj = np.zeros(n, dtype=np.uint64)
    for i in range(n):
        m = 0
        if G.has_node(i):
            for j in G.neighbors(i):
                m |= 1 << int(j)
        adj[i] = m
    return adj


def popcount64(x):
    x = x.astype(np.uint64)
    m1, m2, m4 = np.uint64(0x5555555555555555), np.uint64(0x3333333333333333), \
        np.uint64(0x0F0F0F0F0F0F0F0F)
    x = x - ((x >> np.uint64(1)) & m1)
    x = (x & m2) + ((x >> np.uint64(2)) & m2)
    x = (x + (x >> np.uint64(4))) & m4
    return ((x * np.uint64(0x0101010101010101)) >> np.uint64(56)).astype(np.int64)


def indset_mask(xs, adj):
    """ok[x] = all set bits i of x have (x & adj[i]) == 0."""
    xs = xs.astype(np.uint64)
    viol = np.zeros(xs.shape, dtype=bool)
    for i, a in enumerate(adj):
        bit = (xs >> np.uint64(i)) & np.uint64(1)
        viol |= (bit == 1) & ((xs & a) != 0)
    return ~viol


def score_array(probs, n, num_anc, adj, masked=True, lo=0):
    """(sum p, sum p*weight*[indset]) over a slab probs[x - lo] of the (n+num_anc)-bit
    distribution, ancillas = top `num_anc` bits, exactly as strip_ancillas+avg_cost do.
    float64 accumulation with math.fsum-quality pairwise numpy sums."""
    xs = np.arange(lo, lo + probs.shape[0], dtype=np.uint64)
    node_bits = xs & np.uint64((1 << n) - 1)
    # strip_ancillas sums duplicates BEFORE the indset filter; the filter only looks at node
    # bits, so filtering per entry and summing afterwards is the same number.
    w = popcount64(node_bits)
    ok = indset_mask(node_bits, adj) if masked else np.ones(xs.shape, dtype=bool)
    return float(np.sum(probs)), float(np.sum(probs * w * ok))


def topk_array(probs, n, num_anc, k, threshold=1e-5, lo=0):
    """Top-k (node bitstring int, prob) after marginalising ancillas; ties broken by smaller
    integer (documented deviation: the reference's tie order is dict-insertion order)."""
    size = probs.shape[0]
    assert lo == 0 and size == 1 << (n + num_anc)
    marg = probs.reshape(1 << num_anc, 1 << n).sum(axis=0)
    idx = np.nonzero(marg > threshold)[0]
    order = np.lexsort((idx, -marg[idx]))[:k]
    return [(int(idx[i]), float(marg[idx[i]])) for i in order]
