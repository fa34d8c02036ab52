"""String/dict helpers with the reference's names (utils/helper_funcs.py).  Host-side."""
from .graph_funcs import is_indset

__all__ = ["strip_ancillas", "hamming_weight", "brute_force_search",
           "brute_force_search_memory_efficient"]


def strip_ancillas(counts, circ=None, num_anc=None):
    """Drop the first `num_anc` characters (highest wires = ancillas) of every key and add
    up collisions (utils/helper_funcs.py:3-16)."""
    if circ is not None:
        num_anc = len(circ.ancillas)
    elif not num_anc:
        raise Exception('A valid qiskit circuit or number of ancillas must be provided')
    out = {}
    for key, val in counts.items():
        short = key[num_anc:]
        out[short] = out.get(short, 0) + val if short in out else val
    return out


def hamming_weight(bitstr):
    return bitstr.count("1")


def brute_force_search(G, lim=None):
    n = len(list(G.nodes))
    best, strs = 0, []
    for x in range(1 << n):
        b = format(x, f"0{n}b")
        w = hamming_weight(b)
        if lim is not None and w < lim:
            continue
        if w >= best and is_indset(b, G):
            if w > best:
                best, strs = w, []
            strs.append(b)
    return strs, best


def brute_force_search_memory_efficient(G):
    return brute_force_search(G)
