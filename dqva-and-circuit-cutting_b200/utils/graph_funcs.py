"""Graph helpers with the reference's names and semantics (utils/graph_funcs.py).

Host-side only.  Node id == qubit index throughout (utils/graph_funcs.py:107-118).
"""
import networkx as nx

__all__ = ["graph_from_file", "get_subgraphs", "is_indset", "adjacency_masks"]


def graph_from_file(fn):
    """Parse the benchmark edge-list format `(a, b), (c, d), ...` on the first line
    (utils/graph_funcs.py:3-17).  Nodes are added in order of first appearance, exactly as
    `nx.Graph.add_edges_from` on the parsed list does in the reference."""
    with open(fn, "r") as f:
        line = f.readline()
    edges, cur = [], None
    for tok in line.split(","):
        if "(" in tok:
            cur = [int(tok.strip("( \n"))]
        elif ")" in tok:
            cur.append(int(tok.strip(") \n")))
            edges.append(tuple(cur))
    G = nx.Graph()
    G.add_edges_from(edges)
    return G


def get_subgraphs(G, partition):
    """Induced subgraph per part plus the list of crossing edges, each crossing edge listed
    once per endpoint-side it leaves (utils/graph_funcs.py:89-105)."""
    subgraphs, cut_edges = [], []
    for part in partition:
        sub = nx.Graph()
        sub.add_nodes_from(part)
        for v1, v2 in G.edges:
            in1, in2 = v1 in part, v2 in part
            if in1 and in2:
                sub.add_edge(v1, v2)
            if in1 and not in2:
                cut_edges.append((v1, v2))
        subgraphs.append(sub)
    return subgraphs, cut_edges


def is_indset(bitstr, G):
    """True iff the '1' positions of the little-endian string form an independent set
    (utils/graph_funcs.py:107-118)."""
    chosen = []
    for idx, bit in enumerate(reversed(bitstr)):
        if bit != "1":
            continue
        nbrs = set(G.neighbors(idx))
        if any(node in nbrs for node in chosen):
            return False
        chosen.append(idx)
    return True


def adjacency_masks(G, n):
    """adj[i] = bit mask of the neighbours of node i; the device form of `is_indset`:
    x is independent  <=>  for every set bit i of x, (x & adj[i]) == 0."""
    masks = []
    for i in range(n):
        m = 0
        if G.has_node(i):
            for j in G.neighbors(i):
                m |= 1 << int(j)
        masks.append(m)
    return masks
