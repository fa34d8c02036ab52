"""Ad-hoc device timing of the statevector engine on the C4 / C2 fragment shapes."""
import importlib, os, sys
import numpy as np, torch, networkx as nx
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
importlib.import_module("dqva-and-circuit-cutting_b200")
from qdc_b200.engine import B200Backend
from qdc_b200.circuit import Parameter


def dqva_fragment(n_nodes, seed, symbolic=True):
    """DQVA-structured fragment: one open-controlled RX per node (controls = its neighbours in a
    random 3-regular graph), then the RZ phase separator; last qubit = ancilla (idle when fused)."""
    G = nx.random_regular_graph(3, n_nodes, seed=seed)
    rng = np.random.default_rng(seed)
    pars = [Parameter(f"var_{i}") for i in range(n_nodes + 1)]
    insts = []
    for q in rng.permutation(n_nodes):
        q = int(q)
        insts.append(("zcrx", tuple(int(v) for v in G.neighbors(q)) + (q,), (2 * pars[q],), None))
    for q in range(n_nodes):
        insts.append(("rz", (q,), (2 * pars[n_nodes],), None))
    return n_nodes + 1, insts, pars


def ev_time(fn, reps=5, warm=2):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e-3)
    return min(ts)


def case(name, n_nodes, prep, meas, n_sets, tile_bits, coalesce, basis=None):
    from qdc_b200.engine import BASIS_KETS, SIC_KETS
    if basis is None:
        basis = bool(prep) and len(prep) + len(meas) <= 4
    dev = torch.device("cuda:0")
    be = B200Backend(device=dev, tile_bits=tile_bits, coalesce_bits=coalesce)
    n, insts, pars = dqva_fragment(n_nodes, 7)
    zero = ((1 << n) - 1) & ~sum(1 << q for q in prep)
    prog, auto = be.compile_instructions(n, insts, pars, zero)
    table = np.random.default_rng(1).uniform(0, 2 * np.pi, (n_sets, len(pars)))
    n_init = (2 if basis else 4) ** len(prep)
    kets = BASIS_KETS if basis else SIC_KETS
    states = torch.empty(n_init * n_sets, 1 << n, dtype=torch.complex128, device=dev)
    t_run = ev_time(lambda: be.run(prog, auto, table, prep_qubits=prep, kets=kets, out=states))
    T = be.pauli_tensor(states, prog, n_sets, len(prep), meas, basis_inputs=basis)
    t_ten = ev_time(lambda: be.pauli_tensor(states, prog, n_sets, len(prep), meas, out=T, basis_inputs=basis))
    sims = n_init * n_sets
    ref_variants = 4 ** len(prep) * n_sets * 3 ** len(meas)
    name = name + (" [basis inputs]" if basis else "")
    gbytes = prog.n_groups * 32.0 * (1 << n) * sims / 1e9
    print(f"{name}: n={n} p={len(prep)} m={len(meas)} sets={n_sets} t={prog.tile_bits} c={coalesce} "
          f"groups={prog.n_groups} ops={prog.n_ops} | run {t_run*1e3:.3f} ms ({sims/t_run:.3e} states/s, "
          f"{gbytes/t_run:.0f} GB/s group traffic) | tensor {t_ten*1e3:.3f} ms "
          f"({(states.numel()*16+T.numel()*8)/t_ten/1e9:.0f} GB/s) | ref-equivalent "
          f"{ref_variants/(t_run+t_ten):.3e} variants/s", flush=True)


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "sweep":
        for t, c in [(12, 4), (11, 4), (10, 4), (9, 4), (11, 3), (10, 3), (13, 4)]:
            case("C4 prep side", 16, [0, 5, 9, 13], [], 1, t, c)
        for t, c in [(12, 4), (11, 4), (10, 4), (9, 4), (11, 3), (10, 3), (13, 4)]:
            case("C4 meas side", 20, [], [3, 8, 12, 17], 1, t, c)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "one":
        case("C4 prep side", 16, [0, 5, 9, 13], [], 1, 12, 4, basis=False)
        case("C4 prep side", 16, [0, 5, 9, 13], [], 1, 12, 4)
        case("C4 meas side", 20, [], [3, 8, 12, 17], 1, 12, 4)
        case("C2 like", 14, [2, 9], [5], 64, 12, 4)
        case("C2 prep", 14, [2, 9, 11], [], 64, 12, 4)
        case("C2 meas", 14, [], [5, 8, 12], 64, 12, 4)
        sys.exit(0)
    for t, c in [(12, 4), (13, 4), (11, 4), (12, 5), (12, 3), (10, 4)]:
        case("C4 prep side", 16, [0, 5, 9, 13], [], 1, t, c)
    for t, c in [(12, 4), (13, 4), (11, 4), (12, 5)]:
        case("C4 meas side", 20, [], [3, 8, 12, 17], 1, t, c)
    for t, c in [(12, 4), (13, 4), (11, 4)]:
        case("C2 like", 14, [2, 9], [5], 64, t, c)
    case("C1 like", 10, [4], [], 64, 0, -1)
    case("C1 like meas", 10, [], [4], 64, 0, -1)
