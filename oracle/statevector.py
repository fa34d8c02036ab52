"""ORACLE (test infrastructure, not product code): dense CPU statevector simulator.

Stands in for qiskit-aer 0.7.2 `statevector_simulator` (cutting_requirements.txt:89), which the
reference calls through `qiskit.execute` (qsplit/qsplit_mlrecon_methods.py:25; mis.py:649) and
which is NOT vendored under /root/reference and NOT installable here.  Gate matrices restate
qiskit-terra 0.16.1's published definitions:

  U3(t,p,l) = [[cos t/2, -e^{il} sin t/2], [e^{ip} sin t/2, e^{i(p+l)} cos t/2]]
  RX(t) = [[cos t/2, -i sin t/2], [-i sin t/2, cos t/2]],  RZ(t) = diag(e^{-it/2}, e^{it/2})
  multi-qubit matrices little-endian (first qarg = least significant bit), integer basis index
  x has bit i <-> qubit i.

parity unpinned at the Aer boundary itself (no Aer here, no golden amplitudes in the reference);
pinned only through the known answers listed in tests/test_oracle_golden.py.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline/reference legs may import this.
"""
from __future__ import annotations

import numpy as np

SQ2 = np.sqrt(2.0)


def u3_matrix(theta, phi, lam):
    c, s = np.cos(theta / 2), np.sin(theta / 2)
    return np.array([[c, -np.exp(1j * lam) * s],
                     [np.exp(1j * phi) * s, np.exp(1j * (phi + lam)) * c]], dtype=np.complex128)


def gate_matrix(name, params=()):
    """2x2 matrix of a named single-qubit gate."""
    if name == "x":
        return np.array([[0, 1], [1, 0]], dtype=np.complex128)
    if name == "y":
        return np.array([[0, -1j], [1j, 0]], dtype=np.complex128)
    if name == "z":
        return np.diag([1, -1]).astype(np.complex128)
    if name == "h":
        return np.array([[1, 1], [1, -1]], dtype=np.complex128) / SQ2
    if name == "s":
        return np.diag([1, 1j]).astype(np.complex128)
    if name == "sdg":
        return np.diag([1, -1j]).astype(np.complex128)
    if name == "t":
        return np.diag([1, np.exp(1j * np.pi / 4)]).astype(np.complex128)
    if name == "tdg":
        return np.diag([1, np.exp(-1j * np.pi / 4)]).astype(np.complex128)
    if name == "id":
        return np.eye(2, dtype=np.complex128)
    if name == "rx":
        (t,) = params
        c, s = np.cos(t / 2), np.sin(t / 2)
        return np.array([[c, -1j * s], [-1j * s, c]], dtype=np.complex128)
    if name == "ry":
        (t,) = params
        c, s = np.cos(t / 2), np.sin(t / 2)
        return np.array([[c, -s], [s, c]], dtype=np.complex128)
    if name == "rz":
        (t,) = params
        return np.diag([np.exp(-0.5j * t), np.exp(0.5j * t)]).astype(np.complex128)
    if name == "u1":
        (l,) = params
        return np.diag([1, np.exp(1j * l)]).astype(np.complex128)
    if name == "u3":
        return u3_matrix(*params)
    raise ValueError(f"unknown 1q gate {name}")


def apply_controlled_1q(state, n, target, mat, ctrl_mask=0, ctrl_val=0):
    """state[x] <- (mat on bit `target`) for all x with (x & ctrl_mask) == ctrl_val.
    The state is viewed as [bits above target][target bit][bits below target]; the control test
    separates into one over the high part and one over the low part, so no 2^n index array is
    built (a 25-qubit circuit stays a seconds-long check)."""
    assert not (ctrl_mask >> target) & 1
    lo = 1 << target
    v = state.reshape(-1, 2, lo)
    if not ctrl_mask:
        a0 = v[:, 0, :].copy()
        a1 = v[:, 1, :].copy()
        v[:, 0, :] = mat[0, 0] * a0 + mat[0, 1] * a1
        v[:, 1, :] = mat[1, 0] * a0 + mat[1, 1] * a1
        return state
    hi_m = ctrl_mask & ~((lo << 1) - 1)
    lo_m = ctrl_mask & (lo - 1)
    rows = np.arange(v.shape[0], dtype=np.int64) << (target + 1)
    rows = np.nonzero((rows & hi_m) == (ctrl_val & hi_m))[0]
    if lo_m:
        cols = np.arange(lo, dtype=np.int64)
        cols = np.nonzero((cols & lo_m) == (ctrl_val & lo_m))[0]
        i0, i1 = np.ix_(rows, [0], cols), np.ix_(rows, [1], cols)
    else:
        i0, i1 = (rows, 0), (rows, 1)
    a0 = v[i0].copy() if lo_m else v[rows, 0, :].copy()
    a1 = v[i1].copy() if lo_m else v[rows, 1, :].copy()
    v[i0] = mat[0, 0] * a0 + mat[0, 1] * a1
    v[i1] = mat[1, 0] * a0 + mat[1, 1] * a1
    return state


def apply_unitary(state, n, qubits, mat):
    """Dense k-qubit unitary, little-endian in `qubits` (first = least significant)."""
    k = len(qubits)
    # tensor axes: numpy axis j <-> qubit n-1-j
    psi = state.reshape((2,) * n)
    axes = [n - 1 - q for q in qubits]          # axis of each gate qubit
    # gate tensor: row index bits (msb..lsb) = qubits[k-1]..qubits[0]
    g = mat.reshape((2,) * (2 * k))
    in_axes = list(range(k, 2 * k))             # gate column axes msb..lsb
    psi_axes = [axes[k - 1 - j] for j in range(k)]   # msb..lsb
    out = np.tensordot(g, psi, axes=(in_axes, psi_axes))
    # out axes: gate row axes (msb..lsb) first, then remaining psi axes in order
    rest = [a for a in range(n) if a not in psi_axes]
    cur = psi_axes + rest
    perm = [cur.index(a) for a in range(n)]
    return np.ascontiguousarray(out.transpose(perm)).reshape(-1)


def apply_instruction(state, n, inst):
    name, qs, params, matrix = inst
    if name in ("barrier", "Id", "id"):
        return state
    if name == "unitary":
        return apply_unitary(state, n, list(qs), np.asarray(matrix, dtype=np.complex128))
    if name == "cx":
        return apply_controlled_1q(state, n, qs[1], gate_matrix("x"), 1 << qs[0], 1 << qs[0])
    if name == "cz":
        return apply_controlled_1q(state, n, qs[1], gate_matrix("z"), 1 << qs[0], 1 << qs[0])
    if name == "mcx":
        mask = 0
        for c in qs[:-1]:
            mask |= 1 << c
        return apply_controlled_1q(state, n, qs[-1], gate_matrix("x"), mask, mask)
    if name == "crx":
        return apply_controlled_1q(state, n, qs[1], gate_matrix("rx", params),
                                   1 << qs[0], 1 << qs[0])
    if name == "zcrx":
        mask = 0
        for c in qs[:-1]:
            mask |= 1 << c
        return apply_controlled_1q(state, n, qs[-1], gate_matrix("rx", params), mask, 0)
    return apply_controlled_1q(state, n, qs[0], gate_matrix(name, params))


def simulate(n, instructions, init=None):
    """Run a list of plain-tuple instructions on |0..0> (or `init`) and return amplitudes."""
    if init is None:
        state = np.zeros(1 << n, dtype=np.complex128)
        state[0] = 1.0
    else:
        state = np.array(init, dtype=np.complex128).copy()
    for inst in instructions:
        state = apply_instruction(state, n, inst)
    return state


def probabilities(state):
    return state.real ** 2 + state.imag ** 2


def probabilities_dict(state, n, drop_zero=True):
    """{qiskit bitstring: prob}; char j of the key <-> qubit n-1-j."""
    p = probabilities(state)
    out = {}
    for x in range(1 << n):
        if drop_zero and p[x] == 0.0:
            continue
        out[format(x, f"0{n}b")] = float(p[x])
    return out
