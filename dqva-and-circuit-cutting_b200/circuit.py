"""Qiskit-free circuit container for the QDC hot path.

The reference builds its circuits with qiskit (`QuantumCircuit`, `AncillaRegister`,
`Parameter`; ansatz/dqv_cut_ansatz.py:252-356) and hands `QuantumCircuit` fragments plus a
`wire_path_map` keyed by qiskit `Qubit` objects to the hot path
(qsplit/qsplit_circuit_cutter.py:86-246, qsplit/qsplit_mlrecon_methods.py:216-232).  qiskit is
not available in this image, so this module provides the small subset of that object model the
hot path touches:

* `Qubit` with `.register` / `.index` (the recombination code sorts wires by `.index`,
  qsplit_mlrecon_methods.py:127,527) and value-based hashing,
* `QuantumCircuit` with `.qubits`, `.ancillas`, `.num_qubits`, `.parameters`,
  `.bind_parameters`, the gate methods the ansatz modules call, and `.data`
  (list of `Instruction`),
* `Parameter` / `ParameterExpression` restricted to affine expressions `scale*param + offset`
  (all the reference ever forms is `2*alpha`, `2*gamma`, `gamma/2`, `-gamma/2`, `-2*beta`).

Nothing here does arithmetic on amplitudes: gate application lives in the CUDA engine
(`csrc/statevec.cu`); this is host-side bookkeeping only.
"""
from __future__ import annotations

import itertools
import numbers
from typing import Iterable, List, Optional, Sequence

import numpy as np

__all__ = ["Register", "QuantumRegister", "AncillaRegister", "Qubit", "Parameter",
           "ParameterExpression", "Instruction", "QuantumCircuit"]


class Register:
    """Named group of wires.  Mirrors qiskit's QuantumRegister/AncillaRegister."""
    kind = "q"
    _counter = itertools.count()

    def __init__(self, size: int, name: Optional[str] = None):
        self.size = int(size)
        self.name = name if name is not None else f"{self.kind}{next(Register._counter)}"
        self.prefix = self.name
        self._bits = [Qubit(self, i) for i in range(self.size)]

    def __len__(self):
        return self.size

    def __iter__(self):
        return iter(self._bits)

    def __getitem__(self, key):
        return self._bits[key]

    def __eq__(self, other):
        return (isinstance(other, Register) and self.kind == other.kind
                and self.name == other.name and self.size == other.size)

    def __hash__(self):
        return hash((self.kind, self.name, self.size))

    def __repr__(self):
        return f"{type(self).__name__}({self.size}, '{self.name}')"


class QuantumRegister(Register):
    kind = "q"


class AncillaRegister(Register):
    kind = "anc"


class Qubit:
    """A wire: (register, index).  Hashable by value like qiskit's Bit."""
    __slots__ = ("register", "index")

    def __init__(self, register: Register, index: int):
        self.register = register
        self.index = int(index)

    def __eq__(self, other):
        return (isinstance(other, Qubit) and self.index == other.index
                and self.register == other.register)

    def __hash__(self):
        return hash((self.register, self.index))

    def __repr__(self):
        return f"Qubit({self.register.name}[{self.index}])"


class ParameterExpression:
    """Affine function of at most one Parameter: value = scale * param + offset."""
    __slots__ = ("param", "scale", "offset")

    def __init__(self, param, scale=1.0, offset=0.0):
        self.param = param
        self.scale = float(scale)
        self.offset = float(offset)

    # arithmetic the ansatz modules need (2*alpha, gamma/2, -1*gamma/2, -2*beta)
    def __mul__(self, other):
        if isinstance(other, numbers.Real):
            return ParameterExpression(self.param, self.scale * other, self.offset * other)
        return NotImplemented
    __rmul__ = __mul__

    def __truediv__(self, other):
        if isinstance(other, numbers.Real):
            return ParameterExpression(self.param, self.scale / other, self.offset / other)
        return NotImplemented

    def __neg__(self):
        return ParameterExpression(self.param, -self.scale, -self.offset)

    def __add__(self, other):
        if isinstance(other, numbers.Real):
            return ParameterExpression(self.param, self.scale, self.offset + other)
        return NotImplemented
    __radd__ = __add__

    def __sub__(self, other):
        if isinstance(other, numbers.Real):
            return ParameterExpression(self.param, self.scale, self.offset - other)
        return NotImplemented

    @property
    def parameters(self):
        return {self.param}

    def bind(self, value: float) -> float:
        return self.scale * float(value) + self.offset

    def __repr__(self):
        return f"{self.scale}*{self.param.name}+{self.offset}"


class Parameter(ParameterExpression):
    """Symbolic circuit parameter (qiskit.circuit.Parameter stand-in; mis.py:64)."""
    __slots__ = ("name",)

    def __init__(self, name: str):
        self.name = str(name)
        super().__init__(self, 1.0, 0.0)

    def __eq__(self, other):
        return isinstance(other, Parameter) and other.name == self.name

    def __hash__(self):
        return hash(("Parameter", self.name))

    def __repr__(self):
        return f"Parameter({self.name})"


def _param_sort_key(p: Parameter):
    digits = "".join(ch for ch in p.name if ch.isdigit())
    return (int(digits) if digits else -1, p.name)


class Instruction:
    """One operation: gate name, wires it acts on, and its (possibly symbolic) parameters.

    `matrix` is set for name == "unitary" (little-endian, first qarg least significant, as
    qiskit's UnitaryGate).  `num_ctrl` is the number of leading control qargs for "mcx".
    """
    __slots__ = ("name", "qargs", "params", "matrix")

    def __init__(self, name: str, qargs: Sequence[Qubit], params: Sequence = (), matrix=None):
        self.name = name
        self.qargs = list(qargs)
        self.params = list(params)
        self.matrix = matrix

    def copy(self):
        return Instruction(self.name, list(self.qargs), list(self.params), self.matrix)

    def __repr__(self):
        return f"{self.name}({self.params}) @ {[ (q.register.name, q.index) for q in self.qargs ]}"


# gate name -> number of real parameters
_GATE_PARAMS = {"x": 0, "y": 0, "z": 0, "h": 0, "s": 0, "sdg": 0, "t": 0, "tdg": 0, "id": 0,
                "rx": 1, "ry": 1, "rz": 1, "u1": 1, "u3": 3, "cx": 0, "cz": 0, "crx": 1,
                "mcx": 0, "Id": 0, "unitary": 0, "barrier": 0, "zcrx": 1}


class QuantumCircuit:
    """Minimal gate-list circuit with qiskit's calling conventions for the gates the
    reference's ansatz modules use (ansatz/dqv_cut_ansatz.py:127-171, ansatz/qaoa_plus.py:8-37).
    """

    def __init__(self, *regs, name: Optional[str] = None):
        self.name = name
        self.qregs: List[Register] = []
        self.qubits: List[Qubit] = []
        self.ancillas: List[Qubit] = []
        self.clbits: list = []
        self.data: List[Instruction] = []
        for reg in regs:
            if isinstance(reg, numbers.Integral):
                # qiskit: QuantumCircuit(n, name='q') creates QuantumRegister(n, 'q')
                reg = QuantumRegister(int(reg), "q")
            self.add_register(reg)

    # ---- structure -------------------------------------------------------------------
    def add_register(self, reg: Register):
        self.qregs.append(reg)
        self.qubits.extend(reg)
        if isinstance(reg, AncillaRegister):
            self.ancillas.extend(reg)

    @property
    def num_qubits(self) -> int:
        return len(self.qubits)

    def _q(self, q) -> Qubit:
        if isinstance(q, Qubit):
            return q
        return self.qubits[int(q)]

    def _qs(self, qs) -> List[Qubit]:
        if isinstance(qs, (Qubit, numbers.Integral)):
            return [self._q(qs)]
        return [self._q(q) for q in qs]

    def append(self, inst, qargs=None, cargs=None):
        if isinstance(inst, Instruction) and qargs is None:
            self.data.append(inst)
        else:
            new = inst.copy()
            new.qargs = self._qs(qargs)
            self.data.append(new)
        return self

    def copy(self):
        new = QuantumCircuit(name=self.name)
        new.qregs = list(self.qregs)
        new.qubits = list(self.qubits)
        new.ancillas = list(self.ancillas)
        new.data = [inst.copy() for inst in self.data]
        return new

    # ---- parameters ------------------------------------------------------------------
    @property
    def parameters(self) -> List[Parameter]:
        """Unbound parameters, ordered by the integer in their name (`var_3` < `var_10`).

        qiskit 0.16 returned an unordered `set` here (so the reference's positional
        `_bind`, mis.py:101-103, is order-unstable); we fix a deterministic order.
        """
        found = set()
        for inst in self.data:
            for p in inst.params:
                if isinstance(p, ParameterExpression):
                    found.add(p.param)
        return sorted(found, key=_param_sort_key)

    def bind_parameters(self, binding: dict) -> "QuantumCircuit":
        new = self.copy()
        for inst in new.data:
            inst.params = [p.bind(binding[p.param])
                           if isinstance(p, ParameterExpression) and p.param in binding else p
                           for p in inst.params]
        return new

    # ---- gates -----------------------------------------------------------------------
    def _gate(self, name, qargs, params=()):
        self.data.append(Instruction(name, qargs, params))
        return self

    def _bcast(self, name, qubits, params=()):
        for q in self._qs(qubits):
            self._gate(name, [q], params)
        return self

    def x(self, q): return self._bcast("x", q)
    def y(self, q): return self._bcast("y", q)
    def z(self, q): return self._bcast("z", q)
    def h(self, q): return self._bcast("h", q)
    def s(self, q): return self._bcast("s", q)
    def sdg(self, q): return self._bcast("sdg", q)
    def t(self, q): return self._bcast("t", q)
    def tdg(self, q): return self._bcast("tdg", q)
    def id(self, q): return self._bcast("id", q)
    def rx(self, theta, q): return self._bcast("rx", q, [theta])
    def ry(self, theta, q): return self._bcast("ry", q, [theta])
    def rz(self, theta, q): return self._bcast("rz", q, [theta])
    def u1(self, lam, q): return self._bcast("u1", q, [lam])
    def u3(self, theta, phi, lam, q): return self._bcast("u3", q, [theta, phi, lam])

    def cx(self, c, t): return self._gate("cx", [self._q(c), self._q(t)])
    def cz(self, c, t): return self._gate("cz", [self._q(c), self._q(t)])
    def crx(self, theta, c, t): return self._gate("crx", [self._q(c), self._q(t)], [theta])

    def mcx(self, controls, target):
        ctrls = self._qs(controls)
        if len(ctrls) == 0:
            return self.x(target)
        return self._gate("mcx", ctrls + [self._q(target)])

    def zcrx(self, theta, controls, target):
        """Open-controlled RX: RX(theta) on `target` iff every control is |0>.

        This is the *fused* form of the reference's partial mixer
        X^N . MCX(N->anc) . X^N . CRX(2a; anc->q) . X^N . MCX . X^N
        (ansatz/dqv_cut_ansatz.py:127-151) with the ancilla returned to |0>.
        """
        return self._gate("zcrx", self._qs(controls) + [self._q(target)], [theta])

    def identity_block(self, qubits, name="Id"):
        """Multi-qubit identity instruction (`tie_subgraph_qubits_together`,
        ansatz/dqv_cut_ansatz.py:234-249): a no-op that counts as one operation on each
        wire and keeps those wires in one connected component for the cutter."""
        return self._gate(name, self._qs(qubits))

    def unitary(self, matrix, qubits):
        qs = self._qs(qubits)
        mat = np.asarray(matrix, dtype=np.complex128)
        assert mat.shape == (2 ** len(qs), 2 ** len(qs))
        self.data.append(Instruction("unitary", qs, (), mat))
        return self

    def barrier(self, *qubits):
        self.data.append(Instruction("barrier", list(self.qubits)))
        return self

    # ---- export ----------------------------------------------------------------------
    def to_instructions(self) -> list:
        """Plain-tuple form `(name, (local qubit indices), (params), matrix|None)` consumed
        by the engine compiler and (independently) by the CPU oracle."""
        pos = {q: i for i, q in enumerate(self.qubits)}
        out = []
        for inst in self.data:
            if inst.name == "barrier":
                continue
            out.append((inst.name, tuple(pos[q] for q in inst.qargs), tuple(inst.params),
                        inst.matrix))
        return out

    def __repr__(self):
        return f"<QuantumCircuit {self.num_qubits}q, {len(self.data)} ops>"
