"""qdc_b200 -- B200-native Quantum Divide-and-Conquer hot path.

Drop-in for the hot path of teaguetomesh/dqva-and-circuit-cutting: fragment-variant
simulation -> 4^K recombination -> MIS scoring, behind the reference's own Python names
(`qsplit.qsplit_mlrecon_methods`, `qsplit.qsplit_circuit_cutter`, `utils.cutting_funcs`,
`cutqc`), computed by hand-written sm_100a CUDA kernels in `csrc/` through the C-ABI declared
in `include/qdc_b200.h`.  There is no CPU fallback: every compute entry point raises if
`libqdc_b200.so` cannot be loaded.

The directory name contains '-' (it mirrors the reference repository's name), so the package
registers itself under the importable alias `qdc_b200`.
"""
import sys as _sys

_sys.modules["qdc_b200"] = _sys.modules[__name__]

__version__ = "0.1.0"
