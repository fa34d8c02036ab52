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
import importlib as _importlib
import importlib.abc as _abc
import importlib.util as _util
import sys as _sys

_REAL, _ALIAS = __name__, "qdc_b200"
_sys.modules[_ALIAS] = _sys.modules[__name__]


class _AliasFinder(_abc.MetaPathFinder, _abc.Loader):
    """`qdc_b200.x.y` IS `<this package>.x.y` (one module object under two names), so classes
    such as B200Backend or QdcError are identical whichever name a caller imported them by."""

    def find_spec(self, fullname, path=None, target=None):
        if fullname.startswith(_ALIAS + "."):
            return _util.spec_from_loader(fullname, self, origin=_REAL + fullname[len(_ALIAS):])
        return None

    _real_specs = {}

    def create_module(self, spec):
        module = _importlib.import_module(spec.origin)
        self._real_specs[spec.origin] = module.__spec__
        return module

    def exec_module(self, module):
        # the import machinery has just pointed __spec__ at the alias; put the real one back so
        # that relative imports inside the module keep resolving against its real package
        real = self._real_specs.get(getattr(module.__spec__, "origin", None))
        if real is not None:
            module.__spec__ = real


if not any(type(f).__name__ == "_AliasFinder" for f in _sys.meta_path):
    _sys.meta_path.insert(0, _AliasFinder())

__version__ = "0.1.0"
