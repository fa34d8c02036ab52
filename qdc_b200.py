"""Import shim: `import qdc_b200` from the repo root loads `dqva-and-circuit-cutting_b200/`
(whose name is not a Python identifier) and aliases it."""
import importlib as _importlib
import os as _os
import sys as _sys

_root = _os.path.dirname(_os.path.abspath(__file__))
if _root not in _sys.path:
    _sys.path.insert(0, _root)
_importlib.import_module("dqva-and-circuit-cutting_b200")   # re-points sys.modules["qdc_b200"]
