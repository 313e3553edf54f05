"""Import shim: `import b200ps` -> the package in ./tensorflow-k8s_b200/ (not an identifier).

Every submodule is aliased too, so `import b200ps.kernels` and `from b200ps import kernels` give
the one module object (one B200PSError class, one library handle).
"""
import importlib
import os
import sys

_REAL = "tensorflow-k8s_b200"
_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
_pkg = importlib.import_module(_REAL)
_RUNNABLE = ("launcher", "rank")  # `python -m b200ps.launcher`: runpy must find these through the package path, not an alias
for _name, _mod in list(sys.modules.items()):
    if _name.startswith(_REAL + ".") and _name[len(_REAL) + 1:] not in _RUNNABLE:
        sys.modules["b200ps" + _name[len(_REAL):]] = _mod
sys.modules["b200ps"] = _pkg
