"""Import shim: `import presto_b200` == the package in ./presto-1_b200/ (whose name is not a Python identifier)."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
_pkg = importlib.import_module("presto-1_b200")
for _name in ("operators", "exchange", "tpch", "plans", "expressions"):
    try:
        importlib.import_module("presto-1_b200." + _name)
    except ModuleNotFoundError:
        pass
for _k in list(sys.modules):
    if _k.startswith("presto-1_b200."):
        sys.modules["presto_b200." + _k[len("presto-1_b200."):]] = sys.modules[_k]
sys.modules[__name__] = _pkg
