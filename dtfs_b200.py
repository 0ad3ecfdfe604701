"""Import shim: the package directory is ``distributed-tf-serving_b200`` (hyphen, the project's
given name), which the ``import`` statement cannot spell.  ``import dtfs_b200`` gives that package."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
_pkg = importlib.import_module("distributed-tf-serving_b200")
for _name in ("binding", "build", "dcn_client", "weights"):
    sys.modules[__name__ + "." + _name] = getattr(_pkg, _name)
sys.modules[__name__] = _pkg
