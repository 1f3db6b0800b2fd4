"""Import shim: ``import b200dp`` loads the package that lives in
``parallel-and-distributed-deep-learning_b200/`` (a directory name Python cannot import)."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "parallel-and-distributed-deep-learning_b200")
_spec = importlib.util.spec_from_file_location(
    "b200dp", os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["b200dp"] = _mod
_spec.loader.exec_module(_mod)
