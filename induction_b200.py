"""Import alias: the product package lives in ``induction-network-on-fewrel_b200/`` (a directory name
Python cannot import directly); ``import induction_b200`` loads it under this importable name."""
import importlib.util
import os
import sys

_here = os.path.dirname(os.path.abspath(__file__))
_pkg_dir = os.path.join(_here, "induction-network-on-fewrel_b200")
_spec = importlib.util.spec_from_file_location(
    "induction_b200", os.path.join(_pkg_dir, "__init__.py"), submodule_search_locations=[_pkg_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["induction_b200"] = _mod
_spec.loader.exec_module(_mod)
