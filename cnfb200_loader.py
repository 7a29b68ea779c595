"""Loads the package directory `continuousnf.jl_b200/` (its name is not a Python identifier)
under the module name `cnfb200`."""
import importlib.util
import os
import sys

_ROOT = os.path.dirname(os.path.abspath(__file__))


def load_package():
    if "cnfb200" in sys.modules:
        return sys.modules["cnfb200"]
    pkg_dir = os.path.join(_ROOT, "continuousnf.jl_b200")
    spec = importlib.util.spec_from_file_location(
        "cnfb200", os.path.join(pkg_dir, "__init__.py"), submodule_search_locations=[pkg_dir])
    mod = importlib.util.module_from_spec(spec)
    sys.modules["cnfb200"] = mod
    spec.loader.exec_module(mod)
    return mod
