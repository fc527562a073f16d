"""Import shim: the package directory is named `waterlily.jl_b200/` (after the reference repo),
which is not a valid Python identifier, so it is loaded here under the name `waterlily_jl_b200`.
Usage: `import waterlily_jl_b200 as wl; sim = wl.Simulation(...)`."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "waterlily.jl_b200")
_spec = importlib.util.spec_from_file_location(
    "waterlily_jl_b200", os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["waterlily_jl_b200"] = _mod
_spec.loader.exec_module(_mod)
