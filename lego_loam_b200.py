"""Import shim: the package directory is named `lego-loam_comment_b200` (not a valid Python identifier), so this
module loads it under the importable name `lego_loam_b200`.  `import lego_loam_b200` from the repo root."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lego-loam_comment_b200")
_spec = importlib.util.spec_from_file_location(
    "lego_loam_b200", os.path.join(_dir, "__init__.py"), submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["lego_loam_b200"] = _mod
_spec.loader.exec_module(_mod)
