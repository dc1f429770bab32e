"""Import alias: the package directory name (mandated by the repo layout) is not a Python identifier,
so `import vfi_b200` loads it under this name.  `vfi_b200.models.base`, `vfi_b200.engine`, ... resolve
inside that directory."""
import importlib.util
import os
import sys

_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)),
                    "video-frame-interpolation-a-benchmark-study-and-lessons-from-failed-novel-approache_b200")
_spec = importlib.util.spec_from_file_location("vfi_b200", os.path.join(_DIR, "__init__.py"),
                                               submodule_search_locations=[_DIR])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["vfi_b200"] = _mod
_spec.loader.exec_module(_mod)
