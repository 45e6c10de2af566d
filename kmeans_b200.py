"""Import shim: the package directory is named `k-means-demo_b200/` (not a legal Python
identifier), so `import kmeans_b200` loads it from there."""
import importlib.util
import os
import sys

_root = os.path.join(os.path.dirname(os.path.abspath(__file__)), "k-means-demo_b200")
_spec = importlib.util.spec_from_file_location("kmeans_b200", os.path.join(_root, "__init__.py"),
                                               submodule_search_locations=[_root])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["kmeans_b200"] = _mod
_spec.loader.exec_module(_mod)
