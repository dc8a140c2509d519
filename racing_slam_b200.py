"""Loader shim: makes `import racing_slam_b200` resolve to the package directory
`racing-slam_b200/` (a hyphen is not a valid module name)."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "racing-slam_b200")
_spec = importlib.util.spec_from_file_location("racing_slam_b200", os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["racing_slam_b200"] = _mod
_spec.loader.exec_module(_mod)
