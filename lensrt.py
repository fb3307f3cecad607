"""Import shim: `import lensrt as mi` loads the package in ./mitsuba3-camera_b200 (hyphenated directory
names are not importable by name)."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "mitsuba3-camera_b200")
_name = "mitsuba3_camera_b200"
if _name not in sys.modules:
    _spec = importlib.util.spec_from_file_location(_name, os.path.join(_dir, "__init__.py"),
                                                   submodule_search_locations=[_dir])
    _mod = importlib.util.module_from_spec(_spec)
    sys.modules[_name] = _mod
    _spec.loader.exec_module(_mod)
sys.modules[__name__] = sys.modules[_name]
