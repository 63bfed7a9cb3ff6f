"""Import shim: the package directory is ``image-restoration_b200/`` (not a Python identifier);
``import b200ir`` loads it under the module name ``b200ir``."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "image-restoration_b200")
_spec = importlib.util.spec_from_file_location("b200ir", os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["b200ir"] = _mod
_spec.loader.exec_module(_mod)
