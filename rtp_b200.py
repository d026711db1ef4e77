"""Import shim: the package directory is named after the reference ("ray-tracing-practice_b200"), which is not a
valid Python identifier.  `import rtp_b200` loads it by path and re-exports everything."""
import importlib.util
import os
import sys

_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ray-tracing-practice_b200")
_spec = importlib.util.spec_from_file_location("rtp_b200_pkg", os.path.join(_dir, "__init__.py"),
                                               submodule_search_locations=[_dir])
_mod = importlib.util.module_from_spec(_spec)
sys.modules["rtp_b200_pkg"] = _mod
_spec.loader.exec_module(_mod)
globals().update({k: v for k, v in vars(_mod).items() if not k.startswith("__")})
