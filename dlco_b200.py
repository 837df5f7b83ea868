"""Import shim: the package directory is `opencv-dlco_b200/` (not a valid Python identifier),
so `import dlco_b200` loads it under the module name `opencv_dlco_b200`."""
import importlib.util
import os
import sys

_pkg = os.path.join(os.path.dirname(os.path.abspath(__file__)), "opencv-dlco_b200")
if "opencv_dlco_b200" not in sys.modules:
    _spec = importlib.util.spec_from_file_location(
        "opencv_dlco_b200", os.path.join(_pkg, "__init__.py"), submodule_search_locations=[_pkg])
    _mod = importlib.util.module_from_spec(_spec)
    sys.modules["opencv_dlco_b200"] = _mod
    _spec.loader.exec_module(_mod)
from opencv_dlco_b200 import *  # noqa: F401,F403,E402
from opencv_dlco_b200 import Context, Bank, DlcoError, load_library, LIB_PATH, SIGNATURES  # noqa: F401,E402
import opencv_dlco_b200.learn as learn  # noqa: F401,E402
import opencv_dlco_b200.shard as shard  # noqa: F401,E402
