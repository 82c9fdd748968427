"""Import shim: the package directory is named `optix-renderer_b200` (hyphen), so plain `import`
cannot name it. `import norib200 as nb` gives the same module."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
_mod = importlib.import_module("optix-renderer_b200")
sys.modules[__name__] = _mod
