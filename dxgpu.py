"""Import shim: `import dxgpu` -> the package in ./data-accelerator_b200 (hyphenated directory)."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
sys.modules[__name__] = importlib.import_module("data-accelerator_b200")
