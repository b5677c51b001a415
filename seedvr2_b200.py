"""Import alias: the product package directory is ``comfyui-seedvr2_b200`` (a hyphen cannot appear in an ``import``
statement), so ``import seedvr2_b200`` loads it through importlib and registers it under this name."""
import importlib
import os
import sys

_root = os.path.dirname(os.path.abspath(__file__))
if _root not in sys.path:
    sys.path.insert(0, _root)
_pkg = importlib.import_module("comfyui-seedvr2_b200")
sys.modules[__name__] = _pkg
