"""Import helper: makes `torchsnapshot-1_b200/` importable as the package `torchsnapshot_b200`."""

import importlib.util
import os
import sys

_NAME = "torchsnapshot_b200"
_ROOT = os.path.dirname(os.path.abspath(__file__))
PACKAGE_DIR = os.path.join(_ROOT, "torchsnapshot-1_b200")


def load():
    mod = sys.modules.get(_NAME)
    if mod is not None:
        return mod
    spec = importlib.util.spec_from_file_location(
        _NAME, os.path.join(PACKAGE_DIR, "__init__.py"), submodule_search_locations=[PACKAGE_DIR]
    )
    mod = importlib.util.module_from_spec(spec)
    sys.modules[_NAME] = mod
    spec.loader.exec_module(mod)
    return mod
