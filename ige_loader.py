"""Import helper: the package directory is named ``isaacgymenvs-ma_b200`` (not a valid Python
identifier), so it is registered under the module name ``isaacgymenvs_ma_b200``."""
import importlib.util
import os
import sys

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG_DIR = os.path.join(ROOT, "isaacgymenvs-ma_b200")
PKG_NAME = "isaacgymenvs_ma_b200"


def load_package():
    if PKG_NAME in sys.modules:
        return sys.modules[PKG_NAME]
    spec = importlib.util.spec_from_file_location(
        PKG_NAME, os.path.join(PKG_DIR, "__init__.py"), submodule_search_locations=[PKG_DIR])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[PKG_NAME] = mod
    spec.loader.exec_module(mod)
    return mod
