"""Registers the package directory `gradient-sdf_b200/` (not a Python identifier) as module `gradient_sdf_b200`."""
import importlib.util
import sys
from pathlib import Path

ROOT = Path(__file__).resolve().parent
PKG_DIR = ROOT / "gradient-sdf_b200"


def load():
    name = "gradient_sdf_b200"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, PKG_DIR / "__init__.py",
                                                  submodule_search_locations=[str(PKG_DIR)])
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def load_oracle():
    """TEST INFRASTRUCTURE: the CPU oracle's ctypes wrapper (tests, smoke() and bench's CPU-baseline legs only)."""
    name = "pyoracle"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, ROOT / "oracle" / "pyoracle.py")
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod
