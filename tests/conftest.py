import os
import sys
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
# the reference build (oracle/_ref) is bit-reproducible only single-threaded
os.environ.setdefault("OMP_THREAD_LIMIT", "1")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    import gsdf_loader
    return gsdf_loader.load()


@pytest.fixture(scope="session")
def po():
    import gsdf_loader
    mod = gsdf_loader.load_oracle()
    if not mod.ORACLE_SO.exists():
        mod.build()
    return mod


@pytest.fixture(scope="session")
def synth(pkg):
    import importlib
    return importlib.import_module("gradient_sdf_b200.synth")


GOLDEN = ROOT / "tests" / "golden"


def load_golden(name):
    import numpy as np
    z = np.load(GOLDEN / f"{name}.npz")
    g = {k: z[k] for k in z.files}
    g["depth"] = g["depth_mm"].astype(np.float32) * np.float32(0.001)
    if "free_depth_mm" in g:
        g["free_depth"] = g["free_depth_mm"].astype(np.float32) * np.float32(0.001)
    return g
