"""CPU suite: the C-ABI library loads, exports every symbol include/gsdf_b200.h declares, and refuses to run
without a GPU (no CPU fallback)."""
import ctypes as C
import re
from pathlib import Path

import pytest

ROOT = Path(__file__).resolve().parents[1]


def test_exports_match_header(pkg):
    header = (ROOT / "include" / "gsdf_b200.h").read_text()
    declared = set(re.findall(r"^(?:gsdf_status|const char\*|void\*?|int64_t|int32_t)\s+(gsdf_[a-z0-9_]+)\s*\(", header, re.M))
    L = pkg.lib()
    for sym in sorted(declared):
        assert hasattr(L, sym), f"{sym} declared in include/gsdf_b200.h but not exported"
    assert declared == set(pkg.EXPORTS), declared ^ set(pkg.EXPORTS)
    assert b"sm_100a" in L.gsdf_version()


def test_config_struct_layout(pkg):
    # gsdf_config: 5 x 4 B + 2 x 4 B + 9 x 4 B + 5 x 4 B
    assert C.sizeof(pkg.Config) == 4 * (5 + 2 + 9 + 5)
    assert C.sizeof(pkg.IntegrateStats) == 24


def test_invalid_arguments_are_rejected(pkg):
    L = pkg.lib()
    h = C.c_void_p()
    assert L.gsdf_create(None, C.byref(h)) == 1
    cfg = pkg.Config()
    cfg.variant = 7
    assert L.gsdf_create(C.byref(cfg), C.byref(h)) == 1
    assert L.gsdf_destroy(None) == 1
    assert L.gsdf_integrate(None, None, None, None) == 1


def test_no_cpu_fallback(pkg):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import numpy as np
    nest = pkg.NormalEstimator(64, 48, np.array([[50, 0, 32], [0, 50, 24], [0, 0, 1]], np.float32))
    with pytest.raises(pkg.GsdfError) as e:
        pkg.MapGradPixelSdf(0.02, 0.1, nest, log2_capacity=10)
    assert e.value.status == 2  # GSDF_ERR_CUDA


def test_header_is_plain_c():
    """The boundary is a C ABI: the header must compile as C99 and as C++11 on its own (no torch, no CUDA types)."""
    import subprocess
    hdr = str(ROOT / "include" / "gsdf_b200.h")
    for cmd in (["gcc", "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-fsyntax-only", "-x", "c", hdr],
                ["g++", "-std=c++11", "-Wall", "-Werror", "-fsyntax-only", "-x", "c++", hdr]):
        r = subprocess.run(cmd, capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
