"""Converged count of the noisy C1 grad-sdf loop (tests/test_gpu_parity.py::test_noisy_grad_sdf_loop...), repeated: python profiles/r2_c1_repeat.py [n]"""
import sys
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import gsdf_loader
pkg = gsdf_loader.load()
import importlib
synth = importlib.import_module("gradient_sdf_b200.synth")
W, H, N = 640, 480, 100
frames, poses, K, _ = synth.make_sequence(N, W, H)
nest = pkg.NormalEstimator(W, H, K, 11)
vs = np.float32(0.02)
ident = np.array([0, 0, 0, 0, 0, 0, 1], np.float32)
for rep in range(int(sys.argv[1]) if len(sys.argv) > 1 else 4):
    sdf = pkg.MapGradPixelSdf(vs, np.float32(10) * vs, nest, log2_capacity=17)
    sdf.scan_set_pose(ident)
    sdf.scan_set_tracker(25, 1e-3, 1.0)
    sdf.scan_submit(depth_host=frames[0], track=False)
    sdf.scan_collect()
    res = sdf.scan_run([frames[i].ctypes.data for i in range(1, N)], on_device=False)
    conv = [r.converged for r in res]
    print("converged", sum(conv), "of", len(conv), "first failure at", conv.index(0) + 1 if 0 in conv else None, "iters head", [r.iters for r in res[:12]])
    sdf.close()
