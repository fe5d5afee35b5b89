"""Small driver for ncu captures: builds a map from a few 640x480 frames and runs the hot kernels a few times.
Usage: python profiles/run_hotpath.py [grad|base] [voxel_size] [trunc]"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import gsdf_loader  # noqa: E402

pkg = gsdf_loader.load()
import importlib  # noqa: E402

synth = importlib.import_module("gradient_sdf_b200.synth")
variant = sys.argv[1] if len(sys.argv) > 1 else "grad"
vs = np.float32(sys.argv[2]) if len(sys.argv) > 2 else np.float32(0.01)
trunc = np.float32(sys.argv[3]) if len(sys.argv) > 3 else np.float32(5)
W, H = 640, 480
frames, poses, K, _ = synth.make_sequence(100, W, H)
frames, poses = frames[:8], poses[:8]
nest = pkg.NormalEstimator(W, H, K, 11)
cls = pkg.MapGradPixelSdf if variant == "grad" else pkg.MapPixelSdf
sdf = cls(vs, trunc * vs, nest, log2_capacity=17)
tot = 0
for i in range(6):
    pts, upd, blocks = sdf.update(frames[i], poses[i])
    tot += upd
opt = pkg.RigidPointOptimizer(sdf)
opt.set_pose(poses[5])
conv = opt.optimize(frames[6])
R, pn = pkg.pose_matrix(poses[6])
out = sdf.accumulate(R, pn[:3], 1)
print("variant", variant, "updates/frame", tot / 6, "blocks", blocks, "voxels", sdf.voxel_count()[0], "conv", conv,
      "iters", opt.last_iterations, "valid", out[28])
