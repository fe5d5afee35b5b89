"""Isolated, warm-cache kernel durations through the synchronous entry points (no overlap between kernels):
python profiles/kernel_times.py [grad|base] [voxel] [trunc] [W H]"""
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
W, H = (int(sys.argv[4]), int(sys.argv[5])) if len(sys.argv) > 5 else (640, 480)
N = 24
frames, poses, K, _ = synth.make_sequence(100, W, H)
frames, poses = frames[:N], poses[:N]
nest = pkg.NormalEstimator(W, H, K, 11)
cls = pkg.MapGradPixelSdf if variant == "grad" else pkg.MapPixelSdf
sdf = cls(vs, trunc * vs, nest, log2_capacity=18)
opt = pkg.RigidPointOptimizer(sdf)
ident = np.array([0, 0, 0, 0, 0, 0, 1], np.float32)
sdf.setup(frames[0])
opt.set_pose(ident)
upd = iters = conv_n = 0
for i in range(1, N):
    if i == 4:
        sdf.profile_enable(True)
        it0, v0 = sdf.track_stats()
    conv = opt.optimize(frames[i])
    if conv:
        st = sdf.update(None, opt.pose(), stored=True)
        if i >= 4:
            upd += st[1]
            conv_n += 1
names = ["normals", "fuse", "track", "prepare"]
it1, v1 = sdf.track_stats()
for k, nm in enumerate(names):
    n, ms = sdf.profile_read(k)
    extra = ""
    if nm == "track":
        extra = f"  passes/launch {(it1 - it0) / max(1, n):.2f}  us/pass {1e3 * ms / max(1, it1 - it0):.2f}  valid/pass {(v1 - v0) / max(1, it1 - it0):.0f}"
    if nm == "fuse":
        extra = f"  updates/launch {upd / max(1, n):.0f}"
    print(f"{variant} {W}x{H} vs={vs} {nm:8s} launches {n:3d}  us/launch {1e3 * ms / max(1, n):8.2f}{extra}")
print("voxels", sdf.voxel_count())
