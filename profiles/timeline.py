"""Device timeline of the frame pipeline (labelled CUDA events): python profiles/timeline.py [grad|base]"""
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import gsdf_loader  # noqa: E402

pkg = gsdf_loader.load()
import importlib  # noqa: E402
import torch  # noqa: E402

synth = importlib.import_module("gradient_sdf_b200.synth")
variant = sys.argv[1] if len(sys.argv) > 1 else "grad"
W, H = 640, 480
frames, poses, K, _ = synth.make_sequence(100, W, H)
N = 30
dev = torch.from_numpy(frames[:N]).cuda()
nest = pkg.NormalEstimator(W, H, K, 11)
cls = pkg.MapGradPixelSdf if variant == "grad" else pkg.MapPixelSdf
sdf = cls(np.float32(0.01), np.float32(0.05), nest, log2_capacity=17)
sdf.scan_set_pose(np.array([0, 0, 0, 0, 0, 0, 1], np.float32))
sdf.scan_submit(depth_dev=dev[0].data_ptr(), track=False)
sdf.scan_collect()
for i in range(1, N):
    if i == 10:
        sdf.trace_enable(True)
    sdf.scan_submit(depth_dev=dev[i].data_ptr(), track=True)
    if sdf.scan_pending() == 4:
        sdf.scan_collect()
while sdf.scan_pending():
    sdf.scan_collect()
tr = sdf.trace_dump()
t0 = None
for label, us in tr[: 7 * 8]:
    print(f"{us:9.2f}  {label}")
main_begin = [us for l, us in tr if l == "main:result"]
print("mean frame period (us):", np.diff(main_begin).mean())
