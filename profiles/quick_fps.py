"""Frames/s of the frame loop (gsdf_scan_run, frames resident in HBM), median of R repetitions — for A/B runs of
environment switches: python profiles/quick_fps.py [grad|base] [N frames] [R]"""
import sys
import time
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
N = int(sys.argv[2]) if len(sys.argv) > 2 else 150
R = int(sys.argv[3]) if len(sys.argv) > 3 else 5
W, H = 640, 480
frames, poses, K, _ = synth.make_sequence(max(100, N + 1), W, H, turn=2 * np.pi * max(100, N + 1) / 100)
dev = torch.from_numpy(frames[:N + 1]).cuda()
nest = pkg.NormalEstimator(W, H, K, 11)
cls = pkg.MapGradPixelSdf if variant == "grad" else pkg.MapPixelSdf
sdf = cls(np.float32(0.01), np.float32(0.05), nest, log2_capacity=17)
ident = np.array([0, 0, 0, 0, 0, 0, 1], np.float32)
ext = torch.cuda.ExternalStream(sdf.stream)
fps, conv = [], 0
for rep in range(R + 1):
    sdf.reset()
    sdf.scan_set_pose(ident)
    sdf.scan_submit(depth_dev=dev[0].data_ptr(), track=False)
    sdf.scan_collect()
    sdf.scan_run([dev[i].data_ptr() for i in range(1, 6)], on_device=True)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record(ext)
    res = sdf.scan_run([dev[i].data_ptr() for i in range(6, N + 1)], on_device=True)
    e1.record(ext)
    e1.synchronize()
    wall = time.perf_counter() - t0
    if rep:
        fps.append((N - 5) / (e0.elapsed_time(e1) * 1e-3))
    conv = sum(r.converged for r in res)
    iters = sum(r.track_iterations for r in res)
print(f"{variant} {N - 5} frames x {R}: median {np.median(fps):.0f} frames/s (min {min(fps):.0f} max {max(fps):.0f}), "
      f"{1e6 / np.median(fps):.2f} us/frame, converged {conv}, passes/frame {iters / (N - 5):.2f}, wall of last {wall * 1e3:.2f} ms")
