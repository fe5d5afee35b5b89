"""Driver for ncu captures of the pipelined path: 14 frames through gsdf_scan_run (device-resident frames).
Usage: python profiles/run_pipeline.py [grad|base]"""
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
dev = torch.from_numpy(frames[:15]).cuda()
nest = pkg.NormalEstimator(W, H, K, 11)
cls = pkg.MapGradPixelSdf if variant == "grad" else pkg.MapPixelSdf
sdf = cls(np.float32(0.01), np.float32(0.05), nest, log2_capacity=17)
sdf.scan_set_pose(np.array([0, 0, 0, 0, 0, 0, 1], np.float32))
sdf.scan_submit(depth_dev=dev[0].data_ptr(), track=False)
sdf.scan_collect()
res = sdf.scan_run([dev[i].data_ptr() for i in range(1, 15)], on_device=True)
print(variant, "converged", sum(r.converged for r in res), "of", len(res), "iters", [r.iters for r in res],
      "updates/frame", sum(r.updates for r in res) / len(res))
