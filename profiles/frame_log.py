"""Steady-state device-clock spans of the pipeline's two kernels per frame (in-kernel %globaltimer, no events):
python profiles/frame_log.py [grad|base] [dev|host]"""
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
mode = sys.argv[2] if len(sys.argv) > 2 else "dev"
W, H = 640, 480
N = 100
frames, poses, K, _ = synth.make_sequence(100, W, H)
frames = frames[:N]
host = torch.from_numpy(frames).pin_memory()
dev = host.cuda()
nest = pkg.NormalEstimator(W, H, K, 11)
cls = pkg.MapGradPixelSdf if variant == "grad" else pkg.MapPixelSdf
sdf = cls(np.float32(0.01), np.float32(0.05), nest, log2_capacity=17)
sdf.scan_set_pose(np.array([0, 0, 0, 0, 0, 0, 1], np.float32))
sdf.scan_submit(depth_dev=dev[0].data_ptr(), track=False)
sdf.scan_collect()
sdf._check(sdf.L.gsdf_debug_frame_log(sdf.h, None))
hn = host.numpy()
ptrs = [dev[i].data_ptr() if mode == "dev" else hn[i].ctypes.data for i in range(1, N)]
sdf.scan_run(ptrs, on_device=(mode == "dev"))  # the frame loop in C (batches of frames per tracker launch)
log = np.zeros((256, 8), np.uint64)
sdf._check(sdf.L.gsdf_debug_frame_log(sdf.h, log.ctypes.data))
t = log[1:N].astype(np.int64)
t0 = t[40, 2]
print("frame   side: start    end   len |  main: start    end   len | gap to previous main end")
for i in range(40, 56):
    s0, s1, m0, m1 = (t[i, :4] - t0) / 1e3
    print(f"{i + 1:5d}   {s0:12.2f} {s1:7.2f} {s1 - s0:5.2f} | {m0:12.2f} {m1:7.2f} {m1 - m0:5.2f} | {(t[i, 2] - t[i - 1, 3]) / 1e3:6.2f}")
m = t[20:N - 5]
p = m[m[:, 6] > 0]
loop = (p[:, 5] - p[:, 2]) / 1e3
print(f"steady state per frame: passes {p[:, 6].mean():.2f}, first pass {(p[:, 4] - p[:, 2]).mean() / 1e3:.2f} us, Gauss-Newton loop "
      f"{loop.mean():.2f} us ({(loop / p[:, 6]).mean():.2f} us/pass), fusion + frame end {(p[:, 3] - p[:, 5]).mean() / 1e3:.2f} us")
print(f"steady state: period {np.diff(m[:, 2]).mean() / 1e3:.2f} us, main kernel {(m[:, 3] - m[:, 2]).mean() / 1e3:.2f} us, "
      f"side kernel {(m[:, 1] - m[:, 0]).mean() / 1e3:.2f} us, main-to-main gap {(m[1:, 2] - m[:-1, 3]).mean() / 1e3:.2f} us")
