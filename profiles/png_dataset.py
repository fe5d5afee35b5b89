"""End-to-end frame rate of the Scan3D driver on a PNG dataset (SynthLoader convention, 640x480 16-bit depth PNGs written
by cv2): sequential decode against decode-ahead on T host threads (host/prefetch.h).  python profiles/png_dataset.py [n=160]"""
import re
import subprocess
import sys
import tempfile
import time
from pathlib import Path

import cv2
import numpy as np

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import gsdf_loader  # noqa: E402

gsdf_loader.load()
import importlib  # noqa: E402

synth = importlib.import_module("gradient_sdf_b200.synth")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 160
W, H = 640, 480
frames, poses, K, _ = synth.make_sequence(100, W, H)
frames = np.concatenate([frames, frames[::-1]])[:n] if n > 100 else frames[:n]  # there and back again: stays trackable
d = Path(tempfile.mkdtemp())
(d / "depth").mkdir()
(d / "rgb").mkdir()
np.savetxt(d / "intrinsics.txt", K.reshape(3, 3), fmt="%.9g")
for i, f in enumerate(frames):
    cv2.imwrite(str(d / "depth" / f"{i + 1:03d}.png"), np.round(f.astype(np.float64) * 1000).astype(np.uint16))
    (d / "rgb" / f"{i + 1:03d}.png").write_bytes(b"x")
size = sum(p.stat().st_size for p in (d / "depth").iterdir()) / n
print(f"{n} frames 640x480, {size / 1e3:.0f} kB per depth PNG")
BIN = ROOT / "gradient-sdf_b200" / "host" / "scan3d_b200"
base = [str(BIN), "--input", str(d), "--results", str(d / "res"), "--scan-type", "base-sdf", "--data-type", "synth",
        "--voxel-size", "0.01", "--trunc", "5", "--last", str(n - 1)]
for label, extra in (("sequential decode, synchronous calls", []), ("sequential decode, pipelined", ["--pipelined"]),
                     ("1 thread ahead", ["--pipelined", "--threads", "1"]), ("4 threads", ["--pipelined", "--threads", "4"]),
                     ("16 threads", ["--pipelined", "--threads", "16"])):
    t0 = time.perf_counter()
    r = subprocess.run(base + extra, capture_output=True, text=True, timeout=600)
    wall = time.perf_counter() - t0
    m = re.search(r"prefetched loop: (\d+) frames in ([\d.e-]+) s = ([\d.e+-]+) frames/s", r.stdout)
    conv = r.stdout.count("Convergence after")
    print(f"{label:40s} rc={r.returncode} converged={conv:4d} whole process {wall:6.2f} s" +
          (f"; frame loop {float(m.group(3)):8.1f} frames/s" if m else f"; ~{n / wall:6.1f} frames/s including start-up and writers"))
