"""BASELINE config C5: isolated fusion and tracker timings over voxel size x depth resolution on one B200, against
the HBM roofline (algorithmic bytes of SURVEY.md 8d / measured copy peak).  Per-kernel times come from the library's
CUDA-event brackets with 4 frames in flight (kernels queued back to back).
    python profiles/sweep.py [grad|base] > profiles/r1_sweep_<variant>.md"""
import json
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
peak = 6548.5
try:
    peak = float(json.loads((ROOT / "MEASURED_PEAKS.json").read_text())["hbm_gbs"])
except Exception:
    pass
N = 24
ident = np.array([0, 0, 0, 0, 0, 0, 1], np.float32)
print(f"| resolution | voxel | trunc | frames/s | updates/frame | fuse us | fuse GB/s (frac) | track us/pass | track GB/s (frac) | voxels |")
print("|---|---|---|---|---|---|---|---|---|---|")
for (W, H) in [(320, 240), (640, 480), (1280, 720)]:
    frames, poses, K, _ = synth.make_sequence(200, W, H)
    frames = frames[:N]
    dev = torch.from_numpy(frames).cuda()
    for vs, trunc in [(0.02, 10), (0.01, 5), (0.008, 5), (0.005, 5), (0.004, 5)]:
        nest = pkg.NormalEstimator(W, H, K, 11)
        cls = pkg.MapGradPixelSdf if variant == "grad" else pkg.MapPixelSdf
        vsf = np.float32(vs)
        sdf = cls(vsf, np.float32(trunc) * vsf, nest, log2_capacity=19)
        sdf.scan_set_pose(ident)
        res = []

        def run(profile):
            sdf.reset()
            sdf.scan_set_pose(ident)
            sdf.profile_enable(False)
            sdf.scan_submit(depth_dev=dev[0].data_ptr(), track=False)
            out = [sdf.scan_collect()]
            out += sdf.scan_run([dev[i].data_ptr() for i in range(1, 4)], on_device=True)
            if profile:
                sdf.profile_enable(True)
            torch.cuda.synchronize()
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            ext = torch.cuda.ExternalStream(sdf.stream)
            e0.record(ext)
            out += sdf.scan_run([dev[i].data_ptr() for i in range(4, N)], on_device=True)  # the frame loop in C
            e1.record(ext)
            e1.synchronize()
            return out, e0.elapsed_time(e1)

        out, ms = run(False)
        fps = 1e3 * (N - 4) / ms
        outp, _ = run(True)
        timed = outp[4:]
        n_f, ms_f = sdf.profile_read(1)
        n_a, ms_a = sdf.profile_read(2)
        upd = sum(r.updates for r in timed)
        passes = sum(r.track_iterations for r in timed)
        valid = sum(r.track_valid for r in timed)
        b_v, b_t = (40, 20) if variant == "grad" else (16, 36)
        fuse_bytes = n_f * W * H * 4 + upd * b_v
        track_bytes = passes * W * H * 4 + valid * b_t
        fg = fuse_bytes / max(ms_f, 1e-9) / 1e6
        tg = track_bytes / max(ms_a, 1e-9) / 1e6
        nv = sdf.voxel_count()[0]
        print(f"| {W}x{H} | {vs * 1000:g} mm | {trunc} | {fps:.0f} | {upd / max(1, n_f):.0f} | {1e3 * ms_f / max(1, n_f):.1f} | "
              f"{fg:.0f} ({fg / peak:.3f}) | {1e3 * ms_a / max(1, passes):.2f} | {tg:.0f} ({tg / peak:.3f}) | {nv} |", flush=True)
        sdf.close()
print(f"\npeak = {peak} GB/s (measured copy); fuse bytes = W*H*4 + N_upd*b_v, track bytes per pass = W*H*4 + N_valid*b_t; variant {variant}")
