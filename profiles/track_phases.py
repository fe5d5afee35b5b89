"""Phase timing of the persistent tracker from in-kernel %globaltimer stamps: python profiles/track_phases.py [grad|base]"""
import ctypes as C
import os
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
W, H = 640, 480
frames, poses, K, _ = synth.make_sequence(100, W, H)
nest = pkg.NormalEstimator(W, H, K, 11)
cls = pkg.MapGradPixelSdf if variant == "grad" else pkg.MapPixelSdf
sdf = cls(np.float32(0.01), np.float32(0.05), nest, log2_capacity=17)
for i in range(6):
    sdf.update(frames[i], poses[i])
opt = pkg.RigidPointOptimizer(sdf)
opt.set_pose(poses[5])
opt.optimize(frames[6])  # warm
n = C.c_int32(0)
sdf._check(sdf.L.gsdf_debug_track_timers(sdf.h, None, C.byref(n)))
opt.set_pose(poses[5])
conv = opt.optimize(frames[6])
raw = np.zeros(n.value * 220, np.uint64)
sdf._check(sdf.L.gsdf_debug_track_timers(sdf.h, raw.ctypes.data, C.byref(n)))
t = raw[: n.value * 24].reshape(n.value, 4, 6).astype(np.int64)
print(variant, "ctas", n.value, "converged", conv, "iters", opt.last_iterations)
names = ["lookups", "cta reduce", "exchange", "peer exch", "solve"]
for k in range(min(4, opt.last_iterations + 1)):
    tk = t[:, k, :]
    ok = tk[:, 0] > 0
    t0 = tk[ok, 0].min()
    print(f" pass {k}: begin spread {(tk[ok, 0].max() - t0) / 1e3:.2f} us; pass length {(tk[ok, 5].max() - t0) / 1e3:.2f} us")
    for i, nm in enumerate(names):
        d = (tk[ok, i + 1] - tk[ok, i]) / 1e3
        print(f"   {nm:12s} mean {d.mean():6.2f}  max {d.max():6.2f} us   (phase ends at {(tk[ok, i + 1].max() - t0) / 1e3:6.2f} us)")


# anatomy of the fused track + fuse launch inside the frame pipeline (entry, loop left, fusion done, exit per CTA)
def anatomy(tag, inflight):
    sdf.reset()
    sdf.scan_set_pose(np.array([0, 0, 0, 0, 0, 0, 1], np.float32))
    sdf.scan_set_tracker(25, 1e-3, 1.0)
    sdf.scan_submit(depth_host=frames[0], track=False)
    sdf.scan_collect()
    for i in range(1, 20):
        sdf.scan_submit(depth_host=frames[i], track=True)
        sdf.scan_collect()
    sdf._check(sdf.L.gsdf_debug_track_timers(sdf.h, None, C.byref(n)))
    for i in range(20, 20 + inflight):
        sdf.scan_submit(depth_host=frames[i], track=True)
    res = [sdf.scan_collect() for _ in range(inflight)]
    raw = np.zeros(n.value * 220, np.uint64)
    sdf._check(sdf.L.gsdf_debug_track_timers(sdf.h, raw.ctypes.data, C.byref(n)))
    x = raw[n.value * 24:n.value * 28].reshape(n.value, 4).astype(np.int64)
    fw = raw[n.value * 28:].reshape(-1, 8).astype(np.int64)
    p = raw[: n.value * 24].reshape(n.value, 4, 6).astype(np.int64)
    t0 = x[:, 0].min()
    f = lambda a: f"first {(a.min() - t0) / 1e3:6.2f}  mean {(a.mean() - t0) / 1e3:6.2f}  last {(a.max() - t0) / 1e3:6.2f}"
    print(f"{tag}: last frame iters {res[-1].iters} converged {res[-1].converged}")
    print("   CTA entry      ", f(x[:, 0]))
    for k in range(min(4, res[-1].iters + 1)):
        print(f"   pass {k} begin   ", f(p[:, k, 0]), "  solved", f(p[:, k, 5]))
    print("   loop left      ", f(x[:, 1]))
    print("   fusion done    ", f(x[:, 2]))
    print("   exit           ", f(x[:, 3]))
    # per-warp anatomy of the fusion phase (warps that had work)
    if os.environ.get("GSDF_SAVE_FUSE_STAMPS"):
        np.save(os.environ["GSDF_SAVE_FUSE_STAMPS"] + "_" + tag.split()[0] + ".npy", fw)
    fw = fw[fw[:, 0] > 0]
    if len(fw):
        units, probes, reds = fw[:, 6] & 0xFFFFFFFF, fw[:, 6] >> 32, fw[:, 7]
        busy = fw[units > 0]
        d = lambda a, b: (busy[:, b] - busy[:, a]) / 1e3
        print(f"   fusion per warp: {len(busy)} of {len(fw)} warps had units; units/warp {units.mean():.2f} (max {units.max()}), "
              f"probes/warp {probes.mean():.2f}, reductions/warp {reds.mean():.1f}")
        print(f"      warp total        mean {d(0, 5).mean():6.2f}  max {d(0, 5).max():6.2f} us")
        busy = busy[busy[:, 4] > 0]  # warps that committed something (a chunk of invalid normals commits nothing)
        if len(busy):
            print(f"      pixel data        mean {d(0, 1).mean():6.2f}  max {d(0, 1).max():6.2f} us")
            print(f"      first slot known  mean {d(1, 2).mean():6.2f}  max {d(1, 2).max():6.2f} us")
            print(f"      first merge       mean {d(2, 3).mean():6.2f}  max {d(2, 3).max():6.2f} us")
            print(f"      first reds issued mean {d(3, 4).mean():6.2f}  max {d(3, 4).max():6.2f} us")
            print(f"      remaining units   mean {d(4, 5).mean():6.2f}  max {d(4, 5).max():6.2f} us")
        else:
            print("      (no first-unit stamps in this launch)")
        print(f"      entry spread {(fw[:, 0].max() - fw[:, 0].min()) / 1e3:.2f} us, end spread {(fw[:, 5].max() - fw[:, 5].min()) / 1e3:.2f} us")


anatomy("one frame in flight", 1)
anatomy("four frames in flight (stamps of the last)", 4)


# the same stamps from a launch in the MIDDLE of a long gsdf_scan_run (GSDF_DBG_SKIP): steady state, the frame-only kernels of
# later frames run beside the tracker
import os  # noqa: E402
import torch  # noqa: E402
if "steady" in sys.argv:
    os.environ["GSDF_DBG_SKIP"] = "50"  # read by the library when the stamps are armed
    dev = torch.from_numpy(frames[:100]).cuda()
    sdf.reset()
    sdf.scan_set_pose(np.array([0, 0, 0, 0, 0, 0, 1], np.float32))
    sdf.scan_submit(depth_dev=dev[0].data_ptr(), track=False)
    sdf.scan_collect()
    sdf.scan_run([dev[i].data_ptr() for i in range(1, 10)], on_device=True)
    sdf._check(sdf.L.gsdf_debug_track_timers(sdf.h, None, C.byref(n)))
    res = sdf.scan_run([dev[i].data_ptr() for i in range(10, 100)], on_device=True)
    raw = np.zeros(n.value * 220, np.uint64)
    sdf._check(sdf.L.gsdf_debug_track_timers(sdf.h, raw.ctypes.data, C.byref(n)))
    t = raw[: n.value * 24].reshape(n.value, 4, 6).astype(np.int64)
    print("steady state (launch in the middle of a 90-frame run), passes of the launch's last frame:")
    for k in range(4):
        tk = t[:, k, :]
        ok = tk[:, 0] > 0
        if not ok.any():
            continue
        t0 = tk[ok, 0].min()
        print(f" pass {k}: begin spread {(tk[ok, 0].max() - t0) / 1e3:.2f} us; pass length {(tk[ok, 5].max() - t0) / 1e3:.2f} us")
        for i, nm in enumerate(names):
            d = (tk[ok, i + 1] - tk[ok, i]) / 1e3
            print(f"   {nm:12s} mean {d.mean():6.2f}  max {d.max():6.2f} us   (phase ends at {(tk[ok, i + 1].max() - t0) / 1e3:6.2f} us)")
