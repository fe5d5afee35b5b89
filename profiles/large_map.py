"""Room-scale stress run (BASELINE.json configs[3] style): grad-sdf track+fuse over a long synthetic trajectory inside
a bumpy spherical room, ~10^8 allocated voxels on ONE B200.  The frames are rendered on the GPU with torch (input
generation, not the product), everything else goes through the C ABI (gsdf_scan_run).
    python profiles/large_map.py [n_frames=2000] [voxel_mm=4] [trunc=10] [grad|base]
Under torchrun (N ranks, one GPU each) the map is sharded by block key and the tracker sums are exchanged over NVLink
peer memory, as in bench.py."""
import json
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import gsdf_loader  # noqa: E402

pkg = gsdf_loader.load()
import importlib  # noqa: E402

synth = importlib.import_module("gradient_sdf_b200.synth")
n_frames = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
vs = np.float32((float(sys.argv[2]) if len(sys.argv) > 2 else 4.0) / 1000.0)
trunc = np.float32(float(sys.argv[3]) if len(sys.argv) > 3 else 10.0)
variant = sys.argv[4] if len(sys.argv) > 4 else "grad"
W, H = 640, 480
K = synth.scaled_intrinsics(W, H)
import os  # noqa: E402
world, rank, local = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0")), int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    import torch.distributed as dist
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)

# scene: a spherical shell of radius 3.3 m seen from inside, 64 bumps (spheres) sticking out of it
rng = np.random.default_rng(42)
R_ROOM = 3.3
d = rng.standard_normal((64, 3))
d[:, 2] *= 0.45  # keep the bumps near the band the camera looks at
d /= np.linalg.norm(d, axis=1, keepdims=True)
bumps = np.concatenate([d * 3.15, rng.uniform(0.25, 0.55, (64, 1))], axis=1)


def trajectory(n):
    """One turn on a circle of radius 0.8 m, looking outward, pitch wobbling +-15 degrees: the wall 2.5 m away moves
    ~1 cm per frame at n = 2000."""
    poses = np.zeros((n, 7))
    for k in range(n):
        th = 2 * np.pi * k / n
        pitch = np.deg2rad(15.0) * np.sin(6 * th)
        c = np.array([0.8 * np.cos(th), 0.8 * np.sin(th), 0.0])
        fwd = np.array([np.cos(th) * np.cos(pitch), np.sin(th) * np.cos(pitch), np.sin(pitch)])
        Rm = synth._look_at(c, c + fwd)
        poses[k, :3] = c
        poses[k, 3:] = synth.rot_to_quat(Rm)
    return poses


def render(pose7):
    u, v = torch.meshgrid(torch.arange(W, device=dev, dtype=torch.float64), torch.arange(H, device=dev, dtype=torch.float64),
                          indexing="xy")
    cu, cv = (u - K[0, 2]) / K[0, 0], (v - K[1, 2]) / K[1, 1]
    A = cu * cu + cv * cv + 1.0
    Rm = torch.from_numpy(synth.quat_to_rot(pose7[3:])).to(dev)
    t = torch.from_numpy(pose7[:3]).to(dev)
    # shell (far root, centre = origin)
    c0 = (-t) @ Rm
    B = -2.0 * (cu * c0[0] + cv * c0[1] + c0[2])
    C = c0 @ c0 - R_ROOM ** 2
    depth = 0.5 * (-B + torch.sqrt(B * B - 4.0 * A * C)) / A
    cs = (torch.from_numpy(bumps[:, :3]).to(dev) - t) @ Rm
    for s in range(bumps.shape[0]):
        B = -2.0 * (cu * cs[s, 0] + cv * cs[s, 1] + cs[s, 2])
        C = cs[s] @ cs[s] - bumps[s, 3] ** 2
        disc = B * B - 4.0 * A * C
        z = torch.where(disc >= 0, 0.5 * (-B - torch.sqrt(disc.clamp(min=0))) / A, torch.full_like(A, float("inf")))
        z = torch.where(z > 0, z, torch.full_like(z, float("inf")))
        depth = torch.minimum(depth, z)
    q = torch.clamp(torch.round(depth * 1000.0), 0, 65535)  # the loaders' uint16 millimetre round trip
    return (q.to(torch.float32) * np.float32(1.0 / 1000.0))


poses = trajectory(n_frames)
t0 = time.time()
frames = torch.stack([render(poses[k]) for k in range(n_frames)])
torch.cuda.synchronize()
if rank == 0:
    print(f"rendered {n_frames} frames in {time.time() - t0:.1f} s ({frames.numel() * 4 / 2**30:.2f} GiB in HBM)", flush=True)

nest = pkg.NormalEstimator(W, H, K.astype(np.float32), 11)
cls = pkg.MapGradPixelSdf if variant == "grad" else pkg.MapPixelSdf
sdf = cls(vs, trunc * vs, nest, log2_capacity=21 if world == 1 else 20, device=local, world_size=world, rank=rank)
if world > 1:
    import ctypes as C
    idt = torch.zeros(128, dtype=torch.uint8, device=dev)
    if rank == 0:
        buf = (C.c_uint8 * 128)()
        assert pkg.lib().gsdf_comm_unique_id(buf) == 0
        idt.copy_(torch.tensor(list(buf), dtype=torch.uint8))
    dist.broadcast(idt, 0)
    sdf.comm_init(bytes(idt.cpu().tolist()))
    hd = torch.tensor(list(sdf.p2p_export()), dtype=torch.uint8, device=dev)
    allh = [torch.zeros_like(hd) for _ in range(world)]
    dist.all_gather(allh, hd)
    sdf.p2p_connect(bytes(torch.cat(allh).cpu().tolist()))
    dist.barrier()
sdf.scan_set_pose(poses[0].astype(np.float32))
sdf.scan_set_tracker(25, 1e-3, 1.0)
sdf.scan_submit(depth_dev=frames[0].data_ptr(), track=False)
sdf.scan_collect()
ext = torch.cuda.ExternalStream(sdf.stream)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
e0.record(ext)
ptrs = [frames[k].data_ptr() for k in range(1, n_frames)]
res = sdf.scan_run(ptrs, on_device=True)
e1.record(ext)
torch.cuda.synchronize()
ms = e0.elapsed_time(e1)
upd_all = float(sum(r.updates for r in res))
if world > 1:
    t = torch.tensor([ms], dtype=torch.float64, device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t)
    t = torch.tensor([upd_all], dtype=torch.float64, device=dev); dist.all_reduce(t); upd_all = float(t)
# anatomy: the last 200 frames once more over the finished map (in-kernel spans, outside the timed region)
sdf._check(sdf.L.gsdf_debug_frame_log(sdf.h, None))
sdf.scan_run(ptrs[-200:], on_device=True)
log = np.zeros((256, 4), np.uint64)
sdf._check(sdf.L.gsdf_debug_frame_log(sdf.h, log.ctypes.data))
tl = log[(log[:, 3] > 0) & (log[:, 1] > 0)].astype(np.int64)
tl = tl[np.argsort(tl[:, 2])][2:-2]
anatomy = {"period": float(np.diff(tl[:, 2]).mean() / 1e3), "track_fuse_kernel": float((tl[:, 3] - tl[:, 2]).mean() / 1e3),
           "frame_only_kernel_in_shadow": float((tl[:, 1] - tl[:, 0]).mean() / 1e3),
           "gap_between_track_fuse_kernels": float((tl[1:, 2] - tl[:-1, 3]).mean() / 1e3)}
conv = sum(r.converged for r in res)
err_t = max(float(np.abs(np.array(r.pose7[:3]) - poses[k + 1, :3]).max()) for k, r in enumerate(res))
last = np.array(res[-1].pose7[:])
voxels, blocks = sdf.voxel_count()
if world > 1:
    t = torch.tensor([voxels, blocks], dtype=torch.float64, device=dev); dist.all_reduce(t); voxels, blocks = int(t[0]), int(t[1])
free, total = torch.cuda.mem_get_info()
out = {
    "workload": f"C4-style: {variant}-sdf track+fuse, {n_frames} frames 640x480 inside a bumpy 3.3 m spherical room, "
                f"{float(vs) * 1e3:.0f} mm voxels, trunc {float(trunc):.0f}, {world} B200" + (" (hash sharded by block key)" if world > 1 else ""),
    "n_gpus": world,
    "frames_per_s": 1e3 * (n_frames - 1) / ms, "ms_per_frame": ms / (n_frames - 1),
    "converged": conv, "tracked": len(res), "gn_iterations_per_frame": sum(r.iters for r in res) / len(res),
    "updates_per_frame": upd_all / len(res),
    "voxel_updates_per_s": 1e3 * upd_all / ms,
    "voxels": int(voxels), "blocks": int(blocks), "hash_load_per_rank": blocks / float(world) / float(1 << (21 if world == 1 else 20)),
    "max_translation_error_m": err_t, "final_pose": last.tolist(), "final_pose_true": poses[-1].tolist(),
    "hbm_used_GiB": (total - free) / 2**30, "anatomy_us_200_frames_replayed": anatomy,
}
if rank == 0:
    print(json.dumps(out))
if world > 1:
    dist.destroy_process_group()
