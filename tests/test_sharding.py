"""Multi-GPU sharding protocol: the map is partitioned by voxel-block key, every rank sees the
whole frame and commits / counts only what it owns, and the 29 tracker sums are added across ranks.

CPU suite: the protocol run by two gloo processes over the oracle's shard emulation (torch.distributed, 127.0.0.1).
GPU suite: two shards on ONE device (world_size = 2 handles, partial sums through gsdf_accumulate_local) against the
unsharded map; the real NCCL path is exercised by tests/multigpu_check.py under torchrun when >= 2 GPUs exist."""
import os
import socket
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]


def test_shard_owner_host_restatement(pkg, po):
    rng = np.random.default_rng(5)
    keys = rng.integers(-4000, 4000, (3000, 3)).astype(np.int32)
    o = po.Oracle(0, 0.02, 5.0, 32, 24, np.array([[50, 0, 16], [0, 50, 12], [0, 0, 1]], np.float32))
    for world in (2, 3, 8):
        own = pkg.shard_owner(keys, world)
        assert np.array_equal(own, o.voxel_owner(keys, world))
        assert own.min() == 0 and own.max() == world - 1
        # all voxels of one 8^3 block have one owner; shards are balanced to a few percent
        blocks = {}
        for k, r in zip(map(tuple, keys >> 3), own):
            assert blocks.setdefault(k, r) == r
        counts = np.bincount(own, minlength=world)
        assert counts.min() > 0.8 * len(keys) / world


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


WORKER = r'''
import os, sys
sys.path.insert(0, os.environ["GSDF_ROOT"])
os.environ["OMP_THREAD_LIMIT"] = "1"
import numpy as np, torch, torch.distributed as dist
import gsdf_loader
pkg = gsdf_loader.load(); po = gsdf_loader.load_oracle()
import importlib; synth = importlib.import_module("gradient_sdf_b200.synth")
dist.init_process_group("gloo", init_method=f"tcp://127.0.0.1:{os.environ['GSDF_PORT']}", rank=int(os.environ["RANK"]), world_size=2)
rank, world = dist.get_rank(), 2
W, H = 160, 120
frames, poses, K, _ = synth.make_sequence(100, W, H)
V = int(os.environ["GSDF_VARIANT"])  # 0 grad-sdf, 1 base-sdf (halo copies of the +1 neighbours' low faces)
full = po.Oracle(V, 0.02, 5.0, W, H, K)
mine = po.Oracle(V, 0.02, 5.0, W, H, K); mine.set_shard(world, rank)
upd_full = upd_mine = 0
for i in range(3):
    upd_full += full.update(frames[i], poses[i])[2]
    upd_mine += mine.update(frames[i], poses[i])[2]
# every sample is committed by exactly one rank
t = torch.tensor([upd_mine], dtype=torch.int64); dist.all_reduce(t)
assert int(t) == upd_full, (int(t), upd_full)
# union of the shards' key sets == the unsharded key set, shards disjoint, owner function respected
k, d, w, g = mine.download()
assert (pkg.shard_owner(k, world) == rank).all()
gathered = [None, None]
dist.all_gather_object(gathered, k.tobytes())
ks = [np.frombuffer(b, np.int32).reshape(-1, 3) for b in gathered]
union = {tuple(x) for x in ks[0]} | {tuple(x) for x in ks[1]}
assert len(union) == len(ks[0]) + len(ks[1]) == full.voxel_count()
assert union == {tuple(x) for x in full.download()[0]}
# tracker: partial sums added across ranks == unsharded sums (counter exactly), then identical poses on every rank
R, pn = pkg.pose_matrix(poses[3])
part = torch.from_numpy(mine.accumulate(frames[3], R, pn[:3]).astype(np.float64)); dist.all_reduce(part)
ref = full.accumulate(frames[3], R, pn[:3])
assert part[28].item() == ref[28]
assert np.abs(part.numpy()[:27] - ref[:27]).max() <= 1e-4 * np.abs(ref[:27]).max()
H6 = np.zeros((6, 6), np.float32); o = 6
for a in range(6):
    for b in range(a, 6):
        H6[a, b] = H6[b, a] = part[o]; o += 1
xi = pkg.llt_solve6(H6, part[:6].numpy().astype(np.float32))
pose = pkg.pose_apply_increment(pn, xi)
both = [None, None]; dist.all_gather_object(both, pose.tobytes())
assert both[0] == both[1]
dist.destroy_process_group()
print("rank", rank, "ok")
'''


@pytest.mark.parametrize("variant", [0, 1])
def test_sharding_protocol_two_gloo_ranks(tmp_path, variant):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    port = _free_port()
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), GSDF_PORT=str(port), GSDF_ROOT=str(ROOT), MASTER_ADDR="127.0.0.1",
                   GSDF_VARIANT=str(variant))
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.STDOUT, text=True))
    outs = [p.communicate(timeout=240)[0] for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0 and "ok" in o, o[-3000:]


@pytest.mark.gpu
@pytest.mark.parametrize("variant", ["grad", "base"])
def test_two_shards_on_one_device_reproduce_the_unsharded_map(pkg, po, synth, variant):
    W, H = 320, 240
    cls = pkg.MapGradPixelSdf if variant == "grad" else pkg.MapPixelSdf
    frames, poses, K, _ = synth.make_sequence(100, W, H)
    nest = pkg.NormalEstimator(W, H, K, 11)
    vs, T = np.float32(0.01), np.float32(0.05)
    full = cls(vs, T, nest, log2_capacity=16)
    shards = [cls(vs, T, nest, log2_capacity=16, world_size=2, rank=r) for r in range(2)]
    for i in range(3):
        pts, upd, _ = full.update(frames[i], poses[i])
        parts = [s.update(frames[i], poses[i]) for s in shards]
        assert sum(p[1] for p in parts) == upd           # every sample committed exactly once
        assert all(p[0] == pts for p in parts)           # every rank walks every valid pixel
    kf, df, wf, gf = full.download()
    ks = [s.download() for s in shards]
    for r, (k, d, w, g) in enumerate(ks):
        assert (pkg.shard_owner(k, 2) == r).all()
    k = np.concatenate([ks[0][0], ks[1][0]])
    order = np.lexsort((k[:, 2], k[:, 1], k[:, 0]))
    assert np.array_equal(k[order], kf)                  # union == unsharded key set, bit-exact
    w = np.concatenate([ks[0][2], ks[1][2]])[order]
    d = np.concatenate([ks[0][1], ks[1][1]])[order]
    assert np.allclose(w, wf, rtol=1e-5) and np.allclose(d, df, rtol=1e-4, atol=1e-6)
    # tracker sums: shard partials add up to the unsharded sums
    R, pn = pkg.pose_matrix(poses[3])
    for s in shards + [full]:
        s.upload_frame(frames[3])
    ref = full.accumulate(R, pn[:3])
    part = sum(s.accumulate_local(R, pn[:3]).astype(np.float64) for s in shards)
    assert part[28] == ref[28]
    assert np.abs(part[:27] - ref[:27]).max() <= 1e-4 * np.abs(ref[:27]).max()
    with pytest.raises(pkg.GsdfError):                   # collective entry point without a communicator
        shards[0].accumulate(R, pn[:3])


@pytest.mark.gpu
def test_nccl_sharded_tracking_under_torchrun():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (run by hand: gpurun --gpus 2 -- python -m torch.distributed.run ... tests/multigpu_check.py)")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
                        "--master-addr", "127.0.0.1", "--master-port", str(_free_port()),
                        str(ROOT / "tests" / "multigpu_check.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "MULTIGPU OK" in r.stdout, (r.stdout[-2000:], r.stderr[-2000:])
