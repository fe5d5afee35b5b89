"""Run under torchrun with N >= 2 GPUs: the NCCL-sharded pipeline must reproduce the single-GPU run — identical
converged flags, poses within 1e-5, and the union of the ranks' voxel key sets equal to the single-GPU key set.
(Not a pytest file: tests/test_sharding.py launches it when 2 GPUs are present.)"""
import ctypes as C
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parents[1]
sys.path.insert(0, str(ROOT))
import gsdf_loader  # noqa: E402

pkg = gsdf_loader.load()
import importlib  # noqa: E402

synth = importlib.import_module("gradient_sdf_b200.synth")
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
W, H = 320, 240
frames, poses, K, _ = synth.make_sequence(800, W, H, noise=False)
frames, poses = frames[:8], poses[:8]
dev_frames = torch.from_numpy(frames).cuda()
nest = pkg.NormalEstimator(W, H, K, 11)
vs, T = np.float32(0.01), np.float32(0.05)
ident = np.array([0, 0, 0, 0, 0, 0, 1], np.float32)


def run(sdf):
    sdf.scan_set_pose(ident)
    sdf.scan_set_tracker(25, 1e-3, 1.0)
    out = []
    sdf.scan_submit(depth_host=frames[0], track=False)
    out.append(sdf.scan_collect())
    for i in range(1, len(frames)):
        sdf.scan_submit(depth_host=frames[i], track=True)
        out.append(sdf.scan_collect())
    return out


def run_loop(sdf):
    """the same frames through gsdf_scan_run (several frames per tracker launch, device-resident frames)"""
    sdf.scan_set_pose(ident)
    sdf.scan_set_tracker(25, 1e-3, 1.0)
    sdf.scan_submit(depth_dev=dev_frames[0].data_ptr(), track=False)
    out = [sdf.scan_collect()]
    return out + sdf.scan_run([dev_frames[i].data_ptr() for i in range(1, len(frames))], on_device=True)


def comm_id():
    idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        buf = (C.c_uint8 * 128)()
        assert pkg.lib().gsdf_comm_unique_id(buf) == 0
        idt.copy_(torch.tensor(list(buf), dtype=torch.uint8))
    dist.broadcast(idt, 0)
    return bytes(idt.cpu().tolist())


def check_variant(cls, tag):
    single = cls(vs, T, nest, log2_capacity=16, device=local)
    ref = run(single)
    assert sum(r.converged for r in ref[1:]) >= len(ref) - 2, tag
    sharded = cls(vs, T, nest, log2_capacity=16, device=local, world_size=world, rank=rank)
    sharded.comm_init(comm_id())

    def check(got, how):
        for a, b in zip(got, ref):
            assert a.converged == b.converged and a.iters == b.iters, (tag, how, a.converged, b.converged, a.iters, b.iters)
            assert np.abs(np.array(a.pose7[:]) - np.array(b.pose7[:])).max() < 5e-4, (tag, how)
        # every rank holds bit-identical poses (same sums in the same order on every rank)
        mine = torch.tensor([list(r.pose7) for r in got], dtype=torch.float32, device="cuda")
        allp = [torch.zeros_like(mine) for _ in range(world)]
        dist.all_gather(allp, mine)
        assert all(torch.equal(allp[0], p) for p in allp), (tag, how)

    got = run(sharded)           # NCCL all-reduce per pass
    check(got, "nccl")
    # NVLink peer exchange inside the cooperative tracker launch
    hd = torch.tensor(list(sharded.p2p_export()), dtype=torch.uint8, device="cuda")
    allh = [torch.zeros_like(hd) for _ in range(world)]
    dist.all_gather(allh, hd)
    sharded.p2p_connect(bytes(torch.cat(allh).cpu().tolist()))
    sharded.reset()
    dist.barrier()
    got = run(sharded)
    check(got, "p2p")
    sharded.reset()
    dist.barrier()
    got = run_loop(sharded)
    check(got, "p2p, frame loop in one call")
    upd = torch.tensor([sum(r.updates for r in got)], dtype=torch.int64, device="cuda")
    dist.all_reduce(upd)
    assert abs(int(upd) - sum(r.updates for r in ref)) <= 1e-3 * sum(r.updates for r in ref), tag
    k = sharded.download()[0]
    assert (pkg.shard_owner(k, world) == rank).all(), tag
    n = torch.tensor([len(k)], dtype=torch.int64, device="cuda")
    dist.all_reduce(n)
    nref = single.voxel_count()[0]
    assert abs(int(n) - nref) <= 1e-3 * nref, (tag, int(n), nref)
    dist.barrier()
    # pose-injected fusion: the union of the ranks' key sets must EQUAL the single-GPU key set and the CPU oracle's, key by
    # key (free-running maps differ at the truncation boundary because the poses differ in the last bits)
    single.reset()
    sharded.reset()
    po = gsdf_loader.load_oracle()
    orc = po.Oracle(cls.variant, vs, np.float32(5), W, H, K)
    for i in range(4):
        single.update(frames[i], poses[i])
        sharded.update(frames[i], poses[i])
        if rank == 0:
            orc.update(frames[i], poses[i])
    mine = sharded.download()[0]
    parts = [None] * world
    dist.all_gather_object(parts, mine)
    union = np.concatenate(parts)
    union = union[np.lexsort((union[:, 2], union[:, 1], union[:, 0]))]
    want = single.download()[0]
    assert union.shape == want.shape and np.array_equal(union, want), (tag, "key set of the shards' union differs from one GPU's")
    if rank == 0:
        assert np.array_equal(union, orc.download()[0]), (tag, "key set of the shards' union differs from the oracle's")
    dist.barrier()
    return int(n), nref, len(union)


res = {tag: check_variant(cls, tag) for tag, cls in (("grad", pkg.MapGradPixelSdf), ("base", pkg.MapPixelSdf))}
if rank == 0:
    print("MULTIGPU OK", world, "ranks; voxels (free-running sharded union, free-running single GPU, pose-injected union == single == oracle):", res)
dist.destroy_process_group()
