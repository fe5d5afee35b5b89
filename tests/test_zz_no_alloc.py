"""gsdf_accumulate_noalloc: the tracker's normal-equation sums for callers that are not allowed to allocate.

Some hosts of this library cannot call cudaMalloc at the moment they need the tracker: the device heap already
belongs to another allocator (a framework's caching pool has taken every free byte), or the call sits in a
real-time scanning loop where cudaMalloc's device-wide synchronisation would stall the frame.  For them the
library carries its reduction scratch as a static device array and this entry point uses only that, the handle's
map and frame, and the caller's own output buffer.

The test calls it through the C ABI and checks it against the regular entry point (gsdf_accumulate) AND against the
CPU oracle's sums for the same map, frame and pose, twice in a row on different poses, because all calls share the one
static array and it keeps whatever the last call left there.
This file sorts last on purpose: if it fails under `pytest -x`, every other test has already run.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_noalloc_matches_regular_entry_point_and_oracle_twice(pkg, po, synth):
    import torch

    W, H = 320, 240
    frames, poses, K, _ = synth.make_sequence(200, W, H)
    nest = pkg.NormalEstimator(W, H, K, 11)
    for cls in (pkg.MapGradPixelSdf, pkg.MapPixelSdf):
        sdf = cls(np.float32(0.01), np.float32(0.05), nest, log2_capacity=15)
        orc = po.Oracle(0 if cls is pkg.MapGradPixelSdf else 1, np.float32(0.01), np.float32(5.0), W, H, K)
        for i in range(3):
            sdf.update(frames[i], poses[i])
            orc.update(frames[i], poses[i])
        sdf.upload_frame(frames[3])
        out_dev = torch.zeros(32, dtype=torch.float32, device="cuda")  # caller-owned output, allocated up front
        torch.cuda.synchronize()
        for call, (pose, sampling) in enumerate([(poses[3], 1), (poses[2], 2)]):  # two calls in a row
            R, pn = pkg.pose_matrix(pose)
            ref = sdf.accumulate(R, pn[:3], sampling)
            sdf.accumulate_noalloc(R, pn[:3], out_dev.data_ptr(), sampling)
            sdf.synchronize()
            got = out_dev[:29].cpu().numpy()
            assert got[28] == ref[28] and got[28] > 500, (call, got[28], ref[28])  # contributing pixels: exact
            # same tolerance as the oracle comparison of these sums (tests/test_gpu_parity.py): 1e-4 of the max norm
            assert np.abs(got[:6] - ref[:6]).max() <= 1e-4 * np.abs(ref[:6]).max(), call
            assert np.abs(got[6:27] - ref[6:27]).max() <= 1e-4 * np.abs(ref[6:27]).max(), call
            assert abs(got[27] - ref[27]) <= 1e-4 * abs(ref[27]), call
            # and against the CPU oracle (RigidPointOptimizerOmp.cpp:69-91 restated), same bars
            cpu = orc.accumulate(frames[3], R, pn[:3], sampling)
            assert got[28] == cpu[28], (call, got[28], cpu[28])
            assert np.abs(got[:6] - cpu[:6]).max() <= 1e-4 * np.abs(cpu[:6]).max(), call
            assert np.abs(got[6:27] - cpu[6:27]).max() <= 1e-4 * np.abs(cpu[6:27]).max(), call
            assert abs(got[27] - cpu[27]) <= 1e-4 * abs(cpu[27]), call
        # refused while pipeline frames are pending, exactly like gsdf_accumulate
        sdf.scan_submit(depth_host=frames[4], track=True)
        with pytest.raises(pkg.GsdfError) as e:
            sdf.accumulate_noalloc(R, pn[:3], out_dev.data_ptr(), 1)
        assert e.value.status == 1
        sdf.scan_collect()
        sdf.close()


def test_noalloc_rejects_frames_beyond_the_static_scratch(pkg, synth):
    import torch

    # 1536 x 1024 = 1,572,864 pixels > 1,048,576: refused at sampling 1 with a clear message, accepted at 2
    W, H = 1536, 1024
    K = synth.scaled_intrinsics(W, H).astype(np.float32)
    nest = pkg.NormalEstimator(W, H, K, 11)
    sdf = pkg.MapGradPixelSdf(np.float32(0.02), np.float32(0.1), nest, log2_capacity=10)
    sdf.upload_frame(np.zeros((H, W), np.float32))
    out_dev = torch.zeros(32, dtype=torch.float32, device="cuda")
    with pytest.raises(pkg.GsdfError) as e:
        sdf.accumulate_noalloc(np.eye(3), np.zeros(3), out_dev.data_ptr(), 1)
    assert e.value.status == 1 and "static scratch limit" in str(e.value)
    sdf.accumulate_noalloc(np.eye(3), np.zeros(3), out_dev.data_ptr(), 2)
    sdf.synchronize()
    assert not out_dev.cpu().numpy().any()  # empty frame: all sums zero
    sdf.close()
