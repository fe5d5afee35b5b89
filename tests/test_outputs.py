"""Output parity: marching cubes, point cloud and sdf dumps against the REFERENCE's own writers
(oracle/_ref: MapGradPixelSdfOmp.cpp:132-303, MapPixelSdfOmp.cpp:197-351, mesh/LayeredMarchingCubesNoColor.cpp) run on
the same pose-injected frames.  Two layers:
  * the reference's writers run on the map the GPU built (uploaded into the reference's hash): any difference is a
    difference of the writers alone -> same vertices in the same order, bit for bit before printing;
  * the reference's writers on the reference's own map: key sets are identical, values differ in the last bits
    (atomic accumulation order) -> same line sets, numbers within the fusion tolerances.
"""
import subprocess
from pathlib import Path

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parents[1]
BIN = ROOT / "gradient-sdf_b200" / "host" / "scan3d_b200"


def read_ply(path):
    lines = Path(path).read_text().splitlines()
    end = lines.index("end_header")
    nv = int([l for l in lines[:end] if l.startswith("element vertex")][0].split()[-1])
    nf = [l for l in lines[:end] if l.startswith("element face")]
    nf = int(nf[0].split()[-1]) if nf else 0
    v = np.array([[float(x) for x in l.split()] for l in lines[end + 1:end + 1 + nv]], np.float64).reshape(nv, -1)
    f = np.array([[int(x) for x in l.split()] for l in lines[end + 1 + nv:end + 1 + nv + nf]], np.int64).reshape(nf, 4)
    return lines[:end + 1], v, f


def build_pair(pkg, po, synth, variant, vs, trunc, W, H, nframes):
    if not po.have_ref():
        pytest.skip("oracle/_ref not built (needs /root/reference where build() runs)")
    frames, poses, K, _ = synth.make_sequence(100, W, H)
    vs, trunc = np.float32(vs), np.float32(trunc)
    nest = pkg.NormalEstimator(W, H, K, 11)
    cls = pkg.MapGradPixelSdf if variant == 0 else pkg.MapPixelSdf
    sdf = cls(vs, trunc * vs, nest, log2_capacity=16)
    ref = po.RefScan(variant, vs, trunc, W, H, K)
    for i in range(nframes):
        sdf.update(frames[i], poses[i])
        ref.update(frames[i], poses[i])
    return sdf, ref, frames, poses, K


@pytest.mark.parametrize("variant,vs,trunc", [(0, 0.02, 5), (1, 0.01, 5), (0, 0.008, 10)])
def test_marching_cubes_against_the_reference(pkg, po, synth, tmp_path, variant, vs, trunc):
    W, H = 320, 240
    sdf, ref, *_ = build_pair(pkg, po, synth, variant, vs, trunc, W, H, 4)
    tris = sdf.extract_mesh()
    assert len(tris) > 2000
    # (1) the reference's marching cubes on the GPU's map: identical vertices in identical order
    twin = po.RefScan(variant, np.float32(vs), np.float32(trunc), W, H, np.eye(3, dtype=np.float32))
    twin.upload(*sdf.download())
    assert twin.write("mesh", tmp_path / "twin.ply")
    head, v, f = read_ply(tmp_path / "twin.ply")
    assert head[2] == f"element vertex {3 * len(tris)}" and head[6] == f"element face {len(tris)}"
    assert np.array_equal(f[:, 1:], np.arange(3 * len(tris)).reshape(-1, 3)) and (f[:, 0] == 3).all()
    mine = tris.reshape(-1, 3).astype(np.float64)
    # the file holds 6 significant digits: compare what would be printed
    printed = np.array([[float("%g" % x) for x in row] for row in mine])
    assert np.array_equal(printed, v), np.abs(printed - v).max()
    # (2) the reference's own map: same cells up to the last bits of dist -> nearly the same triangle soup
    assert ref.write("mesh", tmp_path / "ref.ply")
    _, vr, fr = read_ply(tmp_path / "ref.ply")
    assert abs(len(fr) - len(tris)) <= 2e-3 * len(fr)
    # (edge points are t0 / (t0 - t1) along an edge: where the two distances nearly cancel, last-bit differences of dist move
    # the point by a visible fraction of a voxel, so the bar is "nearly every vertex has a twin within 1e-3 voxels")
    from scipy.spatial import cKDTree
    da, _ = cKDTree(vr).query(mine)
    db, _ = cKDTree(mine).query(vr)
    assert (da < 1e-3 * vs).mean() > 0.99 and (db < 1e-3 * vs).mean() > 0.99, ((da < 1e-3 * vs).mean(), (db < 1e-3 * vs).mean())
    assert np.quantile(da, 0.999) < 0.1 * vs
    # properties of the surface: every triangle lies in one cell, no degenerate triangle (LayeredMarchingCubesNoColor.cpp:676)
    ext = tris.max(axis=1) - tris.min(axis=1)
    assert ext.max() <= vs * (1 + 1e-4)
    assert (np.abs(tris[:, 0] - tris[:, 1]).max(axis=1) > 0).all() and (np.abs(tris[:, 1] - tris[:, 2]).max(axis=1) > 0).all()
    sdf.close()


def test_marching_cubes_edge_cases(pkg, synth):
    W, H = 64, 48
    K = synth.scaled_intrinsics(W, H).astype(np.float32)
    nest = pkg.NormalEstimator(W, H, K, 11)
    sdf = pkg.MapGradPixelSdf(0.02, 0.1, nest, log2_capacity=10)
    assert sdf.extract_mesh().shape == (0, 3, 3)  # empty map: the reference's savePly returns false on no vertices
    frames, poses, K2, _ = synth.make_sequence(50, W, H)
    sdf.update(frames[0], poses[0])
    n = len(sdf.extract_mesh())
    assert n > 0
    import ctypes as C
    out = np.zeros((n - 1, 9), np.float32)
    got = C.c_int64(0)
    rc = sdf.L.gsdf_extract_mesh(sdf.h, n - 1, out.ctypes.data_as(C.POINTER(C.c_float)), C.byref(got))
    assert rc == 1 and got.value == n  # GSDF_ERR_INVALID, and the count needed is reported
    shard = pkg.MapGradPixelSdf(0.02, 0.1, nest, log2_capacity=10, world_size=2, rank=0)
    with pytest.raises(pkg.GsdfError):
        shard.extract_mesh()
    sdf.close(); shard.close()


def write_raw(tmp, frames, K, poses):
    n, h, w = frames.shape
    np.savetxt(tmp / "intrinsics.txt", K.reshape(3, 3), fmt="%.9g")
    (tmp / "meta.txt").write_text(f"{w} {h} {n}\n")
    frames.astype(np.float32).tofile(tmp / "frames.f32")
    with open(tmp / "pose.txt", "w") as f:
        for i, p in enumerate(poses):
            f.write(f"{i:03d} " + " ".join(f"{x:.9g}" for x in p) + "\n")


def table(path):
    d = np.loadtxt(path, ndmin=2)
    o = np.argsort(d[:, 0])
    return d[o, 0].astype(np.int64), d[o, 1]


@pytest.mark.parametrize("stype,variant", [("grad-sdf", 0), ("base-sdf", 1)])
def test_scan3d_output_files_against_the_reference_writers(pkg, po, synth, tmp_path, stype, variant):
    """scan3d_b200 (the C++ shells over the C ABI) in ground-truth-pose mode against the reference's writers on the
    reference's map of the same frames and poses: mesh, cloud, grid info, sdf / weight / normal dumps."""
    if not po.have_ref():
        pytest.skip("oracle/_ref not built")
    W, H, vs, trunc = 320, 240, np.float32(0.01), np.float32(5)
    frames, poses, K, _ = synth.make_sequence(100, W, H)
    frames, poses = frames[:8], poses[:8]  # 8 frames: weights reach the clouds' `weight >= 5` filter
    write_raw(tmp_path, frames, K, poses)
    out = tmp_path / "gpu_"
    r = subprocess.run([str(BIN), "--input", str(tmp_path), "--results", str(out), "--scan-type", stype, "--data-type", "raw",
                        "--voxel-size", "0.01", "--trunc", "5", "--capacity-log2", "16", "--save-sdf"], capture_output=True,
                       text=True, timeout=300)
    assert r.returncode == 0 and "Could not save mesh" not in r.stdout + r.stderr, r.stderr[-1500:]
    ref = po.RefScan(variant, vs, trunc, W, H, K)
    for i in range(len(frames)):
        p = np.array([float(x) for x in (tmp_path / "pose.txt").read_text().splitlines()[i].split()[1:]], np.float32)
        ref.update(frames[i], p)
    rp = tmp_path / "ref_"
    assert ref.write("mesh", str(rp) + "gradient_sdf_mesh_final.ply") and ref.write("pc", str(rp) + "gradient_sdf_cloud_final.ply")
    assert ref.write("sdf", str(rp) + "gradient_sdf")
    g, q = str(out) + "gradient_sdf", str(rp) + "gradient_sdf"
    # same set of files: base-sdf writes no normal dumps (MapPixelSdfOmp.cpp:290-351)
    names = lambda pre: sorted(p.name[len(Path(pre).name):] for p in tmp_path.iterdir() if p.name.startswith(Path(pre).name))
    assert [n for n in names(g) if n != "_poses.txt"] == names(q)
    assert Path(g + "_grid_info.txt").read_text() == Path(q + "_grid_info.txt").read_text()
    T = float(trunc * vs)
    for suffix, tol in [("_sdf_d.txt", 1e-5 * T + 2e-6), ("_sdf_weight.txt", None)] + (
            [(f"_sdf_n{c}.txt", None) for c in range(3)] if variant == 0 else []):
        ia, va = table(g + suffix)
        ib, vb = table(q + suffix)
        assert np.array_equal(ia, ib), suffix  # same voxels at the same linear indices
        if tol is None:
            assert np.allclose(va, vb, rtol=1e-3 if "_n" in suffix else 2e-5, atol=2e-4 if "_n" in suffix else 0), suffix
        else:
            assert np.all(np.abs(va - vb) <= tol + 1e-4 * np.abs(vb)), suffix
    # clouds: same header, same points up to borderline voxels of the two filters
    ha, va, _ = read_ply(g + "_cloud_final.ply")
    hb, vb, _ = read_ply(q + "_cloud_final.ply")
    assert ha[:2] == hb[:2] and ha[3:] == hb[3:] and va.shape[1] == (6 if variant == 0 else 3)
    assert len(vb) > 200
    ka = {tuple(x) for x in np.round(va[:, :3] / (0.05 * vs)).astype(np.int64)}
    kb = {tuple(x) for x in np.round(vb[:, :3] / (0.05 * vs)).astype(np.int64)}
    assert len(ka & kb) / len(ka | kb) > 0.99
    # mesh: same header shape, nearly the same triangle count, faces index three fresh vertices each
    ma, mva, mfa = read_ply(g + "_mesh_final.ply")
    mb, mvb, mfb = read_ply(q + "_mesh_final.ply")
    assert ma[:2] == mb[:2] and ma[3:6] == mb[3:6] and ma[7:] == mb[7:]
    assert abs(len(mfa) - len(mfb)) <= 2e-3 * len(mfb) and len(mva) == 3 * len(mfa)
    # pose file: ts tx ty tz qx qy qz qw with the quaternion rebuilt from R (main_scan_3d.cpp:268-280): unit, w >= 0 branch
    pf = np.loadtxt(str(out) + "_poses.txt")
    assert pf.shape == (len(frames), 8) and np.allclose(np.linalg.norm(pf[:, 4:], axis=1), 1, atol=1e-5)
    assert np.abs(np.abs((pf[:, 4:] * poses[:, 3:]).sum(1)) - 1).max() < 1e-5 and np.abs(pf[:, 1:4] - poses[:, :3]).max() < 1e-5


def test_voxel_visibility_against_the_reference(pkg, po, synth):
    """vis_ (MapGradPixelSdfOmp.cpp:118-120): after the same pose-injected frames the (voxel, frame) pairs equal the `true`
    bits of the reference's per-voxel vectors, pair for pair, also when a frame in between integrates nothing (counter_
    still advances, MapGradPixelSdfOmp.cpp:128) and through the frame pipeline (fusion inside the tracker launch)."""
    if not po.have_ref():
        pytest.skip("oracle/_ref not built")
    W, H = 320, 240
    frames, poses, K, _ = synth.make_sequence(100, W, H)
    vs, trunc = np.float32(0.01), np.float32(5)
    nest = pkg.NormalEstimator(W, H, K, 11)
    sdf = pkg.MapGradPixelSdf(vs, trunc * vs, nest, log2_capacity=16)
    sdf.vis_enable(4_000_000)
    ref = po.RefScan(0, vs, trunc, W, H, K)
    seq = [frames[0], frames[1], np.zeros_like(frames[0]), frames[2], frames[3]]
    pos = [poses[0], poses[1], poses[1], poses[2], poses[3]]
    for f, p in zip(seq, pos):
        sdf.update(f, p)
        ref.update(f, p)
    k, fr, nf = sdf.get_vis()
    kr, frr, size = ref.get_vis()
    assert nf == 5 and size == 5
    assert k.shape == kr.shape and np.array_equal(k, kr) and np.array_equal(fr, frr)
    assert 2 not in set(fr.tolist()) and set(fr.tolist()) == {0, 1, 3, 4}
    # every voxel of the map has at least one entry, and only voxels of the map have entries
    assert {tuple(x) for x in k} == {tuple(x) for x in sdf.download()[0]}
    # the frame pipeline: frame 0 fused at identity, then tracked frames fused iff converged -> entries only at fused frames
    sdf.reset()
    ident = np.array([0, 0, 0, 0, 0, 0, 1], np.float32)
    sdf.scan_set_pose(ident)
    sdf.scan_submit(depth_host=frames[0], track=False)
    sdf.scan_collect()
    res = sdf.scan_run([frames[i].ctypes.data for i in range(1, 7)], on_device=False)
    k2, fr2, nf2 = sdf.get_vis()
    assert nf2 == 1 + sum(r.fused for r in res)
    assert set(fr2.tolist()) == set(range(nf2)) and {tuple(x) for x in k2} == {tuple(x) for x in sdf.download()[0]}
    # log too small: reported, never silently truncated
    small = pkg.MapGradPixelSdf(vs, trunc * vs, nest, log2_capacity=16)
    small.vis_enable(1000)
    small.update(frames[0], poses[0])
    with pytest.raises(pkg.GsdfError) as e:
        small.get_vis()
    assert e.value.status == 3
    sdf.close(); small.close()
