"""The C++ host mirror (gradient-sdf_b200/host): forwarding shells with the reference's class names over the C ABI and
the Scan3D driver that reproduces main_scan_3d.cpp's loop and output files."""
import subprocess
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parents[1]
BIN = ROOT / "gradient-sdf_b200" / "host" / "scan3d_b200"


def write_raw(tmp, frames, K, poses=None):
    n, h, w = frames.shape
    np.savetxt(tmp / "intrinsics.txt", K.reshape(3, 3), fmt="%.9g")
    (tmp / "meta.txt").write_text(f"{w} {h} {n}\n")
    frames.astype(np.float32).tofile(tmp / "frames.f32")
    if poses is not None:
        with open(tmp / "pose.txt", "w") as f:
            for i, p in enumerate(poses):
                f.write(f"{i:03d} " + " ".join(f"{x:.9g}" for x in p) + "\n")


def run(args):
    return subprocess.run([str(BIN)] + args, capture_output=True, text=True, timeout=300)


def test_driver_builds_and_rejects_like_the_reference():
    assert BIN.exists(), "run __graft_entry__.build()"
    r = run(["--scan-type", "voxel-gp", "--data-type", "raw"])
    assert r.returncode == 1 and "Your specified scan type is not supported (yet)." in r.stderr  # main_scan_3d.cpp:112
    r = run(["--scan-type", "grad-sdf", "--data-type", "kitti"])
    assert r.returncode == 1 and "Your specified dataset type is not supported (yet)." in r.stderr  # main_scan_3d.cpp:137
    r = run(["--scan-type", "grad-sdf", "--data-type", "tum", "--input", "/nonexistent"])
    assert r.returncode == 1 and "No intrinsics file found" in r.stderr
    r = run(["--scan-type", "grad-sdf", "--data-type", "raw", "--input", "/nonexistent"])
    assert r.returncode == 1 and "No intrinsics file found" in r.stderr  # main_scan_3d.cpp:163


@pytest.mark.gpu
@pytest.mark.parametrize("stype", ["base-sdf", "grad-sdf"])
def test_scan3d_driver_matches_python_binding(pkg, synth, tmp_path, stype):
    W, H = 320, 240
    noise = stype == "base-sdf"  # grad-sdf: the regime in which its tracker converges with margin
    frames, poses, K, _ = synth.make_sequence(400 if noise else 800, W, H, noise=noise)
    frames = frames[:8]
    write_raw(tmp_path, frames, K)
    out = tmp_path / "res"
    # reference loop through the Python binding
    nest = pkg.NormalEstimator(W, H, K, 11)
    cls = pkg.MapPixelSdf if stype == "base-sdf" else pkg.MapGradPixelSdf
    sdf = cls(np.float32(0.01), np.float32(5) * np.float32(0.01), nest, log2_capacity=15)
    sdf.setup(frames[0])
    opt = pkg.RigidPointOptimizer(sdf)
    ref = [opt.pose()]
    for i in range(1, len(frames)):
        if opt.optimize(frames[i]):
            sdf.update(None, opt.pose(), stored=True)
        ref.append(opt.pose())
    ref = np.stack(ref)
    nv_ref = sdf.voxel_count()[0]
    for extra in ([], ["--pipelined"]):
        r = run(["--input", str(tmp_path), "--results", str(out), "--scan-type", stype, "--data-type", "raw",
                 "--voxel-size", "0.01", "--trunc", "5", "--capacity-log2", "15", "--save-sdf"] + extra)
        assert r.returncode == 0, r.stderr[-2000:]
        assert "points integrated into tSDF volume!" in r.stdout and "Convergence after" in r.stdout
        got = np.loadtxt(str(out) + "_poses.txt")[:, 1:]
        assert got.shape == ref.shape
        assert np.abs(got - ref).max() < (5e-5 if noise else 5e-4)
        # output files of MapGradPixelSdfOmp.cpp:186-303
        info = Path(str(out) + "gradient_sdf_grid_info.txt").read_text().splitlines()
        assert info[0].startswith("voxel size: 0.01") and info[1].startswith("voxel dim:")
        d = np.loadtxt(str(out) + "gradient_sdf_sdf_d.txt")
        wt = np.loadtxt(str(out) + "gradient_sdf_sdf_weight.txt")
        assert abs(len(d) - nv_ref) <= 0.001 * nv_ref and len(wt) == len(d)
        assert np.abs(d[:, 1]).max() <= 0.05 + 1e-6 and wt[:, 1].min() > 0
        dims = [int(x) for x in info[1].split(":")[1].split()]
        assert d[:, 0].max() < dims[0] * dims[1] * dims[2]
        ply = Path(str(out) + "gradient_sdf_cloud_final.ply").read_text().splitlines()
        nvert = int(ply[2].split()[-1])
        # grad-sdf: xyz + normal (10 header lines, MapGradPixelSdfOmp.cpp:205-219); base-sdf: xyz only (7, MapPixelSdfOmp.cpp:272-281)
        assert ply[0] == "ply" and (nvert > 100 if stype == "grad-sdf" else nvert >= 0)
        assert len(ply) == (10 if stype == "grad-sdf" else 7) + nvert
        assert Path(str(out) + "gradient_sdf_sdf_n0.txt").exists() == (stype == "grad-sdf")
        mesh = Path(str(out) + "gradient_sdf_mesh_final.ply").read_text().splitlines()
        assert mesh[0] == "ply" and int(mesh[2].split()[-1]) == 3 * int(mesh[6].split()[-1]) > 0


@pytest.mark.gpu
def test_scan3d_driver_ground_truth_poses(pkg, po, synth, tmp_path):
    """With a pose file every frame is fused at its ground-truth pose without tracking (main_scan_3d.cpp:242,250-254):
    the voxel count must equal the oracle's on the same frames and poses."""
    W, H = 160, 120
    frames, poses, K, _ = synth.make_sequence(100, W, H)
    frames, poses = frames[:4], poses[:4]
    write_raw(tmp_path, frames, K, poses)
    out = tmp_path / "gt"
    r = run(["--input", str(tmp_path), "--results", str(out), "--scan-type", "grad-sdf", "--data-type", "raw",
             "--voxel-size", "0.02", "--trunc", "5", "--capacity-log2", "14", "--save-sdf"])
    assert r.returncode == 0, r.stderr[-2000:]
    assert "4 GT poses are loaded!" in r.stdout
    orc = po.Oracle(0, np.float32(0.02), np.float32(5), W, H, K)
    for i in range(4):
        # the pose file is text: parse it back exactly as the driver does
        p = np.array([float(x) for x in (tmp_path / "pose.txt").read_text().splitlines()[i].split()[1:]], np.float32)
        orc.update(frames[i], p)
    d = np.loadtxt(str(out) + "gradient_sdf_sdf_d.txt")
    assert len(d) == orc.voxel_count()


@pytest.mark.gpu
def test_scan3d_driver_reads_the_reference_dataset_conventions(synth, tmp_path):
    """--data-type synth (16-bit PNGs, depth/%03d.png, SynthLoader.h) must give the poses of --data-type raw on the same
    frames, and label them with the loader's depth timestamps (main_scan_3d.cpp:276)."""
    cv2 = pytest.importorskip("cv2")
    W, H = 320, 240
    frames, poses, K, _ = synth.make_sequence(400, W, H)
    frames = frames[:6]
    raw = tmp_path / "raw"
    png = tmp_path / "png"
    raw.mkdir()
    (png / "depth").mkdir(parents=True)
    (png / "rgb").mkdir()
    write_raw(raw, frames, K)
    np.savetxt(png / "intrinsics.txt", K.reshape(3, 3), fmt="%.9g")
    for i, f in enumerate(frames):
        u16 = np.round(f.astype(np.float64) * 1000).astype(np.uint16)
        assert cv2.imwrite(str(png / "depth" / f"{i + 1:03d}.png"), u16)
        assert cv2.imwrite(str(png / "rgb" / f"{i + 1:03d}.png"), np.zeros((H, W, 3), np.uint8))
    res = {}
    for kind, d in (("raw", raw), ("synth", png)):
        out = tmp_path / f"res_{kind}"
        r = run(["--input", str(d), "--results", str(out), "--scan-type", "base-sdf", "--data-type", kind,
                 "--voxel-size", "0.01", "--trunc", "5", "--capacity-log2", "15", "--last", "5"])
        assert r.returncode == 0, r.stderr[-2000:]
        res[kind] = [l.split() for l in Path(str(out) + "_poses.txt").read_text().splitlines()]
    # the same dataset with frames decoded ahead on host threads and fed to the asynchronous pipeline from pinned memory
    out = tmp_path / "res_prefetch"
    r = run(["--input", str(png), "--results", str(out), "--scan-type", "base-sdf", "--data-type", "synth", "--voxel-size",
             "0.01", "--trunc", "5", "--capacity-log2", "15", "--last", "5", "--pipelined", "--threads", "4"])
    assert r.returncode == 0 and "prefetched loop: 5 frames" in r.stdout, r.stderr[-2000:]
    res["prefetch"] = [l.split() for l in Path(str(out) + "_poses.txt").read_text().splitlines()]
    assert [l[0] for l in res["prefetch"]] == [f"{i + 1:03d}" for i in range(6)]
    c = np.array([[float(x) for x in l[1:]] for l in res["prefetch"]])
    assert [l[0] for l in res["synth"]] == [f"{i + 1:03d}" for i in range(6)]
    a = np.array([[float(x) for x in l[1:]] for l in res["raw"]])
    b = np.array([[float(x) for x in l[1:]] for l in res["synth"]])
    assert a.shape == b.shape == (6, 7) and np.abs(a - b).max() < 5e-5
    assert c.shape == (6, 7) and np.abs(a - c).max() < 5e-5


REF_MAIN = ROOT / "gradient-sdf_b200" / "host" / "scan3d_ref_main"


@pytest.mark.gpu
@pytest.mark.parametrize("stype", ["grad-sdf", "base-sdf"])
def test_the_references_own_driver_runs_on_the_library(synth, tmp_path, stype):
    """cpp/depth_scanning/src/main_scan_3d.cpp of the reference, compiled UNMODIFIED (with its own loaders, mat.h, Timer.h,
    Eigen, Sophus, CLI11) against the drop-in headers in gradient-sdf_b200/dropin and linked with libgsdf_b200.so — built by
    build() where /root/reference exists, shipped as a binary — on a synth-format PNG dataset (img_loader/SynthLoader.h), next
    to this repo's own driver on the same files: same pose file (labels and numbers), same clouds, same sdf dumps, a mesh."""
    cv2 = pytest.importorskip("cv2")
    if not REF_MAIN.exists():
        pytest.skip("host/scan3d_ref_main not built (needs /root/reference where build() runs)")
    W, H = 320, 240
    noise = stype == "base-sdf"
    frames, poses, K, _ = synth.make_sequence(400 if noise else 800, W, H, noise=noise)
    frames = frames[:7]
    data = tmp_path / "data"
    (data / "depth").mkdir(parents=True)
    (data / "rgb").mkdir()
    np.savetxt(data / "intrinsics.txt", K.reshape(3, 3), fmt="%.9g")
    for i, f in enumerate(frames):
        assert cv2.imwrite(str(data / "depth" / f"{i + 1:03d}.png"), np.round(f.astype(np.float64) * 1000).astype(np.uint16))
        assert cv2.imwrite(str(data / "rgb" / f"{i + 1:03d}.png"), np.zeros((H, W, 3), np.uint8))
    outs = {}
    for name, exe, extra in (("ref_main", REF_MAIN, []), ("ours", BIN, ["--capacity-log2", "17"])):
        out = tmp_path / f"{name}_"
        r = subprocess.run([str(exe), "--input", str(data) + "/", "--results", str(out), "--scan-type", stype, "--data-type", "synth",
                            "--voxel-size", "0.01", "--trunc", "5", "--save-sdf"] + extra, capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, (name, r.stderr[-1500:])
        assert "points integrated into tSDF volume!" in r.stdout and "Could not save" not in r.stdout + r.stderr, name
        outs[name] = str(out)
    a = [l.split() for l in Path(outs["ref_main"] + "_poses.txt").read_text().splitlines()]
    b = [l.split() for l in Path(outs["ours"] + "_poses.txt").read_text().splitlines()]
    assert [l[0] for l in a] == [l[0] for l in b] == [f"{i + 1:03d}" for i in range(len(frames))]  # depth_timestamp()
    pa, pb = np.array([[float(x) for x in l[1:]] for l in a]), np.array([[float(x) for x in l[1:]] for l in b])
    assert pa.shape == pb.shape == (len(frames), 7) and np.abs(pa - pb).max() < (5e-5 if noise else 5e-4)
    for suffix in ["gradient_sdf_grid_info.txt"]:
        assert Path(outs["ref_main"] + suffix).read_text() == Path(outs["ours"] + suffix).read_text()
    da, db = np.loadtxt(outs["ref_main"] + "gradient_sdf_sdf_d.txt"), np.loadtxt(outs["ours"] + "gradient_sdf_sdf_d.txt")
    assert abs(len(da) - len(db)) <= 1e-3 * len(db)
    for suffix in ["gradient_sdf_cloud_final.ply", "gradient_sdf_mesh_final.ply"]:
        ha = Path(outs["ref_main"] + suffix).read_text().splitlines()
        hb = Path(outs["ours"] + suffix).read_text().splitlines()
        assert ha[:2] == hb[:2] and abs(int(ha[2].split()[-1]) - int(hb[2].split()[-1])) <= 0.01 * int(hb[2].split()[-1]) + 2
    assert Path(outs["ref_main"] + "gradient_sdf_sdf_n0.txt").exists() == (stype == "grad-sdf")


def test_the_references_own_driver_builds_against_the_drop_in_headers():
    """CPU part: where /root/reference exists, build() has compiled the reference's unmodified main_scan_3d.cpp against
    gradient-sdf_b200/dropin; its own argument checks (main_scan_3d.cpp:105-139) answer before any GPU call."""
    if not Path("/root/reference/cpp/depth_scanning/src/main_scan_3d.cpp").exists() and not REF_MAIN.exists():
        pytest.skip("no reference tree and no prebuilt binary")
    assert REF_MAIN.exists(), "run __graft_entry__.build()"
    r = subprocess.run([str(REF_MAIN), "--scan-type", "voxel-gp", "--data-type", "synth"], capture_output=True, text=True, timeout=60)
    assert r.returncode == 1 and "Your specified scan type is not supported (yet)." in r.stderr
    r = subprocess.run([str(REF_MAIN), "--scan-type", "grad-sdf", "--data-type", "synth", "--input", "/nonexistent/"],
                       capture_output=True, text=True, timeout=60)
    assert r.returncode == 1 and "No intrinsics file found" in r.stderr
