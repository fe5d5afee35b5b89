"""Dataset loaders of the host mirror (gradient-sdf_b200/host/img_loader.h) against the
conventions of the reference's img_loader/*.h.  The expected depth values come from OpenCV itself (python-cv2, the
library the reference decodes with): cv2.imwrite writes the 16-bit PNGs, cv2.imread(..., IMREAD_ANYDEPTH) reads them
back, and `convertTo(depth, CV_32FC1, unit_)` (ImageLoader.h:172) is float(sample) * float(unit).  CPU only."""
import subprocess
from pathlib import Path

import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")
ROOT = Path(__file__).resolve().parents[1]
PROBE = ROOT / "gradient-sdf_b200" / "host" / "loader_probe"
W, H, N = 96, 72, 4


def _frames(synth, unit):
    frames, poses, K, _ = synth.make_sequence(40, W, H, unit=unit)
    return frames[:N], K


def _u16(frame, unit):
    return np.clip(np.round(frame.astype(np.float64) / unit), 0, 65535).astype(np.uint16)


def _write_depth(path, u16, compression=None):
    params = [] if compression is None else [cv2.IMWRITE_PNG_COMPRESSION, compression]
    assert cv2.imwrite(str(path), u16, params)


def _probe(kind, d, tmp_path, threads=0, slots=6):
    out = tmp_path / f"probe_{kind}_{threads}"
    extra = [str(threads), str(slots)] if threads else []
    r = subprocess.run([str(PROBE), kind, str(d) + "/", str(out)] + extra, capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l.split() for l in r.stdout.strip().splitlines()]
    frames = [l for l in lines if l[0] != "K"]
    k = [l for l in lines if l[0] == "K"]
    data = np.fromfile(str(out) + ".f32", np.float32)
    return frames, (np.array(k[0][1:], np.float32).reshape(3, 3) if k else None), data


def _expected(files, unit):
    out = []
    for f in files:
        img = cv2.imread(str(f), cv2.IMREAD_ANYDEPTH | cv2.IMREAD_ANYCOLOR)
        assert img is not None and img.dtype == np.uint16 and img.ndim == 2
        out.append(img.astype(np.float32) * np.float32(unit))  # convertTo(CV_32FC1, unit_): float product
    return np.stack(out)


def _write_k(d, K):
    np.savetxt(d / "intrinsics.txt", K, fmt="%.6f")


@pytest.fixture(scope="module")
def probe_built():
    r = subprocess.run(["make", "-C", str(ROOT / "gradient-sdf_b200"), str(PROBE)], capture_output=True, text=True)
    assert r.returncode == 0 and PROBE.exists(), r.stdout[-2000:] + r.stderr[-2000:]


def test_synth_convention(probe_built, synth, tmp_path):
    unit = 1.0 / 1000
    frames, K = _frames(synth, unit)
    d = tmp_path / "synth"
    (d / "depth").mkdir(parents=True)
    (d / "rgb").mkdir()
    files = []
    for i in range(N):  # SynthLoader.h:61-82: depth/%03d.png and rgb/%03d.png, counting from 1
        f = d / "depth" / f"{i + 1:03d}.png"
        _write_depth(f, _u16(frames[i], unit), compression=[0, 3, 9, 1][i])
        cv2.imwrite(str(d / "rgb" / f"{i + 1:03d}.png"), np.full((H, W, 3), 40 * i, np.uint8))
        files.append(f)
    _write_k(d, K)
    got, k, data = _probe("synth", d, tmp_path)
    assert [g[0] for g in got] == ["001", "002", "003", "004"] and all(g[0] == g[1] for g in got)
    assert all((int(g[2]), int(g[3])) == (W, H) for g in got)
    exp = _expected(files, np.float32(unit))
    assert np.array_equal(data.reshape(N, H, W).view(np.uint32), exp.view(np.uint32))   # bit-exact depth
    assert np.array_equal(data.reshape(N, H, W), frames)                                # = the generator's frames
    assert np.allclose(k, K, atol=1e-4)


def test_tum_convention(probe_built, synth, tmp_path):
    unit = 1.0 / 5000  # TumrgbdLoader.h:64-68
    frames, K = _frames(synth, unit)
    d = tmp_path / "tum"
    (d / "depth").mkdir(parents=True)
    (d / "rgb").mkdir()
    files, lines = [], ["# comment line", "# another"]
    for i in range(N):
        ts_d, ts_c = f"1305031102.{160407 + 33000 * i}", f"1305031102.{175304 + 33000 * i}"
        f = d / "depth" / f"{ts_d}.png"
        _write_depth(f, _u16(frames[i], unit))
        cv2.imwrite(str(d / "rgb" / f"{ts_c}.png"), np.zeros((H, W, 3), np.uint8))
        lines.append(f"{ts_c} rgb/{ts_c}.png {ts_d} depth/{ts_d}.png")  # TumrgbdLoader.h:88: rgb first
        files.append(f)
    (d / "associated.txt").write_text("\n".join(lines) + "\n")
    _write_k(d, K)
    got, k, data = _probe("tum", d, tmp_path)
    assert len(got) == N and got[1][0] == "1305031102.193407" and got[1][1] == "1305031102.208304"
    exp = _expected(files, np.float32(unit))
    assert np.array_equal(data.reshape(N, H, W).view(np.uint32), exp.view(np.uint32))
    assert np.array_equal(data.reshape(N, H, W), frames)


def test_printed3d_and_redwood_conventions(probe_built, synth, tmp_path):
    unit = 1.0 / 1000
    frames, K = _frames(synth, unit)
    d = tmp_path / "printed"
    d.mkdir()
    files = []
    for i in range(N):  # Printed3dLoader.h:61-82: depth_%06d.png / color_%06d.png from 0
        f = d / f"depth_{i:06d}.png"
        _write_depth(f, _u16(frames[i], unit))
        cv2.imwrite(str(d / f"color_{i:06d}.png"), np.zeros((H, W, 3), np.uint8))
        files.append(f)
    got, k, data = _probe("printed", d, tmp_path)
    assert [g[0] for g in got] == [f"{i:06d}" for i in range(N)] and k is None  # no intrinsics.txt here
    assert np.array_equal(data.reshape(N, H, W), frames)
    # a missing colour image ends the sequence there, as in the reference (load_color fails -> load_next false)
    (d / "color_000002.png").unlink()
    got, _, data = _probe("printed", d, tmp_path)
    assert len(got) == 2
    # RedwoodLoader.h:57-122: sorted *.png of depth/ paired with sorted *.jpg of rgb/
    r = tmp_path / "rw"
    (r / "depth").mkdir(parents=True)
    (r / "rgb").mkdir()
    order = [3, 0, 2, 1]
    for j, i in enumerate(order):  # written out of order: the loader sorts by name
        _write_depth(r / "depth" / f"{i:07d}.png", _u16(frames[i], unit))
        cv2.imwrite(str(r / "rgb" / f"{i:07d}.jpg"), np.zeros((H, W, 3), np.uint8))
    (r / "depth" / "notes.txt").write_text("not an image")
    got, _, data = _probe("rw", r, tmp_path)
    assert [g[0] for g in got] == [f"{i:07d}.png" for i in range(N)]
    assert [g[1] for g in got] == [f"{i:07d}.jpg" for i in range(N)]
    assert np.array_equal(data.reshape(N, H, W), frames)


def test_png_decoder_rejects_what_is_not_a_depth_map(probe_built, tmp_path):
    d = tmp_path / "bad"
    (d / "depth").mkdir(parents=True)
    (d / "rgb").mkdir()
    cv2.imwrite(str(d / "rgb" / "001.png"), np.zeros((H, W, 3), np.uint8))
    cv2.imwrite(str(d / "depth" / "001.png"), np.zeros((H, W, 3), np.uint8))  # 3-channel "depth"
    r = subprocess.run([str(PROBE), "synth", str(d) + "/", str(tmp_path / "o")], capture_output=True, text=True)
    assert r.returncode == 0 and r.stdout.strip() == "" and "greyscale" in r.stderr
    # 8-bit greyscale is accepted (samples are not rescaled, like IMREAD_ANYDEPTH)
    g8 = (np.arange(H * W, dtype=np.uint32).reshape(H, W) % 251).astype(np.uint8)
    cv2.imwrite(str(d / "depth" / "001.png"), g8)
    r = subprocess.run([str(PROBE), "synth", str(d) + "/", str(tmp_path / "o")], capture_output=True, text=True)
    data = np.fromfile(str(tmp_path / "o.f32"), np.float32).reshape(H, W)
    assert np.array_equal(data, g8.astype(np.float32) * np.float32(1.0 / 1000))
    # a truncated file is an error, not garbage
    raw = (d / "depth" / "001.png").read_bytes()
    (d / "depth" / "001.png").write_bytes(raw[: len(raw) // 2])
    r = subprocess.run([str(PROBE), "synth", str(d) + "/", str(tmp_path / "o")], capture_output=True, text=True)
    assert r.stdout.strip() == "" and "Error: empty depth image" in r.stderr


def test_parallel_prefetch_delivers_the_same_frames_in_order(probe_built, synth, tmp_path):
    """FramePrefetcher (host/prefetch.h): frames enumerated in order, decoded by a thread pool, delivered in order and
    held by the consumer while later ones are decoded — identical to the sequential loader; a frame that cannot be
    loaded ends the sequence at that frame."""
    unit = 1.0 / 1000
    frames, poses, K, _ = synth.make_sequence(60, W, H, unit=unit)
    frames = frames[:23]
    d = tmp_path / "printed"
    d.mkdir()
    for i, fr in enumerate(frames):
        _write_depth(d / f"depth_{i:06d}.png", _u16(fr, unit), compression=[1, 6, 9][i % 3])
        (d / f"color_{i:06d}.png").write_bytes(b"x")
    seq, _, data_seq = _probe("printed", d, tmp_path)
    for threads, slots in ((1, 2), (4, 6), (8, 12)):
        got, _, data = _probe("printed", d, tmp_path, threads=threads, slots=slots)
        assert got == seq and len(got) == len(frames)
        assert np.array_equal(data, data_seq)
    assert np.array_equal(data_seq.reshape(len(frames), H, W), frames)
    (d / "color_000015.png").unlink()
    got, _, data = _probe("printed", d, tmp_path, threads=8, slots=12)
    assert [g[0] for g in got] == [f"{i:06d}" for i in range(15)]
