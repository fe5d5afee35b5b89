"""gradient-sdf_b200 — B200-native Scan3D track+fuse hot path (ctypes host binding over include/gsdf_b200.h).

The directory name is not a Python identifier; load it with `gsdf_loader.load()` at the repo root (tests,
bench.py and __graft_entry__.py do) — it registers this package as `gradient_sdf_b200`.

The classes mirror the reference's C++ surface for this path:
    MapGradPixelSdf / MapPixelSdf   sdf_tracker/MapGradPixelSdf.h, MapPixelSdf.h  (update, setup, tsdf, weights)
    RigidPointOptimizer             sdf_tracker/RigidOptimizer.h:90-110           (optimize, pose, set_pose, setters)
    NormalEstimator                 normals/NormalEstimator.h:158-210             (compute, x0/y0/n_sq_inv planes)
The C++ forwarding shells with the same names are in host/.  Everything computes on the GPU through
libgsdf_b200.so; if the library or a CUDA device is missing the constructors raise (no CPU fallback).
"""
from __future__ import annotations

import ctypes as C
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
LIB_PATH = HERE / "libgsdf_b200.so"

GSDF_GRAD_SDF, GSDF_BASE_SDF = 0, 1
_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int32)

STATUS = {0: "GSDF_OK", 1: "GSDF_ERR_INVALID", 2: "GSDF_ERR_CUDA", 3: "GSDF_ERR_CAPACITY", 4: "GSDF_ERR_NCCL",
          5: "GSDF_ERR_NO_FRAME"}

# every symbol include/gsdf_b200.h declares (tests check the library exports them all)
EXPORTS = [
    "gsdf_create", "gsdf_destroy", "gsdf_last_error", "gsdf_reset", "gsdf_get_normal_cache", "gsdf_compute_normals",
    "gsdf_upload_frame", "gsdf_set_frame_device", "gsdf_integrate", "gsdf_integrate_pose", "gsdf_accumulate",
    "gsdf_track", "gsdf_query", "gsdf_voxel_count", "gsdf_download", "gsdf_pose_matrix", "gsdf_pose_apply_increment",
    "gsdf_llt_solve6", "gsdf_synchronize", "gsdf_stream", "gsdf_launch_count", "gsdf_profile_enable",
    "gsdf_profile_read", "gsdf_track_stats", "gsdf_accumulate_noalloc", "gsdf_scan_set_pose", "gsdf_scan_set_tracker",
    "gsdf_scan_submit", "gsdf_scan_collect", "gsdf_scan_pending", "gsdf_scan_run", "gsdf_upload", "gsdf_host_alloc", "gsdf_host_free", "gsdf_accumulate_local", "gsdf_trace_enable", "gsdf_trace_dump", "gsdf_debug_track_timers", "gsdf_debug_frame_log", "gsdf_p2p_export", "gsdf_p2p_connect", "gsdf_comm_unique_id", "gsdf_comm_init",
    "gsdf_extract_mesh", "gsdf_vis_enable", "gsdf_vis_download", "gsdf_get_voxels", "gsdf_grow", "gsdf_version",
]


class GsdfError(RuntimeError):
    def __init__(self, status: int, msg: str):
        super().__init__(f"{STATUS.get(status, status)}: {msg}")
        self.status = status


class Config(C.Structure):
    _fields_ = [("variant", C.c_int32), ("voxel_size", C.c_float), ("truncation", C.c_float), ("z_min", C.c_float),
                ("z_max", C.c_float), ("width", C.c_int32), ("height", C.c_int32), ("K", C.c_float * 9),
                ("normal_window", C.c_int32), ("log2_capacity", C.c_int32), ("device", C.c_int32),
                ("world_size", C.c_int32), ("rank", C.c_int32)]


class FrameResult(C.Structure):
    _fields_ = [("pose7", C.c_float * 7), ("converged", C.c_int32), ("iters", C.c_int32), ("fused", C.c_int32),
                ("n_valid", C.c_int32), ("points", C.c_int64), ("updates", C.c_int64), ("blocks", C.c_int64),
                ("track_iterations", C.c_int64), ("track_valid", C.c_int64)]


class IntegrateStats(C.Structure):
    _fields_ = [("points", C.c_int64), ("updates", C.c_int64), ("blocks", C.c_int64)]


def build(verbose: bool = False) -> None:
    """Compile libgsdf_b200.so for sm_100a in-tree (nvcc cross-compiles without a GPU)."""
    subprocess.run(["make", "-C", str(HERE), "all"], check=True, stdout=None if verbose else subprocess.DEVNULL)


_lib = None


def lib() -> C.CDLL:
    """The C-ABI library.  Raises if it has not been built: the product never falls back to a CPU path."""
    global _lib
    if _lib is not None:
        return _lib
    if not LIB_PATH.exists():
        raise FileNotFoundError(f"{LIB_PATH} is missing: run __graft_entry__.build() (make -C gradient-sdf_b200)")
    L = C.CDLL(str(LIB_PATH))
    H = C.c_void_p
    L.gsdf_create.argtypes = [C.POINTER(Config), C.POINTER(H)]
    L.gsdf_destroy.argtypes = [H]
    L.gsdf_last_error.restype = C.c_char_p
    L.gsdf_last_error.argtypes = [H]
    L.gsdf_reset.argtypes = [H]
    L.gsdf_get_normal_cache.argtypes = [H, _f32p]
    L.gsdf_compute_normals.argtypes = [H, _f32p, _f32p, _f32p]
    L.gsdf_upload_frame.argtypes = [H, C.c_void_p]
    L.gsdf_set_frame_device.argtypes = [H, C.c_void_p]
    L.gsdf_integrate.argtypes = [H, _f32p, _f32p, C.POINTER(IntegrateStats)]
    L.gsdf_integrate_pose.argtypes = [H, _f32p, C.POINTER(IntegrateStats)]
    L.gsdf_accumulate.argtypes = [H, _f32p, _f32p, C.c_int32, _f32p]
    L.gsdf_accumulate_local.argtypes = [H, _f32p, _f32p, C.c_int32, _f32p]
    L.gsdf_track.argtypes = [H, _f32p, C.c_int32, C.c_float, C.c_float, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]
    L.gsdf_query.argtypes = [H, _f32p, C.c_int64, _f32p, _f32p, _f32p]
    L.gsdf_voxel_count.argtypes = [H, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    L.gsdf_download.argtypes = [H, C.c_int64, _i32p, _f32p, _f32p, _f32p, C.POINTER(C.c_int64)]
    L.gsdf_extract_mesh.argtypes = [H, C.c_int64, _f32p, C.POINTER(C.c_int64)]
    L.gsdf_vis_enable.argtypes = [H, C.c_int64]
    L.gsdf_get_voxels.argtypes = [H, C.c_int64, _i32p, _f32p, _f32p, _f32p]
    L.gsdf_grow.argtypes = [H, C.c_int32]
    L.gsdf_vis_download.argtypes = [H, C.c_int64, _i32p, _i32p, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    L.gsdf_pose_matrix.restype = None
    L.gsdf_pose_matrix.argtypes = [_f32p, _f32p, _f32p]
    L.gsdf_pose_apply_increment.restype = None
    L.gsdf_pose_apply_increment.argtypes = [_f32p, _f32p, _f32p]
    L.gsdf_llt_solve6.restype = None
    L.gsdf_llt_solve6.argtypes = [_f32p, _f32p, _f32p]
    L.gsdf_synchronize.argtypes = [H]
    L.gsdf_stream.restype = C.c_void_p
    L.gsdf_stream.argtypes = [H]
    L.gsdf_launch_count.restype = C.c_int64
    L.gsdf_launch_count.argtypes = [H]
    L.gsdf_scan_set_pose.argtypes = [H, _f32p]
    L.gsdf_scan_set_tracker.argtypes = [H, C.c_int32, C.c_float, C.c_float]
    L.gsdf_scan_submit.argtypes = [H, C.c_void_p, C.c_void_p, C.c_int32]
    L.gsdf_scan_collect.argtypes = [H, C.POINTER(FrameResult)]
    L.gsdf_scan_pending.restype = C.c_int32
    L.gsdf_scan_pending.argtypes = [H]
    L.gsdf_accumulate_noalloc.argtypes = [H, _f32p, _f32p, C.c_int32, C.c_void_p]
    L.gsdf_debug_track_timers.argtypes = [H, C.c_void_p, C.POINTER(C.c_int32)]
    L.gsdf_debug_frame_log.argtypes = [H, C.c_void_p]
    L.gsdf_upload.argtypes = [H, C.c_int64, _i32p, _f32p, _f32p, _f32p]
    L.gsdf_scan_run.argtypes = [H, C.POINTER(C.c_void_p), C.c_int64, C.c_int32, C.c_int32, C.POINTER(FrameResult),
                                C.POINTER(C.c_int64)]
    L.gsdf_trace_enable.argtypes = [H, C.c_int32]
    L.gsdf_trace_dump.restype = C.c_int64
    L.gsdf_trace_dump.argtypes = [H, C.c_char_p, C.c_int64]
    L.gsdf_track_stats.argtypes = [H, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]
    L.gsdf_profile_enable.argtypes = [H, C.c_int32]
    L.gsdf_profile_read.argtypes = [H, C.c_int32, C.POINTER(C.c_int64), C.POINTER(C.c_double)]
    L.gsdf_p2p_export.argtypes = [H, C.POINTER(C.c_uint8)]
    L.gsdf_p2p_connect.argtypes = [H, C.POINTER(C.c_uint8)]
    L.gsdf_comm_unique_id.argtypes = [C.POINTER(C.c_uint8)]
    L.gsdf_comm_init.argtypes = [H, C.POINTER(C.c_uint8)]
    L.gsdf_version.restype = C.c_char_p
    _lib = L
    return L


def _fp(a):
    return a.ctypes.data_as(_f32p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


# ---- sharding (host-side restatement of block_owner() in csrc/gsdf_device.cuh) ----------------------------------
def shard_owner(voxel_keys, world_size: int):
    """Rank that owns each voxel's 8^3 block: (mix64(packed 21-bit block coordinates) >> 40) % world_size."""
    k = np.asarray(voxel_keys, dtype=np.int64).reshape(-1, 3) >> 3
    bias = 1 << 20
    m = np.uint64(0x1FFFFF)
    key = ((k[:, 0] + bias).astype(np.uint64) & m) | (((k[:, 1] + bias).astype(np.uint64) & m) << np.uint64(21)) | \
          (((k[:, 2] + bias).astype(np.uint64) & m) << np.uint64(42))
    x = key.copy()
    with np.errstate(over="ignore"):
        x ^= x >> np.uint64(33); x *= np.uint64(0xff51afd7ed558ccd)
        x ^= x >> np.uint64(33); x *= np.uint64(0xc4ceb9fe1a85ec53)
        x ^= x >> np.uint64(33)
    return ((x >> np.uint64(40)) % np.uint64(world_size)).astype(np.int32)


# ---- pose algebra (no GPU needed) -------------------------------------------------------------------------
def pose_matrix(pose7):
    """(R 3x3, normalised pose7) as Sophus::SE3f(quaternion, t).rotationMatrix() gives them."""
    p = _f32(pose7)
    R, pn = np.zeros(9, np.float32), np.zeros(7, np.float32)
    lib().gsdf_pose_matrix(_fp(p), _fp(R), _fp(pn))
    return R.reshape(3, 3), pn


def pose_apply_increment(pose7, xi):
    p, x, out = _f32(pose7), _f32(xi), np.zeros(7, np.float32)
    lib().gsdf_pose_apply_increment(_fp(p), _fp(x), _fp(out))
    return out


def llt_solve6(H, g):
    H, g, xi = _f32(H).reshape(36), _f32(g), np.zeros(6, np.float32)
    lib().gsdf_llt_solve6(_fp(H), _fp(g), _fp(xi))
    return xi


class NormalEstimator:
    """cv::NormalEstimator<float>(width, height, K, Size(win, win)) — parameters only; the device cache lives in
    the volume the estimator is handed to (as the reference hands NEst to Sdf::update)."""

    def __init__(self, width: int, height: int, K, window: int = 11):
        self.width, self.height, self.window = int(width), int(height), int(window)
        self.K = _f32(K).reshape(3, 3)


class Sdf:
    """Abstract volume (sdf_tracker/Sdf.h).  One instance = one gsdf_handle = one GPU shard."""

    variant = GSDF_GRAD_SDF

    def __init__(self, voxel_size: float, T: float, NEst: NormalEstimator, log2_capacity: int = 17, device: int = 0,
                 world_size: int = 1, rank: int = 0, z_min: float = 0.5, z_max: float = 3.5):
        cfg = Config()
        cfg.variant = self.variant
        cfg.voxel_size = float(np.float32(voxel_size))
        cfg.truncation = float(np.float32(T))
        cfg.z_min, cfg.z_max = z_min, z_max
        cfg.width, cfg.height = NEst.width, NEst.height
        for i, v in enumerate(NEst.K.reshape(9)):
            cfg.K[i] = float(v)
        cfg.normal_window = NEst.window
        cfg.log2_capacity = log2_capacity
        cfg.device, cfg.world_size, cfg.rank = device, world_size, rank
        self.cfg = cfg
        self.width, self.height = NEst.width, NEst.height
        self.voxel_size = cfg.voxel_size
        self.L = lib()
        self.h = C.c_void_p()
        st = self.L.gsdf_create(C.byref(cfg), C.byref(self.h))
        if st != 0:
            raise GsdfError(st, (self.L.gsdf_last_error(None) or b"").decode())

    # -- plumbing
    def _check(self, st: int):
        if st != 0:
            raise GsdfError(st, (self.L.gsdf_last_error(self.h) or b"").decode())

    def close(self):
        if getattr(self, "h", None) and self.h.value:
            self.L.gsdf_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def reset(self):
        self._check(self.L.gsdf_reset(self.h))

    def synchronize(self):
        self._check(self.L.gsdf_synchronize(self.h))

    @property
    def stream(self) -> int:
        return int(self.L.gsdf_stream(self.h) or 0)

    @property
    def launch_count(self) -> int:
        return int(self.L.gsdf_launch_count(self.h))

    def trace_enable(self, on: bool = True):
        self._check(self.L.gsdf_trace_enable(self.h, 1 if on else 0))

    def trace_dump(self):
        """[(label, microseconds since the first mark)] of the frames submitted since trace_enable."""
        n = self.L.gsdf_trace_dump(self.h, None, 0)
        buf = C.create_string_buffer(int(n))
        self.L.gsdf_trace_dump(self.h, buf, n)
        return [(l.rsplit(" ", 1)[0], float(l.rsplit(" ", 1)[1])) for l in buf.value.decode().splitlines()]

    def track_stats(self):
        """(cumulative Gauss-Newton iterations, cumulative contributing pixels) of gsdf_track on this handle."""
        a, b = C.c_int64(0), C.c_int64(0)
        self._check(self.L.gsdf_track_stats(self.h, C.byref(a), C.byref(b)))
        return int(a.value), int(b.value)

    def profile_enable(self, on: bool = True):
        self._check(self.L.gsdf_profile_enable(self.h, 1 if on else 0))

    def profile_read(self, kind: int):
        """(launches, total_ms) of kernel kind 0 = normals, 1 = fuse, 2 = accumulate since profile_enable."""
        n, ms = C.c_int64(0), C.c_double(0.0)
        self._check(self.L.gsdf_profile_read(self.h, kind, C.byref(n), C.byref(ms)))
        return int(n.value), float(ms.value)

    def p2p_export(self) -> bytes:
        buf = (C.c_uint8 * 64)()
        self._check(self.L.gsdf_p2p_export(self.h, buf))
        return bytes(buf)

    def p2p_connect(self, handles: bytes):
        """handles: world_size * 64 bytes, the ranks' p2p_export() results in rank order."""
        buf = (C.c_uint8 * len(handles)).from_buffer_copy(handles)
        self._check(self.L.gsdf_p2p_connect(self.h, buf))

    def comm_init(self, id128: bytes):
        buf = (C.c_uint8 * 128).from_buffer_copy(id128)
        self._check(self.L.gsdf_comm_init(self.h, buf))

    # -- frames
    def upload_frame(self, depth):
        """depth: host float32 array (height, width), metres; a pinned torch tensor's numpy view works too."""
        d = _f32(depth)
        assert d.size == self.width * self.height
        self._keep = d
        self._check(self.L.gsdf_upload_frame(self.h, d.ctypes.data))

    def set_frame_device(self, dev_ptr: int):
        self._check(self.L.gsdf_set_frame_device(self.h, C.c_void_p(dev_ptr)))

    # -- NormalEstimator
    def normal_cache(self):
        out = np.zeros((11, self.height, self.width), np.float32)
        self._check(self.L.gsdf_get_normal_cache(self.h, _fp(out)))
        return out

    def compute_normals(self):
        n = np.zeros((3, self.height, self.width), np.float32)
        self._check(self.L.gsdf_compute_normals(self.h, _fp(n[0]), _fp(n[1]), _fp(n[2])))
        return n

    # -- Sdf::update / setup
    def update(self, depth, pose7, stats: bool = True, stored: bool = False):
        """Sdf::update(color, depth, K, pose, NEst): depth=None fuses the frame already set.
        stored=False: pose7 is raw and goes through the SE3(quaternion, t) constructor (normalised once);
        stored=True: pose7 is an SE3 as the tracker holds it (RigidPointOptimizer.pose()), used as is."""
        if depth is not None:
            self.upload_frame(depth)
        p = _f32(pose7) if stored else pose_matrix(pose7)[1]
        st = IntegrateStats()
        self._check(self.L.gsdf_integrate_pose(self.h, _fp(p), C.byref(st) if stats else None))
        return (st.points, st.updates, st.blocks) if stats else None

    def update_Rt(self, depth, R, t, stats: bool = True):
        if depth is not None:
            self.upload_frame(depth)
        R, t = _f32(R).reshape(9), _f32(t)
        st = IntegrateStats()
        self._check(self.L.gsdf_integrate(self.h, _fp(R), _fp(t), C.byref(st) if stats else None))
        return (st.points, st.updates, st.blocks) if stats else None

    def setup(self, depth):
        """Sdf::setup = update at the identity pose (Sdf.h:119-121)."""
        return self.update(depth, np.array([0, 0, 0, 0, 0, 0, 1], np.float32))

    # -- the asynchronous frame pipeline (main_scan_3d.cpp:208-266 without the host in the loop)
    def scan_set_pose(self, pose7):
        p = _f32(pose7)
        self._check(self.L.gsdf_scan_set_pose(self.h, _fp(p)))

    def scan_set_tracker(self, max_iters=25, conv_threshold=1e-3, damping=1.0):
        self._check(self.L.gsdf_scan_set_tracker(self.h, max_iters, conv_threshold, damping))

    def scan_submit(self, depth_host=None, depth_dev: int = 0, track: bool = True):
        """Enqueue one frame (returns at once).  depth_host: float32 array (pinned for true overlap), kept alive by
        the caller until collected; depth_dev: device pointer."""
        hp = None
        if depth_host is not None:
            d = depth_host if (isinstance(depth_host, np.ndarray) and depth_host.dtype == np.float32
                               and depth_host.flags["C_CONTIGUOUS"]) else _f32(depth_host)
            self._pending_keep = getattr(self, "_pending_keep", [])[-8:] + [d]
            hp = d.ctypes.data
        self._check(self.L.gsdf_scan_submit(self.h, hp, C.c_void_p(depth_dev) if depth_dev else None,
                                            1 if track else 0))

    def scan_collect(self) -> FrameResult:
        r = FrameResult()
        self._check(self.L.gsdf_scan_collect(self.h, C.byref(r)))
        return r

    def scan_pending(self) -> int:
        return int(self.L.gsdf_scan_pending(self.h))

    def scan_run(self, frame_ptrs, on_device: bool, track: bool = True):
        """The Scan3D frame loop in one call (gsdf_scan_run): frame_ptrs = host or device addresses of the frames, which
        the caller keeps alive.  Returns the list of FrameResult in frame order."""
        n = len(frame_ptrs)
        ptrs = (C.c_void_p * n)(*[int(p) for p in frame_ptrs])
        res = (FrameResult * n)()
        done = C.c_int64(0)
        self._check(self.L.gsdf_scan_run(self.h, ptrs, n, 1 if on_device else 0, 1 if track else 0, res, C.byref(done)))
        return list(res)

    # -- Sdf::tsdf / weights (batched)
    def query(self, pts):
        p = _f32(pts).reshape(-1, 3)
        n = p.shape[0]
        phi, grad, w = np.zeros(n, np.float32), np.zeros((n, 3), np.float32), np.zeros(n, np.float32)
        self._check(self.L.gsdf_query(self.h, _fp(p), n, _fp(phi), _fp(grad), _fp(w)))
        return phi, grad, w

    def accumulate(self, R, t, sampling: int = 1):
        R, t, out = _f32(R).reshape(9), _f32(t), np.zeros(29, np.float32)
        self._check(self.L.gsdf_accumulate(self.h, _fp(R), _fp(t), sampling, _fp(out)))
        return out

    def accumulate_noalloc(self, R, t, out29_dev_ptr: int, sampling: int = 1):
        """gsdf_accumulate_noalloc: same sums as accumulate(), no allocation anywhere on the way; the 29 floats
        land in caller-owned DEVICE memory at out29_dev_ptr once the handle's stream gets there."""
        R, t = _f32(R).reshape(9), _f32(t)
        self._check(self.L.gsdf_accumulate_noalloc(self.h, _fp(R), _fp(t), sampling, C.c_void_p(out29_dev_ptr)))

    def accumulate_local(self, R, t, sampling: int = 1):
        """This shard's partial sums (no collective)."""
        R, t, out = _f32(R).reshape(9), _f32(t), np.zeros(29, np.float32)
        self._check(self.L.gsdf_accumulate_local(self.h, _fp(R), _fp(t), sampling, _fp(out)))
        return out

    # -- map access
    def voxel_count(self):
        n, b = C.c_int64(0), C.c_int64(0)
        self._check(self.L.gsdf_voxel_count(self.h, C.byref(n), C.byref(b)))
        return int(n.value), int(b.value)

    def upload(self, keys, dist, weight, grad=None):
        """Checkpoint / resume: set the listed voxels (arrays as returned by download())."""
        k = np.ascontiguousarray(keys, np.int32).reshape(-1, 3)
        d, w = _f32(dist).reshape(-1), _f32(weight).reshape(-1)
        g = None if grad is None else _f32(grad).reshape(-1, 3)
        assert len(d) == len(w) == len(k) and (g is None or len(g) == len(k))
        self._check(self.L.gsdf_upload(self.h, len(k), k.ctypes.data_as(_i32p), _fp(d), _fp(w), _fp(g) if g is not None else None))

    def download(self, sort: bool = True):
        n, _ = self.voxel_count()
        keys, dist = np.zeros((n, 3), np.int32), np.zeros(n, np.float32)
        weight, grad = np.zeros(n, np.float32), np.zeros((n, 3), np.float32)
        got = C.c_int64(0)
        if n:
            self._check(self.L.gsdf_download(self.h, n, keys.ctypes.data_as(_i32p), _fp(dist), _fp(weight), _fp(grad),
                                             C.byref(got)))
            assert got.value == n
        if sort and n:
            order = np.lexsort((keys[:, 2], keys[:, 1], keys[:, 0]))
            keys, dist, weight, grad = keys[order], dist[order], weight[order], grad[order]
        return keys, dist, weight, grad

    def get_voxels(self, keys):
        """getSdf(Vec3i) for many voxels: (dist, weight, grad) at the given keys, zeros where there is no voxel."""
        k = np.ascontiguousarray(keys, np.int32).reshape(-1, 3)
        n = len(k)
        d, w, g = np.zeros(n, np.float32), np.zeros(n, np.float32), np.zeros((n, 3), np.float32)
        if n:
            self._check(self.L.gsdf_get_voxels(self.h, n, k.ctypes.data_as(_i32p), _fp(d), _fp(w), _fp(g)))
        return d, w, g

    def grow(self, new_log2_capacity: int):
        """Rehash the block table into 2^new_log2_capacity slots (gsdf_grow)."""
        self._check(self.L.gsdf_grow(self.h, new_log2_capacity))

    def vis_enable(self, max_entries: int):
        """Start recording vis_ (MapGradPixelSdfOmp.cpp:118-120): which update() calls touched which voxel."""
        self._check(self.L.gsdf_vis_enable(self.h, max_entries))

    def get_vis(self):
        """get_vis() (MapGradPixelSdf.h:140-142) as arrays: keys (n, 3) int32, frames (n,) int32 — one row per (voxel, frame)
        pair, sorted by key then frame — and counter_ (frames fused so far)."""
        n, nf = C.c_int64(0), C.c_int64(0)
        self._check(self.L.gsdf_vis_download(self.h, 0, None, None, C.byref(n), C.byref(nf)))
        keys, frames = np.zeros((n.value, 3), np.int32), np.zeros(n.value, np.int32)
        if n.value:
            self._check(self.L.gsdf_vis_download(self.h, n.value, keys.ctypes.data_as(_i32p), frames.ctypes.data_as(_i32p),
                                                 C.byref(n), C.byref(nf)))
            o = np.lexsort((frames, keys[:, 2], keys[:, 1], keys[:, 0]))
            keys, frames = keys[o], frames[o]
        return keys, frames, nf.value

    def extract_mesh(self):
        """Sdf::extract_mesh (marching cubes, MapGradPixelSdfOmp.cpp:132-183): triangles as an (n, 3, 3) float array in the
        reference's order; vertex i of the reference's PLY is .reshape(-1, 3)[i], face f is (3f, 3f+1, 3f+2)."""
        n = C.c_int64(0)
        self._check(self.L.gsdf_extract_mesh(self.h, 0, None, C.byref(n)))
        v = np.zeros((n.value, 3, 3), np.float32)
        if n.value:
            self._check(self.L.gsdf_extract_mesh(self.h, n.value, _fp(v), C.byref(n)))
        return v


class MapGradPixelSdf(Sdf):
    """MapGradPixelSdf(voxel_size, T) — Gradient-SDF volume (sdf_tracker/MapGradPixelSdf.h:99-103)."""
    variant = GSDF_GRAD_SDF


class MapPixelSdf(Sdf):
    """MapPixelSdf(voxel_size, T) — plain SDF volume (sdf_tracker/MapPixelSdf.h:97-101)."""
    variant = GSDF_BASE_SDF


class RigidPointOptimizer:
    """RigidPointOptimizer(tSDF) / (num_iterations, conv_threshold, damping, tSDF)
    (sdf_tracker/RigidPointOptimizer.h:55-61, RigidOptimizer.h:68-83)."""

    def __init__(self, tSDF: Sdf, num_iterations: int = 25, conv_threshold: float = 1e-3, damping: float = 1.0):
        self.tSDF = tSDF
        self.num_iterations, self.conv_threshold, self.damping = num_iterations, conv_threshold, damping
        self._pose = np.array([0, 0, 0, 0, 0, 0, 1], np.float32)
        self.last_iterations = 0

    def set_num_iterations(self, n):
        self.num_iterations = int(n)

    def set_conv_threshold(self, t):
        self.conv_threshold = float(t)

    def set_damping(self, d):
        self.damping = float(d)

    def set_pose(self, pose7):
        self._pose = pose_matrix(pose7)[1]  # SE3(q, t) normalises the quaternion

    def pose(self):
        return self._pose.copy()

    def optimize(self, depth) -> bool:
        """optimize(depth, K): depth=None tracks against the frame already set on the volume."""
        s = self.tSDF
        if depth is not None:
            s.upload_frame(depth)
        conv, it = C.c_int32(0), C.c_int32(0)
        s._check(s.L.gsdf_track(s.h, _fp(self._pose), self.num_iterations, self.conv_threshold, self.damping,
                                C.byref(conv), C.byref(it)))
        self.last_iterations = int(it.value)
        return bool(conv.value)
