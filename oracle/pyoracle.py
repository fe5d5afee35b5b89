"""TEST INFRASTRUCTURE — ctypes access to the CPU oracle.  NOT part of the product.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.

  Oracle   -> oracle/liboracle.so        plain-C restatement (gsdf_oracle.c)
  RefScan  -> oracle/_ref/libgsdf_ref.so the reference's own OMP translation units (ref_driver.cpp)

Both expose the same methods so tests can run either against the CUDA path.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
ORACLE_SO = HERE / "liboracle.so"
REF_SO = HERE / "_ref" / "libgsdf_ref.so"

_f32p = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int32)


def _fp(a):
    return a.ctypes.data_as(_f32p)


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def build(force: bool = False) -> None:
    """Compile liboracle.so (always) and _ref (only where /root/reference exists)."""
    args = ["make", "-C", str(HERE), "all"]
    if force:
        args.insert(1, "-B")
    subprocess.run(args, check=True, stdout=subprocess.DEVNULL)


def have_ref() -> bool:
    return REF_SO.exists()


class _Base:
    variant: int
    w: int
    h: int

    def _alloc_map(self, n):
        return (np.zeros((n, 3), np.int32), np.zeros(n, np.float32), np.zeros(n, np.float32),
                np.zeros((n, 3), np.float32))

    @staticmethod
    def sort_map(keys, dist, weight, grad):
        order = np.lexsort((keys[:, 2], keys[:, 1], keys[:, 0]))
        return keys[order], dist[order], weight[order], grad[order]


class Oracle(_Base):
    """Plain-C restatement (single thread, pixel order = the OMP build with one thread)."""

    def __init__(self, variant, voxel_size, trunc_factor, w, h, K, normal_win=11):
        if not ORACLE_SO.exists():
            build()
        L = C.CDLL(str(ORACLE_SO))
        self.L = L
        L.orc_create.restype = C.c_void_p
        L.orc_create.argtypes = [C.c_int, C.c_float, C.c_float, C.c_int, C.c_int, _f32p, C.c_int]
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_set_zrange.argtypes = [C.c_void_p, C.c_float, C.c_float]
        L.orc_set_shard.argtypes = [C.c_void_p, C.c_int, C.c_int]
        L.orc_voxel_owner.restype = C.c_int
        L.orc_voxel_owner.argtypes = [_i32p, C.c_int]
        L.orc_normal_cache.argtypes = [C.c_void_p, _f32p]
        L.orc_normals.argtypes = [C.c_void_p, _f32p, _f32p, _f32p, _f32p]
        L.orc_update.restype = C.c_longlong
        L.orc_update.argtypes = [C.c_void_p, _f32p, _f32p, _f32p, C.POINTER(C.c_longlong)]
        L.orc_voxel_count.restype = C.c_longlong
        L.orc_voxel_count.argtypes = [C.c_void_p]
        L.orc_download.argtypes = [C.c_void_p, _i32p, _f32p, _f32p, _f32p]
        L.orc_query.argtypes = [C.c_void_p, _f32p, C.c_longlong, _f32p, _f32p, _f32p]
        L.orc_accumulate.argtypes = [C.c_void_p, _f32p, _f32p, _f32p, C.c_int, _f32p]
        L.orc_optimize.restype = C.c_int
        L.orc_optimize.argtypes = [C.c_void_p, _f32p, _f32p, C.c_int, C.c_float, C.c_float, C.c_int,
                                   C.POINTER(C.c_int)]
        L.orc_pose_matrix.argtypes = [_f32p, _f32p, _f32p]
        L.orc_apply_increment.argtypes = [_f32p, _f32p, _f32p]
        L.orc_llt_solve.restype = C.c_int
        L.orc_llt_solve.argtypes = [_f32p, _f32p, _f32p]
        self.variant, self.w, self.h = variant, w, h
        self.K = _f32(K).reshape(9)
        self.handle = L.orc_create(variant, voxel_size, trunc_factor, w, h, _fp(self.K), normal_win)
        self.pose = np.array([0, 0, 0, 0, 0, 0, 1], np.float32)
        self.iters, self.thr, self.damping = 25, 1e-3, 1.0

    def close(self):
        if self.handle:
            self.L.orc_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_shard(self, world, rank):
        self.L.orc_set_shard(self.handle, world, rank)

    def voxel_owner(self, keys, world):
        k = np.ascontiguousarray(keys, dtype=np.int32).reshape(-1, 3)
        return np.array([self.L.orc_voxel_owner(k[i].ctypes.data_as(_i32p), world) for i in range(len(k))], np.int32)

    def normal_cache(self):
        out = np.zeros((11, self.h, self.w), np.float32)
        self.L.orc_normal_cache(self.handle, _fp(out))
        return out

    def normals(self, depth):
        d = _f32(depth)
        n = np.zeros((3, self.h, self.w), np.float32)
        self.L.orc_normals(self.handle, _fp(d), _fp(n[0]), _fp(n[1]), _fp(n[2]))
        return n

    def pose_matrix(self, pose7):
        p = _f32(pose7)
        R = np.zeros(9, np.float32)
        pn = np.zeros(7, np.float32)
        self.L.orc_pose_matrix(_fp(p), _fp(R), _fp(pn))
        return R.reshape(3, 3), pn

    def update(self, depth, pose7):
        """Fuse at pose7; returns (R, t, n_updates, n_points)."""
        d = _f32(depth)
        R, pn = self.pose_matrix(pose7)
        t = pn[:3].copy()
        pts = C.c_longlong(0)
        upd = self.L.orc_update(self.handle, _fp(d), _fp(R), _fp(t), C.byref(pts))
        return R, t, int(upd), int(pts.value)

    def update_Rt(self, depth, R, t):
        d, R, t = _f32(depth), _f32(R), _f32(t)
        pts = C.c_longlong(0)
        upd = self.L.orc_update(self.handle, _fp(d), _fp(R), _fp(t), C.byref(pts))
        return int(upd), int(pts.value)

    def voxel_count(self):
        return int(self.L.orc_voxel_count(self.handle))

    def download(self, sort=True):
        n = self.voxel_count()
        k, d, w, g = self._alloc_map(n)
        if n:
            self.L.orc_download(self.handle, k.ctypes.data_as(_i32p), _fp(d), _fp(w), _fp(g))
        return self.sort_map(k, d, w, g) if sort else (k, d, w, g)

    def query(self, pts):
        p = _f32(pts).reshape(-1, 3)
        n = p.shape[0]
        phi, grad, w = np.zeros(n, np.float32), np.zeros((n, 3), np.float32), np.zeros(n, np.float32)
        self.L.orc_query(self.handle, _fp(p), n, _fp(phi), _fp(grad), _fp(w))
        return phi, grad, w

    def accumulate(self, depth, R, t, sampling=1):
        d, R, t = _f32(depth), _f32(R), _f32(t)
        out = np.zeros(29, np.float32)
        self.L.orc_accumulate(self.handle, _fp(d), _fp(R), _fp(t), sampling, _fp(out))
        return out

    def set_pose(self, pose7):
        self.pose = _f32(pose7).copy()

    def get_pose(self):
        return self.pose.copy()

    def set_opt(self, iters, thr, damping):
        self.iters, self.thr, self.damping = iters, thr, damping

    def optimize(self, depth, sampling=1):
        d = _f32(depth)
        it = C.c_int(0)
        conv = self.L.orc_optimize(self.handle, _fp(d), _fp(self.pose), self.iters, self.thr, self.damping,
                                   sampling, C.byref(it))
        return bool(conv), int(it.value)

    def apply_increment(self, pose7, xi):
        p, x = _f32(pose7), _f32(xi)
        out = np.zeros(7, np.float32)
        self.L.orc_apply_increment(_fp(p), _fp(x), _fp(out))
        return out

    def llt_solve(self, H, g):
        H, g = _f32(H).reshape(36), _f32(g)
        xi = np.zeros(6, np.float32)
        ok = self.L.orc_llt_solve(_fp(H), _fp(g), _fp(xi))
        return xi, bool(ok)


class RefScan(_Base):
    """The reference's own OMP classes.  Set OMP_THREAD_LIMIT=1 in the environment BEFORE the first
    construction for the bit-reproducible mode (the tracker calls omp_set_num_threads(4) itself,
    RigidPointOptimizerOmp.cpp:68, so OMP_NUM_THREADS is not enough)."""

    def __init__(self, variant, voxel_size, trunc_factor, w, h, K, normal_win=11, quiet=True):
        if not REF_SO.exists():
            raise FileNotFoundError(f"{REF_SO} not built (needs /root/reference; run `make -C oracle ref`)")
        L = C.CDLL(str(REF_SO))
        self.L = L
        L.gsdf_ref_create.restype = C.c_void_p
        L.gsdf_ref_create.argtypes = [C.c_int, C.c_float, C.c_float, C.c_int, C.c_int, _f32p, C.c_int, C.c_int]
        L.gsdf_ref_destroy.argtypes = [C.c_void_p]
        L.gsdf_ref_normal_cache.argtypes = [C.c_void_p, _f32p]
        L.gsdf_ref_normals.argtypes = [C.c_void_p, _f32p, _f32p, _f32p, _f32p]
        L.gsdf_ref_update.argtypes = [C.c_void_p, _f32p, _f32p, _f32p, _f32p]
        L.gsdf_ref_set_pose.argtypes = [C.c_void_p, _f32p]
        L.gsdf_ref_get_pose.argtypes = [C.c_void_p, _f32p, _f32p]
        L.gsdf_ref_set_opt.argtypes = [C.c_void_p, C.c_int, C.c_float, C.c_float]
        L.gsdf_ref_optimize.restype = C.c_int
        L.gsdf_ref_optimize.argtypes = [C.c_void_p, _f32p, C.POINTER(C.c_int)]
        L.gsdf_ref_voxel_count.restype = C.c_longlong
        L.gsdf_ref_voxel_count.argtypes = [C.c_void_p]
        L.gsdf_ref_download.argtypes = [C.c_void_p, _i32p, _f32p, _f32p, _f32p]
        L.gsdf_ref_query.argtypes = [C.c_void_p, _f32p, C.c_longlong, _f32p, _f32p, _f32p]
        L.gsdf_ref_pose_matrix.argtypes = [_f32p, _f32p, _f32p]
        L.gsdf_ref_apply_increment.argtypes = [_f32p, _f32p, _f32p]
        L.gsdf_ref_llt_solve.argtypes = [_f32p, _f32p, _f32p]
        L.gsdf_ref_write.restype = C.c_int
        L.gsdf_ref_write.argtypes = [C.c_void_p, C.c_int, C.c_char_p]
        L.gsdf_ref_upload.argtypes = [C.c_void_p, C.c_longlong, _i32p, _f32p, _f32p, _f32p]
        L.gsdf_ref_vis.restype = C.c_longlong
        L.gsdf_ref_vis.argtypes = [C.c_void_p, _i32p, _i32p, C.POINTER(C.c_longlong)]
        self.variant, self.w, self.h = variant, w, h
        self.K = _f32(K).reshape(9)
        self.handle = L.gsdf_ref_create(variant, voxel_size, trunc_factor, w, h, _fp(self.K), normal_win,
                                        1 if quiet else 0)

    def close(self):
        if self.handle:
            self.L.gsdf_ref_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def write(self, which, filename):
        """The reference's own writers on the current map: which = 'mesh' | 'pc' | 'sdf' (extract_mesh, extract_pc,
        save_sdf); returns their bool."""
        return bool(self.L.gsdf_ref_write(self.handle, {"mesh": 0, "pc": 1, "sdf": 2}[which], str(filename).encode()))

    def get_vis(self):
        """The reference's vis_ as (keys (n, 3), frames (n,)) sorted by key then frame, and the largest vector size."""
        mx = C.c_longlong(0)
        n = self.L.gsdf_ref_vis(self.handle, None, None, C.byref(mx))
        keys, frames = np.zeros((n, 3), np.int32), np.zeros(n, np.int32)
        if n:
            self.L.gsdf_ref_vis(self.handle, keys.ctypes.data_as(_i32p), frames.ctypes.data_as(_i32p), C.byref(mx))
            o = np.lexsort((frames, keys[:, 2], keys[:, 1], keys[:, 0]))
            keys, frames = keys[o], frames[o]
        return keys, frames, int(mx.value)

    def upload(self, keys, dist, weight, grad):
        """Replace the reference's hash by the given voxels (arrays as download() returns them)."""
        k = np.ascontiguousarray(keys, np.int32); d = _f32(dist); w = _f32(weight); g = _f32(grad)
        self.L.gsdf_ref_upload(self.handle, len(d), k.ctypes.data_as(_i32p), _fp(d), _fp(w), _fp(g))

    def normal_cache(self):
        out = np.zeros((11, self.h, self.w), np.float32)
        self.L.gsdf_ref_normal_cache(self.handle, _fp(out))
        return out

    def normals(self, depth):
        d = _f32(depth)
        n = np.zeros((3, self.h, self.w), np.float32)
        self.L.gsdf_ref_normals(self.handle, _fp(d), _fp(n[0]), _fp(n[1]), _fp(n[2]))
        return n

    def pose_matrix(self, pose7):
        p = _f32(pose7)
        R = np.zeros(9, np.float32)
        pn = np.zeros(7, np.float32)
        self.L.gsdf_ref_pose_matrix(_fp(p), _fp(R), _fp(pn))
        return R.reshape(3, 3), pn

    def update(self, depth, pose7):
        """Fuse at pose7; returns (R, t) the reference used (update counts are not observable here)."""
        d, p = _f32(depth), _f32(pose7)
        R, t = np.zeros(9, np.float32), np.zeros(3, np.float32)
        self.L.gsdf_ref_update(self.handle, _fp(d), _fp(p), _fp(R), _fp(t))
        return R.reshape(3, 3), t, None, None

    def voxel_count(self):
        return int(self.L.gsdf_ref_voxel_count(self.handle))

    def download(self, sort=True):
        n = self.voxel_count()
        k, d, w, g = self._alloc_map(n)
        if n:
            self.L.gsdf_ref_download(self.handle, k.ctypes.data_as(_i32p), _fp(d), _fp(w), _fp(g))
        return self.sort_map(k, d, w, g) if sort else (k, d, w, g)

    def query(self, pts):
        p = _f32(pts).reshape(-1, 3)
        n = p.shape[0]
        phi, grad, w = np.zeros(n, np.float32), np.zeros((n, 3), np.float32), np.zeros(n, np.float32)
        self.L.gsdf_ref_query(self.handle, _fp(p), n, _fp(phi), _fp(grad), _fp(w))
        return phi, grad, w

    def set_pose(self, pose7):
        p = _f32(pose7)
        self.L.gsdf_ref_set_pose(self.handle, _fp(p))

    def get_pose(self):
        p = np.zeros(7, np.float32)
        self.L.gsdf_ref_get_pose(self.handle, _fp(p), None)
        return p

    def set_opt(self, iters, thr, damping):
        self.L.gsdf_ref_set_opt(self.handle, iters, thr, damping)

    def optimize(self, depth):
        d = _f32(depth)
        it = C.c_int(0)
        conv = self.L.gsdf_ref_optimize(self.handle, _fp(d), C.byref(it))
        return bool(conv), int(it.value)

    def apply_increment(self, pose7, xi):
        p, x = _f32(pose7), _f32(xi)
        out = np.zeros(7, np.float32)
        self.L.gsdf_ref_apply_increment(_fp(p), _fp(x), _fp(out))
        return out

    def llt_solve(self, H, g):
        H, g = _f32(H).reshape(36), _f32(g)
        xi = np.zeros(6, np.float32)
        self.L.gsdf_ref_llt_solve(_fp(H), _fp(g), _fp(xi))
        return xi, True


def threads_in_use() -> int:
    lim = os.environ.get("OMP_THREAD_LIMIT")
    n = os.cpu_count() or 1
    return min(n, int(lim)) if lim else n
