"""Seeded synthetic Scan3D input: analytic sphere ray-caster + Kinect disparity noise + orbit poses.

Restates the reference's procedural fixture (matlab/RenderSpheres.m:38-123, matlab/add_kinect_noise.m:43-74)
in numpy; the orbit is generated parametrically with the same shape as matlab/poses.txt (one 360 degree turn of
radius ~2 m about the origin, z rising from -0.125 m to 0.25 m, camera looking at the origin, image "down" =
world +z) so any frame count can be produced.  Host-side input generator only: not on the timed path.
"""
from __future__ import annotations

import numpy as np

K_KINECT = np.array([[525.0, 0.0, 319.5], [0.0, 525.0, 239.5], [0.0, 0.0, 1.0]])  # RenderSpheres.m:39
SCENE_SEED = 1234
NOISE_SEED = 5678


def scaled_intrinsics(width: int, height: int) -> np.ndarray:
    """Kinect intrinsics scaled from 640x480 to width x height (sweep config)."""
    sx, sy = width / 640.0, height / 480.0
    K = K_KINECT.copy()
    K[0, 0] *= sx
    K[1, 1] *= sy
    K[0, 2] = (K_KINECT[0, 2] + 0.5) * sx - 0.5
    K[1, 2] = (K_KINECT[1, 2] + 0.5) * sy - 0.5
    return K


def random_spheres(n: int = 5, seed: int = SCENE_SEED, extent: float = 1.0, rmin: float = 0.0625,
                   rmax: float = 0.5) -> np.ndarray:
    """n non-intersecting spheres [cx cy cz r], centre ~U(-extent/2, extent/2)^3 (RenderSpheres.m:46-53)."""
    rng = np.random.default_rng(seed)
    out = []
    while len(out) < n:
        c = (rng.random(3) - 0.5) * extent
        r = rmin + (rmax - rmin) * rng.random()
        if all(np.linalg.norm(c - s[:3]) > s[3] + r for s in out):
            out.append(np.array([c[0], c[1], c[2], r]))
    return np.stack(out)


def _look_at(c: np.ndarray, target: np.ndarray) -> np.ndarray:
    z = target - c
    z = z / np.linalg.norm(z)
    x = np.cross(np.array([0.0, 0.0, 1.0]), z)
    x = x / np.linalg.norm(x)
    y = np.cross(z, x)
    return np.stack([x, y, z], axis=1)  # columns = camera axes in world


def rot_to_quat(R: np.ndarray) -> np.ndarray:
    """(qx,qy,qz,qw) from a rotation matrix."""
    t = np.trace(R)
    if t > 0:
        s = np.sqrt(t + 1.0) * 2
        q = np.array([(R[2, 1] - R[1, 2]) / s, (R[0, 2] - R[2, 0]) / s, (R[1, 0] - R[0, 1]) / s, 0.25 * s])
    else:
        i = int(np.argmax(np.diag(R)))
        j, k = (i + 1) % 3, (i + 2) % 3
        s = np.sqrt(R[i, i] - R[j, j] - R[k, k] + 1.0) * 2
        q = np.zeros(4)
        q[i] = 0.25 * s
        q[j] = (R[j, i] + R[i, j]) / s
        q[k] = (R[k, i] + R[i, k]) / s
        q[3] = (R[k, j] - R[j, k]) / s
    return q / np.linalg.norm(q)


def quat_to_rot(q: np.ndarray) -> np.ndarray:
    x, y, z, w = q
    return np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                     [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                     [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])


def orbit_poses(n_frames: int, radius: float = 2.0, z0: float = -0.125, z1: float = 0.25,
                turn: float = 2 * np.pi, target=(0.0, 0.0, 0.0)) -> np.ndarray:
    """n_frames camera-to-world poses [tx ty tz qx qy qz qw] (TUM order, ImageLoader.h:231-259)."""
    poses = np.zeros((n_frames, 7))
    tgt = np.asarray(target, dtype=np.float64)
    for k in range(n_frames):
        th = turn * (k + 1) / n_frames
        c = np.array([radius * np.sin(th), -radius * np.cos(th), z0 + (z1 - z0) * k / max(1, n_frames - 1)]) + tgt
        R = _look_at(c, tgt)
        poses[k, :3] = c
        poses[k, 3:] = rot_to_quat(R)
    return poses


def render_depth(spheres: np.ndarray, pose7: np.ndarray, K: np.ndarray = K_KINECT, width: int = 640,
                 height: int = 480) -> np.ndarray:
    """Noise-free z-depth (float64, 0 = miss): nearest root of A z^2 + B z + C = 0 per pixel and sphere
    (RenderSpheres.m:91-123)."""
    u, v = np.meshgrid(np.arange(width, dtype=np.float64), np.arange(height, dtype=np.float64))
    cu = (u - K[0, 2]) / K[0, 0]
    cv = (v - K[1, 2]) / K[1, 1]
    A = cu * cu + cv * cv + 1.0
    R = quat_to_rot(pose7[3:])
    c = (spheres[:, :3] - pose7[:3]) @ R
    depth = np.full((height, width), np.inf)
    for s in range(spheres.shape[0]):
        B = -2.0 * (cu * c[s, 0] + cv * c[s, 1] + c[s, 2])
        C = float(c[s] @ c[s] - spheres[s, 3] ** 2)
        disc = B * B - 4.0 * A * C
        hit = disc >= 0
        z = np.full_like(depth, np.inf)
        z[hit] = 0.5 * (-B[hit] - np.sqrt(disc[hit])) / A[hit]
        z[z <= 0] = np.inf
        depth = np.minimum(depth, z)
    depth[np.isinf(depth)] = 0.0
    return depth


def add_kinect_noise(z: np.ndarray, rng: np.random.Generator, sigma_d: float = 0.5) -> np.ndarray:
    """Disparity-space noise + quantisation (add_kinect_noise.m:43-74)."""
    z = z.copy()
    mask = z > 0
    d = np.zeros_like(z)
    d[mask] = (3.0 - 1.0 / z[mask]) / 2.85e-3
    d[mask] += sigma_d * rng.standard_normal(z.shape)[mask]
    d = np.round(d)
    z_inv = -2.85e-3 * d + 3.0
    z[mask] = 1.0 / z_inv[mask]
    return z


def quantise(z: np.ndarray, unit: float = 1.0 / 1000.0) -> np.ndarray:
    """uint16 PNG round trip of a loader: uint16(z/unit) then float(v) * float(unit) (SynthLoader.h:57-83,
    ImageLoader.h:159-175).  unit = 1/5000 gives the TUM convention (TumrgbdLoader.h:68)."""
    q = np.clip(np.round(z / unit), 0, 65535).astype(np.uint16)
    return q.astype(np.float32) * np.float32(unit)


def make_sequence(n_frames: int, width: int = 640, height: int = 480, n_spheres: int = 5, noise: bool = True,
                  scene_seed: int = SCENE_SEED, noise_seed: int = NOISE_SEED, unit: float = 1.0 / 1000.0,
                  radius: float = 2.0, extent: float = 1.0, rmin: float = 0.0625, rmax: float = 0.5,
                  turn: float = 2 * np.pi):
    """Returns (depth [n,h,w] float32 metres, poses [n,7] float32 camera-to-world, K float32 3x3, spheres)."""
    K = scaled_intrinsics(width, height)
    spheres = random_spheres(n_spheres, scene_seed, extent, rmin, rmax)
    poses = orbit_poses(n_frames, radius=radius, turn=turn)
    rng = np.random.default_rng(noise_seed)
    frames = np.zeros((n_frames, height, width), dtype=np.float32)
    for k in range(n_frames):
        z = render_depth(spheres, poses[k], K, width, height)
        if noise:
            z = add_kinect_noise(z, rng)
        frames[k] = quantise(z, unit)
    return frames, poses.astype(np.float32), K.astype(np.float32), spheres


# --------------------------------------------------------------------------------------------------------------------
# Room-scale scene (BASELINE.json configs[3], "TUM fr1-style"): the camera moves inside a 6 x 6 x 3 m box room that holds
# ~20 spheres (furniture stand-ins), on a smooth hand-held-like path (<= 3 cm and <= 1.5 degrees between frames at 2000
# frames per lap).  Rendering works on numpy arrays or on torch tensors (pass xp=torch and a device) so that thousands of
# frames can be produced on the GPU for the bench; the arithmetic is the same analytic ray casting as render_depth.
ROOM_HALF = np.array([3.0, 3.0, 1.5])
ROOM_SEED = 4242


def room_scene(n_spheres: int = 20, seed: int = ROOM_SEED) -> np.ndarray:
    """Spheres [cx cy cz r] standing in the outer part of the room (between 1.6 m and 2.6 m from the vertical axis), not
    intersecting each other."""
    rng = np.random.default_rng(seed)
    out = []
    while len(out) < n_spheres:
        ang, rad = rng.uniform(0, 2 * np.pi), rng.uniform(1.6, 2.6)
        r = rng.uniform(0.2, 0.5)
        c = np.array([rad * np.cos(ang), rad * np.sin(ang), rng.uniform(-1.5 + r * 0.5, 0.9)])
        if all(np.linalg.norm(c - s[:3]) > s[3] + r + 0.05 for s in out):
            out.append(np.array([c[0], c[1], c[2], r]))
    return np.stack(out)


def room_trajectory(n_frames: int, laps: float = 1.0, frames_per_lap: int = 2000) -> np.ndarray:
    """Camera-to-world poses: the camera walks a 0.6 m circle about the room's axis looking outward, bobbing +-0.15 m and
    pitching +-12 degrees: 0.18 degrees and ~2 mm per frame at 2000 frames per lap."""
    poses = np.zeros((n_frames, 7))
    for k in range(n_frames):
        th = 2 * np.pi * k / frames_per_lap * laps
        pitch = np.deg2rad(12.0) * np.sin(5 * th)
        c = np.array([0.6 * np.cos(th), 0.6 * np.sin(th), 0.15 * np.sin(3 * th)])
        fwd = np.array([np.cos(th) * np.cos(pitch), np.sin(th) * np.cos(pitch), np.sin(pitch)])
        poses[k, :3] = c
        poses[k, 3:] = rot_to_quat(_look_at(c, c + fwd))
    return poses


def render_room(spheres: np.ndarray, pose7: np.ndarray, K: np.ndarray = K_KINECT, width: int = 640, height: int = 480,
                xp=np, device=None):
    """Noise-free z-depth of the box room + spheres seen from pose7 (float64; every pixel hits a wall)."""
    kw = {} if xp is np else {"device": device, "dtype": xp.float64}
    if xp is np:
        u, v = np.meshgrid(np.arange(width, dtype=np.float64), np.arange(height, dtype=np.float64))
    else:
        u, v = xp.meshgrid(xp.arange(width, **kw), xp.arange(height, **kw), indexing="xy")
    cu = (u - float(K[0, 2])) / float(K[0, 0])
    cv = (v - float(K[1, 2])) / float(K[1, 1])
    R = quat_to_rot(np.asarray(pose7[3:], np.float64))
    o = np.asarray(pose7[:3], np.float64)
    inf = float("inf")
    depth = None
    for axis in range(3):
        # world-space ray direction per unit z-depth: R @ [cu, cv, 1]
        dirn = R[axis, 0] * cu + R[axis, 1] * cv + R[axis, 2]
        for sign in (-1.0, 1.0):
            z = (sign * ROOM_HALF[axis] - o[axis]) / dirn
            z = xp.where(z > 0, z, xp.full_like(z, inf)) if xp is not np else np.where(z > 0, z, inf)
            depth = z if depth is None else (xp.minimum(depth, z))
    A = cu * cu + cv * cv + 1.0
    c = (spheres[:, :3] - o) @ R
    for s in range(spheres.shape[0]):
        B = -2.0 * (cu * c[s, 0] + cv * c[s, 1] + c[s, 2])
        C = float(c[s] @ c[s] - spheres[s, 3] ** 2)
        disc = B * B - 4.0 * A * C
        if xp is np:
            z = np.where(disc >= 0, 0.5 * (-B - np.sqrt(np.maximum(disc, 0))) / A, inf)
            z = np.where(z > 0, z, inf)
        else:
            z = xp.where(disc >= 0, 0.5 * (-B - xp.sqrt(disc.clamp(min=0))) / A, xp.full_like(A, inf))
            z = xp.where(z > 0, z, xp.full_like(z, inf))
        depth = xp.minimum(depth, z)
    return depth


def make_room_sequence(n_frames: int, width: int = 640, height: int = 480, noise: bool = True, laps: float = 1.0,
                       frames_per_lap: int = 2000, start: int = 0, noise_seed: int = NOISE_SEED, unit: float = 1.0 / 1000.0):
    """(depth [n,h,w] float32, poses [n,7] float32, K float32, spheres) of the room scene, frames start .. start+n-1 of the
    lap (numpy; for a few frames — the bench renders its frames with render_room on the GPU)."""
    K = scaled_intrinsics(width, height)
    spheres = room_scene()
    poses = room_trajectory(start + n_frames, laps, frames_per_lap)[start:]
    rng = np.random.default_rng(noise_seed)
    frames = np.zeros((n_frames, height, width), dtype=np.float32)
    for k in range(n_frames):
        z = render_room(spheres, poses[k], K, width, height)
        if noise:
            z = add_kinect_noise(z, rng)
        frames[k] = quantise(z, unit)
    return frames, poses.astype(np.float32), K.astype(np.float32), spheres
