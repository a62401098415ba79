"""Seeded synthetic RGB-D trajectories for tests, golden fixtures and bench.py.

Shapes and statistics follow SURVEY.md §8(d): RGB uniform f32 in [0,255] (worst case for colour
parity), depth = smooth field + a rectangular foreground occluder (real depth competition, ~10 %
holes), pinhole K with a NON-integer principal point (N3: exact-integer landings are a
discontinuity of the reference and are tested separately), float32 world-to-camera poses along a
Slerp + natural-cubic path like the reference's `pose_interpolation`
(/root/reference/model/diffusionGS.py:1208-1234).

Nothing here touches the GPU or the oracle; everything is NumPy on the host.
"""
from __future__ import annotations

import dataclasses
import numpy as np

# name -> (S sources, T targets, H, W)   (BASELINE.json `configs`, SURVEY §8d A..E)
CONFIGS = {
    "A_llff": (3, 1, 378, 504),
    "B_svd": (3, 25, 576, 1024),
    "C_dtu": (3, 64, 300, 400),
    "D_dl3dv": (2, 25, 540, 960),      # one of the 8 trajectories
    "E_stress": (32, 100, 2160, 3840),
}


@dataclasses.dataclass
class Trajectory:
    rgb: np.ndarray        # [S,H,W,3] f32 in [0,255]
    depth: np.ndarray      # [S,H,W]   f32 z-depth
    K: np.ndarray          # [3,3]     f64 pixel-unit pinhole
    pose_src: np.ndarray   # [S,4,4]   f32 w2c
    pose_tgt: np.ndarray   # [T,4,4]   f32 w2c
    depth_tgt: np.ndarray  # [T,H,W]   f32 pseudo depth of the target views (consistency pass)

    @property
    def shape(self):
        S, H, W = self.depth.shape
        return S, self.pose_tgt.shape[0], H, W


def euler_deg_to_R(rx, ry, rz):
    rx, ry, rz = np.deg2rad([rx, ry, rz])
    cx, sx, cy, sy, cz, sz = np.cos(rx), np.sin(rx), np.cos(ry), np.sin(ry), np.cos(rz), np.sin(rz)
    Rx = np.array([[1, 0, 0], [0, cx, -sx], [0, sx, cx]])
    Ry = np.array([[cy, 0, sy], [0, 1, 0], [-sy, 0, cy]])
    Rz = np.array([[cz, -sz, 0], [sz, cz, 0], [0, 0, 1]])
    return Rz @ Ry @ Rx


def source_pose(k: float) -> np.ndarray:
    """k-th source w2c pose: off-axis Euler (1.3,2.0,-0.7) deg * k, translation (0.1,0.02,-0.03)*k."""
    T = np.eye(4)
    T[:3, :3] = euler_deg_to_R(1.3 * k, 2.0 * k, -0.7 * k)
    T[:3, 3] = np.array([0.1, 0.02, -0.03]) * k
    return T.astype(np.float32)


def interpolate_poses(pose_a: np.ndarray, pose_b: np.ndarray, n: int) -> np.ndarray:
    """n w2c poses from pose_a to pose_b: Slerp rotations + natural cubic translations, cast f32
    (same recipe as diffusionGS.py:1208-1234, any n instead of the fixed 25)."""
    from scipy.interpolate import CubicSpline
    from scipy.spatial.transform import Rotation, Slerp

    if n == 1:
        ts = np.array([0.5])
    else:
        ts = np.linspace(0.0, 1.0, n)
    key = np.array([0.0, 1.0])
    rots = Rotation.from_matrix(np.stack([pose_a[:3, :3], pose_b[:3, :3]]).astype(np.float64))
    Rm = Slerp(key, rots)(ts).as_matrix()
    tr = CubicSpline(key, np.stack([pose_a[:3, 3], pose_b[:3, 3]]).astype(np.float64), bc_type="natural")(ts)
    out = np.tile(np.eye(4), (n, 1, 1))
    out[:, :3, :3] = Rm
    out[:, :3, 3] = tr
    return out.astype(np.float32)


def depth_field(H: int, W: int, rng: np.random.Generator, phase: float = 0.0) -> np.ndarray:
    y, x = np.mgrid[0:H, 0:W].astype(np.float64)
    d = 2.0 + 0.5 * np.sin(x / 37.0 + phase) + 0.5 * np.cos(y / 23.0 - phase) + 0.05 * rng.random((H, W))
    # foreground occluder (~12 % of the frame) at depth 0.9
    y0, y1 = int(0.30 * H), int(0.30 * H + 0.35 * H)
    x0, x1 = int(0.40 * W), int(0.40 * W + 0.35 * W)
    d[y0:y1, x0:x1] = 0.9
    return d.astype(np.float32)


def make_intrinsics(H: int, W: int) -> np.ndarray:
    f = 0.8 * W
    return np.array([[f, 0.0, W / 2 - 0.37], [0.0, f, H / 2 + 0.21], [0.0, 0.0, 1.0]], dtype=np.float64)


def make_trajectory(S: int, T: int, H: int, W: int, seed: int = 0, smooth_rgb: bool = False,
                    pose_step: float = 1.0) -> Trajectory:
    rng = np.random.default_rng(seed)
    if smooth_rgb:
        y, x = np.mgrid[0:H, 0:W].astype(np.float32)
        rgb1 = np.stack([255 * x / max(W - 1, 1), 255 * y / max(H - 1, 1), 127.5 + 0 * x], -1)
        rgb = np.stack([np.roll(rgb1, s, axis=-1) for s in range(S)]).astype(np.float32)
    else:
        rgb = (rng.random((S, H, W, 3), dtype=np.float32) * np.float32(255.0)).astype(np.float32)
    depth = np.stack([depth_field(H, W, rng, phase=0.3 * s) for s in range(S)])
    K = make_intrinsics(H, W)
    pose_src = np.stack([source_pose(pose_step * s) for s in range(S)])
    pose_tgt = interpolate_poses(pose_src[0], source_pose(pose_step * max(S - 1, 1)), T)
    depth_tgt = np.stack([depth_field(H, W, rng, phase=0.1 + 0.01 * t) * np.float32(1.01) for t in range(T)])
    return Trajectory(rgb, depth, K, pose_src, pose_tgt, depth_tgt)


def all_pairs(S: int, T: int):
    """PAIR mode: every (source, target) pair, source-major."""
    ps = np.repeat(np.arange(S, dtype=np.int32), T)
    pt = np.tile(np.arange(T, dtype=np.int32), S)
    return ps, pt


def target_major(ps, pt):
    """The same pairs ordered by target first: a pair range then touches few target depth maps (pipelined upload)."""
    order = np.lexsort((ps, pt))
    return np.ascontiguousarray(ps[order]), np.ascontiguousarray(pt[order])


def reference_policy_pairs(T: int, S: int = 2):
    """The reference's own source policy: target i<12 from the left anchor, else the right one
    (diffusionGS.py:1555-1567)."""
    pt = np.arange(T, dtype=np.int32)
    ps = np.where(pt < 12, 0, S - 1).astype(np.int32)
    return ps, pt
