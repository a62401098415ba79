"""The benchmark workloads of BASELINE.json (`configs`, SURVEY §8d A..E) as lists of trajectories, and the plan a
rank builds for its share of the target frames.

A workload is a list of independent trajectories (S sources -> T target frames each, one shared H x W and K).  Its
UNITS are the target frames of all trajectories in order; `sharding.partition_units` cuts them into contiguous
blocks, one per GPU (whole trajectories first, then frames: a contiguous cut of the concatenated frame list does
exactly that), and `build_plan` turns a block [f0, f1) into ONE `TrajectoryWarp` over the sources of the trajectories
it touches - a block-diagonal pair list, so several trajectories still run as a handful of launches.

  PAIR   every (source, target) pair of a trajectory is its own output (reference semantics)
  FUSED  all sources of a trajectory accumulate into one output per target frame (extension; config E)

The reference loop being split is /root/reference/model/diffusionGS.py:1554-1599 (:1415-1493 backward); the 8
trajectories of config D are its consecutive view pairs (:244-250).  Everything here is host-side NumPy + plan
construction; inputs are synthetic and seeded.
"""
from __future__ import annotations

import dataclasses
from typing import List, Tuple

import numpy as np
import torch

from . import synth, trajectory, warp as _w


@dataclasses.dataclass
class Workload:
    name: str
    mode: str                       # "pair" | "fused"
    consistency: bool               # reprojection-consistency masks per pair (PAIR mode only)
    H: int
    W: int
    trajs: List[Tuple[int, int]]    # (S, T) per trajectory
    pose_step: float = 1.0
    lazy_sources: bool = False      # generate source by source (config E: 4.25 GB of RGB-D)

    @property
    def frames(self) -> int:
        return sum(t for _, t in self.trajs)

    @property
    def pairs(self) -> int:
        return sum(s * t for s, t in self.trajs)

    def frame_owner(self, f: int) -> Tuple[int, int]:
        """global frame index -> (trajectory, target index inside it)"""
        for k, (_, t) in enumerate(self.trajs):
            if f < t:
                return k, f
            f -= t
        raise IndexError(f)

    def describe(self) -> str:
        n = len(self.trajs)
        S, T = self.trajs[0]
        what = (f"{n} trajectories x ({S} src -> {T} target poses)" if n > 1 else f"{S} src RGB-D f32 -> {T} target poses")
        if self.mode == "pair":
            out = "PAIR mode (%d pair warps), forward soft-z splat (rgb f32 + hole mask + winner idx)" % self.pairs
            if self.consistency:
                out += " + reprojection consistency (mask_reproj u8 + soft_mask f32) per pair"
        else:
            out = "FUSED mode (%d pair warps into %d frames), forward soft-z splat only (rgb f32 + hole mask + winner idx)" % (
                self.pairs, self.frames)
        return f"{self.name}: {what} at {self.H}x{self.W}, {out}"


def for_config(name: str, mode: str = None) -> Workload:
    S, T, H, W = synth.CONFIGS[name]
    if name == "D_dl3dv":
        wl = Workload(name, "pair", True, H, W, [(S, T)] * 8)
    elif name == "E_stress":
        wl = Workload(name, "fused", False, H, W, [(S, T)], pose_step=0.12, lazy_sources=True)
    else:
        wl = Workload(name, "pair", True, H, W, [(S, T)])
    if mode is not None and mode != wl.mode:
        wl = dataclasses.replace(wl, mode=mode, consistency=(mode == "pair"))
    return wl


# ---- synthetic data -----------------------------------------------------------------------------------------------

def lazy_source(H: int, W: int, s: int, seed: int):
    """Source view `s` of a lazily generated trajectory: (rgb [H,W,3] f32, depth [H,W] f32); same statistics as
    synth.make_trajectory, its own generator per source so that 4K x 32 views never sit in host memory at once."""
    rng = np.random.default_rng([seed, s])
    rgb = (rng.random((H, W, 3), dtype=np.float32) * np.float32(255.0)).astype(np.float32)
    return rgb, synth.depth_field(H, W, rng, phase=0.3 * s)


def lazy_cameras(S: int, T: int, H: int, W: int, pose_step: float):
    pose_src = np.stack([synth.source_pose(pose_step * s) for s in range(S)])
    pose_tgt = synth.interpolate_poses(pose_src[0], synth.source_pose(pose_step * max(S - 1, 1)), T)
    return synth.make_intrinsics(H, W), pose_src, pose_tgt


@dataclasses.dataclass
class RankShare:
    """What one rank holds for frames [f0, f1) of a workload."""
    plan: trajectory.TrajectoryWarp
    f0: int
    f1: int
    trajs: list                     # synth.Trajectory per touched trajectory (None when lazy)
    traj_ids: list
    src_base: list                  # first local source index of each touched trajectory
    frame_of_slot: np.ndarray       # global frame of every output slot (pair or frame)
    pair_global: list               # per local pair: (trajectory, source, target) in the workload's own numbering
    host_rgbd: torch.Tensor = None  # host staging of the inputs (pinned when the plan was built with pinned_host=True)
    host_tgt: torch.Tensor = None


def slots_per_frame(wl: Workload) -> List[int]:
    """Output slots of every target frame in order: S pair warps in PAIR mode, one fused frame in FUSED mode."""
    out = []
    for S, T in wl.trajs:
        out += [S if wl.mode == "pair" else 1] * T
    return out


def partition_slots(wl: Workload, world: int):
    """The workload's output slots (pair warps, target-major; frames in FUSED mode) cut into `world` contiguous blocks:
    [(q0, q1)] per rank.  In PAIR mode a (source, target) pair is the independent unit of work, so the cut
    may fall INSIDE a target frame: 25 frames x 3 sources over 8 GPUs is 10,10,10,9,9,9,9,9 pair warps instead of
    4,3,3,3,3,3,3,3 frames (12 pair warps on the heaviest rank).  FUSED mode: slots are frames (13,13,13,13,12,12,12,12)."""
    from . import sharding
    return sharding.partition_units(sum(slots_per_frame(wl)), world)


def frames_of_slots(wl: Workload, q0: int, q1: int):
    """(f0, f1, head, tail): the frames [f0, f1) that slots [q0, q1) touch, and how many slots of frame f0 / f1-1 lie
    OUTSIDE the range (to be trimmed from the frames' slot list)."""
    assert q0 < q1
    spf = slots_per_frame(wl)
    starts = np.concatenate([[0], np.cumsum(spf)])
    f0 = int(np.searchsorted(starts, q0, side="right") - 1)
    f1 = int(np.searchsorted(starts, q1, side="left"))
    return f0, f1, int(q0 - starts[f0]), int(starts[f1] - q1)


def build_plan(wl: Workload, f0: int, f1: int, seed: int = 0, trim=(0, 0), **plan_kw) -> RankShare:
    """Plan + synthetic inputs (uploaded, device resident; pinned host staging filled when pinned_host=True) for frames
    [f0, f1).  Trajectory k of the workload uses seed `seed + k`.  `trim` = (head, tail): that many pair warps of the
    first / last frame belong to a neighbouring rank (`frames_of_slots`; PAIR mode only)."""
    assert 0 <= f0 < f1 <= wl.frames
    H, W = wl.H, wl.W
    ids, base, local_T = [], [], []
    f = 0
    for k, (S, T) in enumerate(wl.trajs):
        a, b = max(f0, f), min(f1, f + T)
        if a < b:
            ids.append(k)
            local_T.append((a - f, b - f))
        f += T
    S_loc = sum(wl.trajs[k][0] for k in ids)
    T_loc = f1 - f0
    ps, pt, pair_global, frame_of_slot = [], [], [], []
    sb, tb = 0, 0
    for k, (t0, t1) in zip(ids, local_T):
        S = wl.trajs[k][0]
        base.append(sb)
        for t in range(t0, t1):                      # target-major: a pair range then needs few target depth maps
            for s in range(S):
                ps.append(sb + s)
                pt.append(tb + (t - t0))
                pair_global.append((k, s, t))
        sb += S
        tb += t1 - t0
    head, tail = int(trim[0]), int(trim[1])
    if head or tail:
        assert wl.mode == "pair" and head + tail < len(ps)
        keep = slice(head, len(ps) - tail)
        ps, pt, pair_global = ps[keep], pt[keep], pair_global[keep]
    ps, pt = np.asarray(ps, np.int32), np.asarray(pt, np.int32)
    fr0 = [sum(T for _, T in wl.trajs[:k]) for k in ids]
    if wl.mode == "fused":
        frame_of_slot = np.concatenate([np.arange(fr + a, fr + b) for fr, (a, b) in zip(fr0, local_T)])
    else:
        frame_of_slot = np.asarray([sum(T for _, T in wl.trajs[:k]) + t for (k, s, t) in pair_global])
    plan = trajectory.TrajectoryWarp(S_loc, T_loc, H, W, ps, pt, consistency=wl.consistency, fused=(wl.mode == "fused"),
                                     **plan_kw)
    trajs = []
    pose_src, pose_tgt, K = [], [], None
    stage_rgbd = plan.pinned["rgbd"] if plan.pinned is not None else torch.empty((S_loc, H, W, 4), dtype=torch.float32)
    stage_tgt = None
    if wl.consistency:
        stage_tgt = plan.pinned["tgt_depth"] if plan.pinned is not None else torch.empty((T_loc, H, W), dtype=torch.float32)
    tb = 0
    for k, sb, (t0, t1) in zip(ids, base, local_T):
        S, T = wl.trajs[k]
        if wl.lazy_sources:
            K, p_s, p_t = lazy_cameras(S, T, H, W, wl.pose_step)
            for s in range(S):
                rgb, dep = lazy_source(H, W, s, seed + k)
                v = stage_rgbd[sb + s].numpy()
                v[..., :3] = rgb
                v[..., 3] = dep
            trajs.append(None)
        else:
            tr = synth.make_trajectory(S, T, H, W, seed=seed + k, pose_step=wl.pose_step)
            K, p_s, p_t = tr.K, tr.pose_src, tr.pose_tgt
            stage_rgbd[sb:sb + S].copy_(torch.from_numpy(_w.pack_rgbd(tr.rgb, tr.depth)))
            if wl.consistency:
                stage_tgt[tb:tb + (t1 - t0)].copy_(torch.from_numpy(tr.depth_tgt[t0:t1]))
            trajs.append(tr)
        pose_src.append(p_s)
        pose_tgt.append(p_t[t0:t1])
        tb += t1 - t0
    plan.set_cameras(np.concatenate(pose_src), np.concatenate(pose_tgt), K)
    plan.upload(stage_rgbd, stage_tgt)
    return RankShare(plan, f0, f1, trajs, ids, base, frame_of_slot, pair_global, stage_rgbd, stage_tgt)
