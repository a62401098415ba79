"""Multi-GPU plumbing of the geometric-conditioning path: one process per GPU, frames partitioned, one gather.

The reference warps the frames of a trajectory in a sequential Python loop on one device
(/root/reference/model/diffusionGS.py:1554-1599, :1415-1493) and has no distributed code on this path
(SURVEY §0 R8).  Target frames (and whole trajectories) are independent units, so the B200 design is:

  * `partition_frames` / `partition_units` — contiguous block partition of the T target frames (or of whole
    trajectories) over the ranks; sources, poses and K are replicated;
  * every rank runs its own `TrajectoryWarp` over its frames — NO collective on the data path;
  * `gather_frames` — the only cross-GPU traffic: the finished frames and masks go to the rank that runs the
    diffusion passes, as one batch of point-to-point transfers (`torch.distributed.batch_isend_irecv`, which the
    NCCL backend turns into one grouped ncclSend/ncclRecv over NVLink; sizes may differ per rank).
    The wire format is whatever tensors the caller selects — `compact_wire` picks what the diffusion side
    consumes (uint8 colour + the two 0/1 masks, 5 B/px instead of 22 B/px).

Everything here works on tensors of any device, so the logic is tested on CPU with the gloo backend
(tests/test_sharding_gloo.py); on the GPU box the same code runs over NCCL (bench.py --gpus N).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist


def partition_units(n: int, world: int) -> List[Tuple[int, int]]:
    """Contiguous block partition of n independent units over `world` ranks: [(begin, end)] per rank; the first
    n % world ranks get one more (100 frames over 8 GPUs -> 13,13,13,13,12,12,12,12)."""
    assert n >= 0 and world >= 1
    base, extra = divmod(n, world)
    out, b = [], 0
    for r in range(world):
        e = b + base + (1 if r < extra else 0)
        out.append((b, e))
        b = e
    return out


def partition_frames(T: int, world: int, rank: int) -> Tuple[int, int]:
    """[t0, t1) of the target frames rank `rank` warps."""
    return partition_units(T, world)[rank]


def pairs_for_frames(S: int, t0: int, t1: int, policy: str = "all"):
    """(pair_src, pair_tgt) of a rank's frames; pair_tgt is LOCAL (0 .. t1-t0).  'all' = every source for every
    frame (PAIR mode); 'reference_policy' = frame i < 12 from the left anchor, else the right one
    (diffusionGS.py:1555-1567), decided on the GLOBAL frame index."""
    import numpy as np
    n = t1 - t0
    if policy == "all":
        ps = np.repeat(np.arange(S, dtype=np.int32), n)
        pt = np.tile(np.arange(n, dtype=np.int32), S)
    elif policy == "reference_policy":
        pt = np.arange(n, dtype=np.int32)
        ps = np.where(pt + t0 < 12, 0, S - 1).astype(np.int32)
    else:
        raise ValueError(policy)
    return ps, pt


def compact_wire(out_fwd: Dict[str, torch.Tensor], out_inv: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """The tensors the diffusion rank consumes: uint8 warped colour, hole mask, reprojection mask (+ the soft
    mask when present).  Leading axis = the rank's own units (pairs)."""
    wire = {}
    if "rgb_u8" in out_fwd:
        wire["rgb_u8"] = out_fwd["rgb_u8"]
    elif "rgb_f32" in out_fwd:
        wire["rgb_f32"] = out_fwd["rgb_f32"]
    wire["mask"] = out_fwd["mask"]
    for k in ("mask_reproj", "soft_mask_reproj"):
        if k in out_inv:
            wire[k] = out_inv[k]
    return wire


def wire_bytes(wire: Dict[str, torch.Tensor]) -> int:
    return sum(t.numel() * t.element_size() for t in wire.values())


class FrameGather:
    """Gather of per-rank result tensors (leading axis = that rank's units) on rank `dst`.

    `counts[r]` = number of leading-axis rows rank r contributes.  The destination owns one preallocated buffer
    per key of shape [sum(counts), ...]; its own rows are copied in place, every other rank's rows arrive by
    point-to-point transfers issued as ONE batch."""

    def __init__(self, template: Dict[str, torch.Tensor], counts: Sequence[int], dst: int = 0, group=None):
        self.counts = [int(c) for c in counts]
        self.dst, self.group = dst, group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        assert len(self.counts) == self.world
        self.offsets = [0]
        for c in self.counts:
            self.offsets.append(self.offsets[-1] + c)
        self.keys = sorted(template.keys())
        self.buffers: Optional[Dict[str, torch.Tensor]] = None
        if self.rank == dst:
            self.buffers = {k: torch.empty((self.offsets[-1],) + tuple(template[k].shape[1:]), dtype=template[k].dtype,
                                           device=template[k].device) for k in self.keys}

    def bytes_received(self, template: Dict[str, torch.Tensor]) -> int:
        """Bytes that cross the interconnect into `dst` per gather."""
        per_row = sum(int(template[k][0].numel()) * template[k].element_size() for k in self.keys)
        return per_row * (self.offsets[-1] - self.counts[self.dst])

    def __call__(self, local: Dict[str, torch.Tensor]) -> Optional[Dict[str, torch.Tensor]]:
        """Blocking from the caller's point of view only in stream order: the transfers are enqueued and waited on
        (NCCL: on the current stream).  Returns the gathered buffers on `dst`, None elsewhere."""
        for k in self.keys:
            assert local[k].shape[0] == self.counts[self.rank], (k, tuple(local[k].shape), self.counts[self.rank])
        if self.world == 1:
            for k in self.keys:
                self.buffers[k].copy_(local[k])
            return self.buffers
        ops = []
        if self.rank == self.dst:
            for r in range(self.world):
                lo, hi = self.offsets[r], self.offsets[r + 1]
                if hi == lo:
                    continue
                for k in self.keys:
                    if r == self.rank:
                        self.buffers[k][lo:hi].copy_(local[k])
                    else:
                        ops.append(dist.P2POp(dist.irecv, self.buffers[k][lo:hi], self._global(r), self.group))
        elif self.counts[self.rank] > 0:
            for k in self.keys:
                ops.append(dist.P2POp(dist.isend, local[k].contiguous(), self._global(self.dst), self.group))
        if ops:
            for w in dist.batch_isend_irecv(ops):
                w.wait()
        return self.buffers

    def _global(self, group_rank: int) -> int:
        return dist.get_global_rank(self.group, group_rank) if self.group is not None else group_rank
