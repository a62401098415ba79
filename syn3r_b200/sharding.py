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


def partition_units(n: int, world: int, light: int = None) -> List[Tuple[int, int]]:
    """Contiguous block partition of n independent units over `world` ranks: [(begin, end)] per rank; the first
    n % world ranks get one more (100 frames over 8 GPUs -> 13,13,13,13,12,12,12,12).
    `light` = a rank that should carry a small share - the diffusion rank, whose HBM also takes every peer's stores during
    the step (§5 of DESIGN.md): the extra units then go to the ranks farthest from it (light=0: 12,12,12,12,13,13,13,13)."""
    assert n >= 0 and world >= 1
    base, extra = divmod(n, world)
    if light is None:
        heavy = set(range(extra))
    else:
        order = sorted(range(world), key=lambda r: (-min((r - light) % world, (light - r) % world), r))
        heavy = set(order[:extra])
    out, b = [], 0
    for r in range(world):
        e = b + base + (1 if r in heavy else 0)
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
    mask when present).  Leading axis = the rank's own units (pairs).  A plan built with pack_masks=True ships the two
    0/1 masks as bits (`mask_bits`, `mask_reproj_bits`: np.packbits order, 1 bit per pixel instead of 8)."""
    wire = {}
    if "rgb_u8" in out_fwd:
        wire["rgb_u8"] = out_fwd["rgb_u8"]
    elif "rgb_f32" in out_fwd:
        wire["rgb_f32"] = out_fwd["rgb_f32"]
    if "mask_bits" in out_fwd:
        wire["mask_bits"] = out_fwd["mask_bits"]
    else:
        wire["mask"] = out_fwd["mask"]
    if "mask_reproj_bits" in out_inv:
        wire["mask_reproj_bits"] = out_inv["mask_reproj_bits"]
    elif "mask_reproj" in out_inv:
        wire["mask_reproj"] = out_inv["mask_reproj"]
    if "soft_mask_reproj" in out_inv:
        wire["soft_mask_reproj"] = out_inv["soft_mask_reproj"]
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


class PipelinedGather:
    """The gather of `FrameGather`, issued per CHUNK of a rank's rows on a side stream so that the transfer of one chunk
    overlaps the kernels of the next ("per-frame ncclSend/Recv on a side stream").

    Every rank cuts its own rows into the same number of chunks; chunk c of every sender is matched by one batch of
    receives on `dst`.  Usage per step, on every rank:

        g.begin()
        for c in range(g.n_chunks):
            a, b = g.bounds(c)            # this rank's rows of chunk c
            ... enqueue the kernels that produce rows [a, b) on the current stream ...
            g.push(c, local)              # records an event, sends / receives rows [a, b) behind it on the side stream
        g.end()                           # the current stream waits for the side stream (gather complete on `dst`)
    """

    def __init__(self, template: Dict[str, torch.Tensor], counts: Sequence[int], n_chunks: int, dst: int = 0, group=None):
        self.base = FrameGather(template, counts, dst, group)
        self.n_chunks = max(1, int(n_chunks))
        b = self.base
        self._bounds = [[(b.counts[r] * c) // self.n_chunks for c in range(self.n_chunks + 1)] for r in range(b.world)]
        dev = next(iter(template.values())).device
        self.side = torch.cuda.Stream(device=dev) if dev.type == "cuda" else None

    @property
    def buffers(self):
        return self.base.buffers

    def bounds(self, c: int) -> Tuple[int, int]:
        r = self.base.rank
        return self._bounds[r][c], self._bounds[r][c + 1]

    def bytes_received(self, template) -> int:
        return self.base.bytes_received(template)

    def begin(self):
        if self.side is not None:
            self.side.wait_stream(torch.cuda.current_stream())

    def push(self, c: int, local: Dict[str, torch.Tensor]):
        b = self.base
        ev = None
        if self.side is not None:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream())
        ctx = torch.cuda.stream(self.side) if self.side is not None else _null()
        with ctx:
            if ev is not None:
                self.side.wait_event(ev)
            ops = []
            if b.rank == b.dst:
                for r in range(b.world):
                    lo, hi = self._bounds[r][c], self._bounds[r][c + 1]
                    if hi == lo:
                        continue
                    for k in b.keys:
                        dstv = b.buffers[k][b.offsets[r] + lo:b.offsets[r] + hi]
                        if r == b.rank:
                            if dstv.data_ptr() != local[k][lo:hi].data_ptr():      # a plan may already write in place
                                dstv.copy_(local[k][lo:hi], non_blocking=True)
                        else:
                            ops.append(dist.P2POp(dist.irecv, dstv, b._global(r), b.group))
            else:
                lo, hi = self._bounds[b.rank][c], self._bounds[b.rank][c + 1]
                if hi > lo:
                    for k in b.keys:
                        ops.append(dist.P2POp(dist.isend, local[k][lo:hi], b._global(b.dst), b.group))
            if ops:
                for wk in dist.batch_isend_irecv(ops):
                    wk.wait()

    def end(self):
        if self.side is not None:
            torch.cuda.current_stream().wait_stream(self.side)
        return self.base.buffers


class _null:
    def __enter__(self):
        return self

    def __exit__(self, *a):
        return False


def bitpack_masks_wire(wire: Dict[str, torch.Tensor]) -> Dict[str, torch.Tensor]:
    """Names of the 0/1 byte masks in a wire (they could travel as bits: 8x smaller); used for reporting only."""
    return {k: v for k, v in wire.items() if v.dtype == torch.uint8 and k.startswith("mask")}


class PeerOutputs:
    """Outputs of every rank written STRAIGHT into the diffusion rank's memory: one symmetric allocation per key
    (`torch.distributed._symmetric_memory`, CUDA IPC / fabric handles over NVLink), rank r's plan gets views of rank
    `dst`'s buffer at its own row offset, so the resolve step of the tile kernel and the consistency kernel store over
    NVLink while they compute - no staging copy, no send/recv kernel; a barrier at the end makes the rows visible.
    Raises if symmetric memory is not available on the box (callers fall back to the NCCL gather)."""

    def __init__(self, template: Dict[str, torch.Tensor], counts: Sequence[int], dst: int = 0):
        import torch.distributed._symmetric_memory as symm
        self.rank, self.world, self.dst = dist.get_rank(), dist.get_world_size(), dst
        self.counts = [int(c) for c in counts]
        self.offsets = [0]
        for c in self.counts:
            self.offsets.append(self.offsets[-1] + c)
        total = self.offsets[-1]
        self.keys = sorted(template.keys())
        self.local_views, self.buffers, self._hdl = {}, {}, {}
        gname = dist.group.WORLD.group_name
        for k in self.keys:
            t = template[k]
            shape = (total,) + tuple(t.shape[1:])
            mine = symm.empty(shape, dtype=t.dtype, device=t.device)
            hdl = symm.rendezvous(mine, gname)
            remote = hdl.get_buffer(dst, shape, t.dtype)
            self._hdl[k] = hdl
            lo, hi = self.offsets[self.rank], self.offsets[self.rank + 1]
            self.local_views[k] = remote[lo:hi]
            if self.rank == dst:
                self.buffers[k] = mine

    def bytes_received(self, template) -> int:
        per_row = sum(int(template[k][0].numel()) * template[k].element_size() for k in self.keys)
        return per_row * (self.offsets[-1] - self.counts[self.dst])

    def barrier(self):
        """All ranks' stores are complete and visible on `dst` (stream-ordered barrier of the symmetric-memory handle)."""
        self._hdl[self.keys[0]].barrier()
