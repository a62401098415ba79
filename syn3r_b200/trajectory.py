"""Whole-trajectory geometric conditioning: S posed RGB-D sources -> T target poses in one batched call.

This is the batched form of the per-frame Python loops of the reference
(/root/reference/model/diffusionGS.py:1554-1599 `warp_images`, :1415-1493 `warp_images_bw`): every
(source, target) pair is one `forward_warp` call (PAIR mode, reference semantics) plus one
`inverse_warp`-style consistency check, all pairs in a handful of kernel launches.

`TrajectoryWarp` owns the device buffers (inputs, outputs, workspace) and pinned host staging, so a
step is: H2D of the step's inputs -> kernels -> D2H of the results, with nothing allocated per step.
"""
from __future__ import annotations

import dataclasses
from typing import Optional, Sequence

import numpy as np
import torch

from . import warp as _w


@dataclasses.dataclass
class TrajectoryOutputs:
    rgb_f32: Optional[torch.Tensor]       # [P,H,W,3] f32  normalised warped colour (0..255 scale)
    rgb_u8: Optional[torch.Tensor]        # [P,H,W,3] u8
    mask: torch.Tensor                    # [P,H,W]   u8   hole mask (1 = covered)
    winner: Optional[torch.Tensor]        # [P,H,W]   i32  winning source pixel, -1 = hole
    mask_reproj: Optional[torch.Tensor]   # [P,H,W]   u8   reprojection-consistency mask
    soft_mask_reproj: Optional[torch.Tensor]  # [P,H,W] f32
    status: torch.Tensor                  # [P] i32


class TrajectoryWarp:
    """Device-resident plan for one trajectory shape (S sources, pairs, H, W)."""

    def __init__(self, S: int, T: int, H: int, W: int, pair_src: Sequence[int], pair_tgt: Sequence[int], *,
                 device=None, consistency: bool = True, want_f32: bool = True, want_u8: bool = False,
                 want_winner: bool = True, bandwidth: float = 20.0, pinned_host: bool = False,
                 host_outputs: Optional[Sequence[str]] = None, post: bool = False, fused: bool = False,
                 pack_masks: bool = False):
        """fused=True (extension): the pairs of a target frame share ONE accumulator, outputs are per target frame
        ([T,H,W,...]); the pairs must be ordered by target (synth.target_major) and there is no per-pair consistency pass."""
        self.S, self.T, self.H, self.W = S, T, H, W
        self.pair_src_h = np.asarray(pair_src, np.int32)
        self.pair_tgt_h = np.asarray(pair_tgt, np.int32)
        self.P = len(self.pair_src_h)
        self.fused = bool(fused)
        if self.fused:
            assert not consistency, "per-pair consistency masks are undefined in FUSED mode (SURVEY §8d config E)"
            assert np.all(np.diff(self.pair_tgt_h) >= 0), "FUSED mode: order the pairs by target frame"
        self.N = T if self.fused else self.P          # output slots: target frames (FUSED) or pairs (PAIR)
        self.dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.consistency, self.bandwidth = consistency, float(bandwidth)
        self.want_f32, self.want_u8, self.want_winner = want_f32, want_u8, want_winner
        d, P = self.dev, self.N
        self.rgbd = torch.empty((S, H, W, 4), dtype=torch.float32, device=d)
        self.tgt_depth = torch.empty((T, H, W), dtype=torch.float32, device=d) if consistency else None
        self.Kinv_src = torch.empty((S, 9), dtype=torch.float64, device=d)
        self.pair_par = torch.empty((self.P, 24), dtype=torch.float64, device=d)
        self.pair_P = torch.empty((self.P, 2, 16), dtype=torch.float32, device=d) if consistency else None
        self.pair_src = torch.from_numpy(self.pair_src_h).to(d)
        self.pair_tgt = torch.from_numpy(self.pair_tgt_h).to(d)
        self._slot_cache = {}
        self._ws_fwd = [None]              # the plan's own forward workspace (kernel state: counters, boxes, redo list)
        self.K_host = None
        self.out_fwd = {}
        self.out_inv = {}
        # outputs (allocated once)
        if want_f32:
            self.out_fwd["rgb_f32"] = torch.empty((P, H, W, 3), dtype=torch.float32, device=d)
        if want_u8:
            self.out_fwd["rgb_u8"] = torch.empty((P, H, W, 3), dtype=torch.uint8, device=d)
        self.out_fwd["mask"] = torch.empty((P, H, W), dtype=torch.uint8, device=d)
        if want_winner:
            self.out_fwd["winner"] = torch.empty((P, H, W), dtype=torch.int32, device=d)
        self.out_fwd["status"] = torch.empty((P,), dtype=torch.int32, device=d)
        if consistency:
            self.out_inv["mask_reproj"] = torch.empty((P, H, W), dtype=torch.uint8, device=d)
            self.out_inv["soft_mask_reproj"] = torch.empty((P, H, W), dtype=torch.float32, device=d)
        # bit-packed copies of the 0/1 masks (np.packbits order) for the trips over PCIe / NVLink: 1 bit per pixel
        self.pack_masks = bool(pack_masks)
        if self.pack_masks:
            assert (H * W) % 16 == 0, "pack_masks needs H * W to be a multiple of 16 (whole bytes per slot, aligned loads)"
            self.out_fwd["mask_bits"] = torch.empty((P, H * W // 8), dtype=torch.uint8, device=d)
            if consistency:
                self.out_inv["mask_reproj_bits"] = torch.empty((P, H * W // 8), dtype=torch.uint8, device=d)
        # the callers' per-frame post-processing (diffusionGS.py:1572-1597, :1452-1491) as part of every run
        self.post = bool(post)
        self.out_post = {}
        if self.post:
            self._alloc_post()
        # outputs that travel to the host in step_host()/download(); None = every output.  The reference's own
        # return values are ("rgb_u8", "mask") from forward_warp (forward_warp.py:182) and the consumed keys
        # ("mask_reproj", "soft_mask_reproj") of inverse_warp (diffusionGS.py:1346-1347); the rest are extensions
        # (fp32 colour, winner index) that device-side consumers read in place.
        allk = list(self.out_fwd) + list(self.out_inv) + list(self.out_post)
        self.host_outputs = allk if host_outputs is None else [k for k in allk if k in set(host_outputs) | {"status"}]
        assert host_outputs is None or set(host_outputs) <= set(allk), f"unknown output in {host_outputs}"
        self.pinned = None
        if pinned_host:
            self._alloc_pinned()

    # ---- sizes ---------------------------------------------------------------------------------
    def algorithmic_bytes(self) -> int:
        """inputs once + outputs once, accumulators zero."""
        hw = self.H * self.W
        out_px = (12 if self.want_f32 else 0) + (3 if self.want_u8 else 0) + 1 + (4 if self.want_winner else 0)
        b = self.S * hw * 16 + self.N * hw * out_px
        if self.consistency:
            b += self.T * hw * 4 + self.P * hw * 5
        return b

    # ---- parameters ----------------------------------------------------------------------------
    def set_cameras(self, pose_src, pose_tgt, K, K_tgt=None):
        """Host-side small-matrix preparation + upload (poses in the caller's dtype, K f64 3x3 or [S,3,3])."""
        Kinv, par = _w.forward_pair_params(pose_src, pose_tgt, K, K_tgt, self.pair_src_h, self.pair_tgt_h)
        self.Kinv_src.copy_(torch.from_numpy(Kinv))
        self.pair_par.copy_(torch.from_numpy(par))
        if self.consistency:
            K2d = np.asarray(K)
            assert K2d.ndim == 2, "the backward path shares one K between all views (forward_warp.py:187)"
            self.K_host = K2d.astype(np.float32)
            self.pair_P.copy_(torch.from_numpy(
                _w.inverse_pair_params(pose_src, pose_tgt, self.pair_src_h, self.pair_tgt_h)))

    # ---- data ----------------------------------------------------------------------------------
    def _alloc_pinned(self):
        S, T, H, W, P = self.S, self.T, self.H, self.W, self.N
        pin = lambda shape, dt: torch.empty(shape, dtype=dt, pin_memory=True)  # noqa: E731
        self.pinned = {"rgbd": pin((S, H, W, 4), torch.float32)}
        if self.consistency:
            self.pinned["tgt_depth"] = pin((T, H, W), torch.float32)
        for k, v in self._host_items():
            self.pinned["out_" + k] = pin(tuple(v.shape), v.dtype)

    def _alloc_post(self):
        assert self.want_u8, "the post-processing blanks the uint8 frame: construct with want_u8=True"
        assert self.H % 8 == 0 and self.W % 8 == 0, "post-processing needs H and W to be multiples of 8"
        P, H, W, d = self.N, self.H, self.W, self.dev
        self.out_post = {"post_mask_png": torch.empty((P, H, W), dtype=torch.uint8, device=d),
                         "post_cond_u8": torch.empty((P, H, W, 3), dtype=torch.uint8, device=d),
                         "post_latent": torch.empty((P, H // 8, W // 8), dtype=torch.uint8, device=d)}
        if self.consistency:
            self.out_post["post_latent_reproj"] = torch.empty((P, H // 8, W // 8), dtype=torch.uint8, device=d)
            self.out_post["post_soft_pooled"] = torch.empty((P, H // 8, W // 8), dtype=torch.float32, device=d)

    def _post(self, sl=slice(None)):
        from . import conditioning as _c
        o = self.out_post
        _c.condition_post_batched(self.out_fwd["mask"][sl], img_u8=self.out_fwd["rgb_u8"][sl],
                                  want=("mask_png", "cond_u8", "latent"),
                                  out={"mask_png": o["post_mask_png"][sl], "cond_u8": o["post_cond_u8"][sl],
                                       "latent": o["post_latent"][sl]})
        if self.consistency:
            _c.condition_post_batched(self.out_inv["mask_reproj"][sl], soft=self.out_inv["soft_mask_reproj"][sl],
                                      want=("latent", "soft_pooled"),
                                      out={"latent": o["post_latent_reproj"][sl], "soft_pooled": o["post_soft_pooled"][sl]})

    def _host_items(self):
        both = {**self.out_fwd, **self.out_inv, **self.out_post}
        return [(k, both[k]) for k in self.host_outputs]

    def upload(self, rgbd_host: torch.Tensor, tgt_depth_host: Optional[torch.Tensor] = None):
        """Async H2D of one step's inputs (pinned host tensors recommended)."""
        self.rgbd.copy_(rgbd_host, non_blocking=True)
        if self.consistency:
            self.tgt_depth.copy_(tgt_depth_host, non_blocking=True)

    def h2d_bytes(self) -> int:
        return self.rgbd.numel() * 4 + (self.tgt_depth.numel() * 4 if self.consistency else 0)

    def d2h_bytes(self) -> int:
        return sum(v.numel() * v.element_size() for _, v in self._host_items())

    def download(self):
        """Async D2H of the host outputs into the pinned staging buffers."""
        assert self.pinned is not None, "construct with pinned_host=True"
        for k, v in self._host_items():
            self.pinned["out_" + k].copy_(v, non_blocking=True)
        return self.pinned

    # ---- the hot path --------------------------------------------------------------------------
    def _forward(self, sl=slice(None), phases=0):
        out_f = self.out_fwd if sl == slice(None) else {k: v[sl] for k, v in self.out_fwd.items()}
        if self.fused:                      # sl is a range of SLOTS (target frames); its pairs are contiguous (target-major)
            t0, t1 = (0, self.T) if sl == slice(None) else (sl.start, sl.stop)
            key = (t0, t1)
            if key not in self._slot_cache:
                p0, p1 = (int(v) for v in np.searchsorted(self.pair_tgt_h, [t0, t1]))
                self._slot_cache[key] = (p0, p1, (self.pair_tgt[p0:p1] - t0).contiguous())
            p0, p1, slot = self._slot_cache[key]
            _w.forward_warp_batched(self.rgbd, self.Kinv_src, self.pair_par[p0:p1], self.pair_src[p0:p1], pair_slot=slot,
                                    n_slots=t1 - t0, want_f32=self.want_f32, want_u8=self.want_u8,
                                    want_winner=self.want_winner, is_image=True, out=out_f, phases=phases,
                                    workspace=self._ws_fwd)
            return
        _w.forward_warp_batched(self.rgbd, self.Kinv_src, self.pair_par[sl], self.pair_src[sl], want_f32=self.want_f32,
                                want_u8=self.want_u8, want_winner=self.want_winner, is_image=True, out=out_f,
                                phases=phases, workspace=self._ws_fwd)

    def _inverse(self, sl=slice(None)):
        out_i = self.out_inv if sl == slice(None) else {k: v[sl] for k, v in self.out_inv.items()}
        _w.inverse_warp_batched(None, None, self.tgt_depth, self.K_host, self.pair_P[sl], self.pair_src[sl],
                                self.pair_tgt[sl], rgbd=self.rgbd, bandwidth=self.bandwidth,
                                want=("mask_reproj", "soft_mask_reproj"), out=out_i)

    def _pack(self, sl, fwd: bool):
        from . import conditioning as _c
        if fwd:
            _c.pack_mask_bits(self.out_fwd["mask"][sl], out=self.out_fwd["mask_bits"][sl].view(-1))
        else:
            _c.pack_mask_bits(self.out_inv["mask_reproj"][sl], out=self.out_inv["mask_reproj_bits"][sl].view(-1))

    def _run_range(self, p0: int, p1: int, before_inverse=None):
        """Forward warp (+ consistency masks) of output slots [p0, p1) (pairs; target frames in FUSED mode) into the
        matching slices of the outputs.  `before_inverse()` runs between the two (step_host: wait for the target depth)."""
        sl = slice(None) if (p0 == 0 and p1 == self.N) else slice(p0, p1)
        # (A/B, round 2: the forward pre-pass on a high-priority side stream UNDER the consistency kernel, which does not
        # depend on the forward warp, measured 1.74 ms at best and 1.80 ms on average against 1.766 ms in stream order: the
        # small pre-pass kernels starve behind the consistency kernel's 43 200 CTAs and the tile kernel then starts late.)
        self._forward(sl)
        if self.pack_masks:
            self._pack(slice(p0, p1), True)
        if self.consistency:
            if before_inverse is not None:
                before_inverse()
            self._inverse(sl)
            if self.pack_masks:
                self._pack(slice(p0, p1), False)
        if self.post:
            self._post(sl)

    def run(self) -> TrajectoryOutputs:
        """Enqueue forward warp (+ consistency masks) for all pairs on the current stream."""
        self._run_range(0, self.N)
        f, i = self.out_fwd, self.out_inv
        return TrajectoryOutputs(f.get("rgb_f32"), f.get("rgb_u8"), f["mask"], f.get("winner"),
                                 i.get("mask_reproj"), i.get("soft_mask_reproj"), f["status"])

    def run_post(self) -> dict:
        """The reference's per-frame host block between the warp and the SVD pipeline (diffusionGS.py:1572-1597,
        :1452-1491), for every pair, on the device (syn3r_b200.conditioning): from the forward outputs the dilated
        hole mask, the blanked conditioning frame and the latent-resolution mask; from the consistency outputs the
        latent mask of `mask_reproj` and the 8x8 pooled uncertainty.  Call after run() (a plan built with post=True
        does it inside every run); needs want_u8."""
        if not self.out_post:
            self._alloc_post()
        self._post()
        o = self.out_post
        res = {"fw": {"mask_png": o["post_mask_png"], "cond_u8": o["post_cond_u8"], "latent": o["post_latent"]}, "bw": {}}
        if self.consistency:
            res["bw"] = {"latent": o["post_latent_reproj"], "soft_pooled": o["post_soft_pooled"]}
        return res

    @staticmethod
    def slot_ranges(n_slots: int, chunks: int, ramp: bool = True):
        """Bounds of the `chunks` contiguous slot ranges of step_host.  ramp: the sizes double from range to range
        (1 : 2 : 4 : ...), so that the first device-to-host copy - PCIe is what bounds the step, and it idles until the
        first range's kernels are done - starts after a small range, while the per-range costs (the pre-pass, the
        launches) are still paid only `chunks` times; ramp=False: equal ranges."""
        n = max(1, min(int(chunks), n_slots))
        if not ramp:
            return [(n_slots * c) // n for c in range(n + 1)]
        w = np.cumsum([0] + [2 ** c for c in range(n)], dtype=np.float64)
        b = [int(round(n_slots * x / w[-1])) for x in w]
        for c in range(1, n + 1):                       # every range keeps at least one slot
            b[c] = min(max(b[c], b[c - 1] + 1), n_slots - (n - c))
        return b

    def step_host(self, rgbd_host: torch.Tensor, tgt_depth_host: Optional[torch.Tensor] = None, chunks: int = 5,
                  ramp: Optional[bool] = None):
        """One end-to-end step with HOST buffers: H2D of the inputs, kernels, D2H of the host outputs into the pinned
        staging buffers.  The pairs are processed in `chunks` contiguous ranges (slot_ranges()) so that the
        device-to-host copy of one range (on a side stream; PCIe is the bottleneck of this step) overlaps the kernels of
        the next.  ramp=None: doubling ranges when the step is bound by the copies (measured on config B: 6.80 ms against
        6.95 ms with equal ranges), equal ranges for a post=True plan, whose outputs are a third of the bytes and whose
        extra launches per range make the small leading ranges cost more than they gain (4.12 against 3.73 ms).
        Everything is joined back into the current stream before returning; nothing synchronises the host."""
        assert self.pinned is not None, "construct with pinned_host=True"
        cur = torch.cuda.current_stream()
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(device=self.dev)
            self._up_stream = torch.cuda.Stream(device=self.dev)
        side, up = self._copy_stream, self._up_stream
        # inputs travel on their own stream: the sources first (the forward kernels need nothing else), then the
        # target depth maps in the order the slot ranges read them.  Only the consistency kernel of a range waits for
        # its maps, so the upload of range c + 1 overlaps the kernels of range c instead of sitting in front of them.
        up.wait_stream(cur)
        bounds = self.slot_ranges(self.N, chunks, (not self.post) if ramp is None else ramp)
        n = len(bounds) - 1
        ev_src, ev_tgt = torch.cuda.Event(), []
        with torch.cuda.stream(up):
            self.rgbd.copy_(rgbd_host, non_blocking=True)
            ev_src.record(up)
            sent = np.zeros(self.T, bool)
            for c in range(n):
                if self.consistency:
                    need = np.unique(self.pair_tgt_h[bounds[c]:bounds[c + 1]])
                    need = need[~sent[need]]
                    sent[need] = True
                    # contiguous runs of target indices -> one copy each
                    for run in np.split(need, np.nonzero(np.diff(need) != 1)[0] + 1) if need.size else []:
                        a, b = int(run[0]), int(run[-1]) + 1
                        self.tgt_depth[a:b].copy_(tgt_depth_host[a:b], non_blocking=True)
                e = torch.cuda.Event()
                e.record(up)
                ev_tgt.append(e)
        cur.wait_event(ev_src)
        for c in range(n):
            p0, p1 = bounds[c], bounds[c + 1]
            self._run_range(p0, p1, before_inverse=lambda c=c: cur.wait_event(ev_tgt[c]))
            ev = torch.cuda.Event()
            ev.record(cur)
            side.wait_event(ev)
            with torch.cuda.stream(side):
                for k, v in self._host_items():
                    self.pinned["out_" + k][p0:p1].copy_(v[p0:p1], non_blocking=True)
        cur.wait_stream(side)
        cur.wait_stream(up)
        return self.pinned

    def launches_per_run(self) -> int:
        """Kernels of libsyn3r_warp.so per run() (memsets not counted): k_src_stats, k_fwd_prep (bounds), k_fwd_prep (pruned max z'),
        k_fwd_finalize, k_fwd_tile_fx, k_fwd_tile (redo list, normally empty) (+ k_reproj_masks2)."""
        return (6 + (1 if self.consistency else 0) + ((2 if self.consistency else 1) if self.post else 0) +
                ((2 if self.consistency else 1) if self.pack_masks else 0))


def from_synth(traj, pairs=None, **kw) -> TrajectoryWarp:
    """Plan + cameras for a `syn3r_b200.synth.Trajectory`; data is uploaded by the caller."""
    from . import synth
    S, T, H, W = traj.shape
    ps, pt = synth.all_pairs(S, T) if pairs is None else pairs
    tw = TrajectoryWarp(S, T, H, W, ps, pt, **kw)
    tw.set_cameras(traj.pose_src, traj.pose_tgt, traj.K)
    return tw
