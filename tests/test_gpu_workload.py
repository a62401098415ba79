"""GPU: a rank's share of a multi-trajectory workload (syn3r_b200/workload.py: ONE plan over the sources of the
trajectories it touches, block-diagonal pair list) equals the trajectories warped one by one; FUSED plans and their
slot ranges equal the whole-batch FUSED call."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from syn3r_b200 import synth, trajectory, warp as w, workload  # noqa: E402


def test_block_diagonal_plan_equals_per_trajectory_plans():
    wl = workload.Workload("mini_D", "pair", True, 48, 64, [(2, 3)] * 3)
    f0, f1 = 2, 8                                     # cuts inside trajectory 0 and trajectory 2
    sh = workload.build_plan(wl, f0, f1, seed=5, want_f32=True, want_u8=True, want_winner=True)
    out = sh.plan.run()
    torch.cuda.synchronize()
    assert sh.traj_ids == [0, 1, 2] and sh.plan.S == 6 and sh.plan.T == 6 and sh.plan.P == 12
    for p, (k, s, t) in enumerate(sh.pair_global):
        tr = synth.make_trajectory(2, 3, 48, 64, seed=5 + k)
        one = trajectory.from_synth(tr, pairs=(np.array([s], np.int32), np.array([t], np.int32)), want_u8=True, want_winner=True)
        one.upload(torch.from_numpy(w.pack_rgbd(tr.rgb, tr.depth)), torch.from_numpy(tr.depth_tgt))
        r = one.run()
        torch.cuda.synchronize()
        assert torch.equal(out.mask[p], r.mask[0]) and torch.equal(out.rgb_u8[p], r.rgb_u8[0]), p
        assert torch.equal(out.rgb_f32[p], r.rgb_f32[0]) and torch.equal(out.mask_reproj[p], r.mask_reproj[0]), p
        assert torch.equal(out.soft_mask_reproj[p], r.soft_mask_reproj[0]), p
        # the winner index carries the plan's LOCAL source number
        s_loc = int(sh.plan.pair_src_h[p])
        hw = 48 * 64
        assert torch.equal(torch.where(out.winner[p] >= 0, out.winner[p] - s_loc * hw, out.winner[p]),
                           torch.where(r.winner[0] >= 0, r.winner[0] - s * hw, r.winner[0])), p


def test_fused_plan_and_slot_ranges():
    wl = workload.Workload("mini_E", "fused", False, 64, 96, [(5, 6)], pose_step=0.3, lazy_sources=True)
    sh = workload.build_plan(wl, 1, 6, seed=2, want_f32=True, want_u8=True, want_winner=True)
    tw = sh.plan
    assert tw.fused and tw.N == 5 and tw.P == 25
    whole = {k: v.clone() for k, v in tw.run().__dict__.items() if isinstance(v, torch.Tensor)}
    torch.cuda.synchronize()
    # in slot ranges (what step_host / the pipelined gather do)
    for v in tw.out_fwd.values():
        v.zero_()
    for a, b in ((0, 2), (2, 3), (3, 5)):
        tw._run_range(a, b)
    torch.cuda.synchronize()
    for k in ("rgb_f32", "rgb_u8", "mask", "winner"):
        assert torch.equal(tw.out_fwd[k], whole[k]), k
    # frame 0 of the share against the PAIR-mode warps of its sources: fused mask == OR of the pair masks
    p0, p1 = (int(v) for v in np.searchsorted(tw.pair_tgt_h, [0, 1]))
    o = w.forward_warp_batched(tw.rgbd, tw.Kinv_src, tw.pair_par[p0:p1].contiguous(), tw.pair_src[p0:p1].contiguous(),
                               want_f32=False, want_u8=False)
    assert torch.equal(o["mask"].bool().any(0), whole["mask"][0].bool())
    assert torch.equal(whole["winner"][0] >= 0, whole["mask"][0].bool())


@pytest.mark.parametrize("world", [2, 3, 4])
def test_slot_partition_inside_a_frame_reassembles_the_workload(world):
    """PAIR mode: the ranks' shares cut at PAIR granularity (workload.partition_slots: the cut may fall inside a target
    frame) and laid out at their slot offsets equal the one-GPU plan, bit for bit."""
    wl = workload.Workload("mini_B", "pair", True, 48, 64, [(3, 5)])                 # 15 pair warps
    full = workload.build_plan(wl, 0, wl.frames, seed=3, want_f32=False, want_u8=True, want_winner=False).plan
    ref = full.run()
    torch.cuda.synchronize()
    parts = workload.partition_slots(wl, world)
    assert sum(q1 - q0 for q0, q1 in parts) == 15 and max(q1 - q0 for q0, q1 in parts) - min(q1 - q0 for q0, q1 in parts) <= 1
    for q0, q1 in parts:
        f0, f1, head, tail = workload.frames_of_slots(wl, q0, q1)
        sh = workload.build_plan(wl, f0, f1, seed=3, trim=(head, tail), want_f32=False, want_u8=True, want_winner=False)
        assert sh.plan.P == q1 - q0 and len(sh.frame_of_slot) == q1 - q0
        o = sh.plan.run()
        torch.cuda.synchronize()
        assert torch.equal(o.mask, ref.mask[q0:q1]) and torch.equal(o.rgb_u8, ref.rgb_u8[q0:q1])
        assert torch.equal(o.mask_reproj, ref.mask_reproj[q0:q1]) and torch.equal(o.soft_mask_reproj, ref.soft_mask_reproj[q0:q1])
