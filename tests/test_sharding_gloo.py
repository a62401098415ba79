"""Multi-process host logic of the frame-partitioned path on CPU (gloo): partition, per-rank pair lists, and the
gather of per-rank frames to the diffusion rank.  The kernels themselves need a GPU (tests/test_gpu_parity.py);
here the per-rank "results" are deterministic CPU tensors so that the plumbing is checked end to end."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from syn3r_b200 import sharding, synth


def test_partition_units_matches_survey_split():
    assert [e - b for b, e in sharding.partition_units(100, 8)] == [13, 13, 13, 13, 12, 12, 12, 12]
    assert sharding.partition_units(25, 2) == [(0, 13), (13, 25)]
    # the diffusion rank (which ingests the peers' stores) gets the light share, the extras go far from it
    lens = [e - b for b, e in sharding.partition_units(100, 8, light=0)]
    assert sum(lens) == 100 and lens[0] == 12 and sorted(lens) == [12] * 4 + [13] * 4
    assert [e - b for b, e in sharding.partition_units(25, 8, light=0)][0] == 3
    assert sharding.partition_units(25, 2, light=0) == [(0, 12), (12, 25)]
    assert sharding.partition_units(3, 4) == [(0, 1), (1, 2), (2, 3), (3, 3)]
    for n in (0, 1, 7, 25, 64, 100):
        for w in (1, 2, 3, 4, 8):
            parts = sharding.partition_units(n, w)
            assert parts[0][0] == 0 and parts[-1][1] == n
            assert all(parts[i][1] == parts[i + 1][0] for i in range(w - 1))
            sizes = [e - b for b, e in parts]
            assert max(sizes) - min(sizes) <= 1


def test_pairs_for_frames_cover_the_single_gpu_pair_list():
    S, T = 3, 25
    ps_all, pt_all = synth.all_pairs(S, T)
    want = set(zip(ps_all.tolist(), pt_all.tolist()))
    got = set()
    for r in range(4):
        t0, t1 = sharding.partition_frames(T, 4, r)
        ps, pt = sharding.pairs_for_frames(S, t0, t1)
        got |= set(zip(ps.tolist(), (pt + t0).tolist()))
    assert got == want
    # reference policy: decided on the GLOBAL frame index (diffusionGS.py:1555-1567)
    ps_ref, pt_ref = synth.reference_policy_pairs(T, S)
    got = []
    for r in range(4):
        t0, t1 = sharding.partition_frames(T, 4, r)
        ps, pt = sharding.pairs_for_frames(S, t0, t1, "reference_policy")
        got += list(zip(ps.tolist(), (pt + t0).tolist()))
    assert got == list(zip(ps_ref.tolist(), pt_ref.tolist()))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _fake_frame(t, H, W):
    """Deterministic stand-in for the warped outputs of global frame t."""
    g = torch.Generator().manual_seed(1000 + t)
    return {"rgb_u8": torch.randint(0, 256, (H, W, 3), generator=g, dtype=torch.uint8),
            "mask": torch.randint(0, 2, (H, W), generator=g, dtype=torch.uint8),
            "soft_mask_reproj": torch.rand((H, W), generator=g, dtype=torch.float32)}


def _worker(rank, world, port, T, dst, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        H, W = 6, 8
        parts = sharding.partition_units(T, world)
        t0, t1 = parts[rank]
        frames = [_fake_frame(t, H, W) for t in range(t0, t1)]
        local = {k: (torch.stack([f[k] for f in frames]) if frames else
                     torch.empty((0,) + tuple(_fake_frame(0, H, W)[k].shape), dtype=_fake_frame(0, H, W)[k].dtype))
                 for k in ("rgb_u8", "mask", "soft_mask_reproj")}
        g = sharding.FrameGather(local, [e - b for b, e in parts], dst=dst)
        for _ in range(2):                       # reusable: two rounds give the same result
            out = g(local)
        ok = True
        if rank == dst:
            for t in range(T):
                ref = _fake_frame(t, H, W)
                ok = ok and all(torch.equal(out[k][t], ref[k]) for k in ref)
            ok = ok and g.bytes_received(local) == (T - (t1 - t0)) * (H * W * 3 + H * W + H * W * 4)
        else:
            ok = out is None
        dist.barrier()
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,T,dst", [(2, 5, 0), (2, 4, 1), (3, 2, 0)])
def test_frame_gather_gloo(world, T, dst):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, T, dst, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == {r: True for r in range(world)}


def test_compact_wire_selects_what_diffusion_consumes():
    P, H, W = 2, 4, 5
    fwd = {"rgb_f32": torch.zeros(P, H, W, 3), "rgb_u8": torch.zeros(P, H, W, 3, dtype=torch.uint8),
           "mask": torch.zeros(P, H, W, dtype=torch.uint8), "winner": torch.zeros(P, H, W, dtype=torch.int32),
           "status": torch.zeros(P, dtype=torch.int32)}
    inv = {"mask_reproj": torch.zeros(P, H, W, dtype=torch.uint8), "soft_mask_reproj": torch.zeros(P, H, W)}
    wire = sharding.compact_wire(fwd, inv)
    assert set(wire) == {"rgb_u8", "mask", "mask_reproj", "soft_mask_reproj"}
    assert sharding.wire_bytes(wire) == P * H * W * (3 + 1 + 1 + 4)
    assert np.isclose(sharding.wire_bytes(wire) / (P * H * W), 9.0)


def _worker_pipelined(rank, world, port, T, chunks, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        H, W = 6, 8
        parts = sharding.partition_units(T, world)
        t0, t1 = parts[rank]
        frames = [_fake_frame(t, H, W) for t in range(t0, t1)]
        proto = _fake_frame(0, H, W)
        local = {k: (torch.stack([f[k] for f in frames]) if frames else torch.empty((0,) + tuple(proto[k].shape), dtype=proto[k].dtype))
                 for k in proto}
        g = sharding.PipelinedGather(local, [e - b for b, e in parts], chunks, dst=0)
        seen = []
        for _ in range(2):
            g.begin()
            for c in range(g.n_chunks):
                a, b = g.bounds(c)
                seen.append((a, b))
                g.push(c, local)
            out = g.end()
        ok = seen[:g.n_chunks] == seen[g.n_chunks:] and seen[0][0] == 0 and seen[g.n_chunks - 1][1] == t1 - t0
        if rank == 0:
            for t in range(T):
                ref = _fake_frame(t, H, W)
                ok = ok and all(torch.equal(out[k][t], ref[k]) for k in ref)
        else:
            ok = ok and out is None
        dist.barrier()
        q.put((rank, bool(ok)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,T,chunks", [(2, 7, 3), (3, 4, 4)])
def test_pipelined_gather_gloo(world, T, chunks):
    """Per-chunk sends (transfer of one range behind the next range's kernels): every rank cuts its own rows
    into the same number of chunks, empty chunks included (3 ranks, 4 frames, 4 chunks)."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_pipelined, args=(r, world, port, T, chunks, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == {r: True for r in range(world)}


def test_workload_partition_cuts_trajectories_first_then_frames():
    """Config D (8 trajectories x 25 frames): 2 / 4 / 8 ranks get whole trajectories; 3 ranks cut inside one.  Config E:
    100 frames over 8 ranks -> 13,13,13,13,12,12,12,12."""
    from syn3r_b200 import workload
    wl = workload.for_config("D_dl3dv")
    assert wl.frames == 200 and wl.pairs == 400 and wl.mode == "pair"
    for world in (2, 4, 8):
        for a, b in sharding.partition_units(wl.frames, world):
            assert a % 25 == 0 and b % 25 == 0
    assert wl.frame_owner(0) == (0, 0) and wl.frame_owner(26) == (1, 1) and wl.frame_owner(199) == (7, 24)
    we = workload.for_config("E_stress")
    assert we.mode == "fused" and not we.consistency and we.frames == 100
    assert [b - a for a, b in sharding.partition_units(we.frames, 8)] == [13, 13, 13, 13, 12, 12, 12, 12]


def test_slot_partition_cuts_pairs_not_frames():
    """PAIR mode: 75 pair warps (25 frames x 3 sources) over 8 GPUs are 10,10,10,9,9,9,9,9 pair warps - the cut may fall
    inside a frame - where a frame partition would leave 12 on one rank; FUSED mode keeps whole frames."""
    from syn3r_b200 import workload
    wb = workload.for_config("B_svd")
    parts = workload.partition_slots(wb, 8)
    assert [q1 - q0 for q0, q1 in parts] == [10, 10, 10, 9, 9, 9, 9, 9]
    touched = [workload.frames_of_slots(wb, q0, q1) for q0, q1 in parts]
    assert touched[0] == (0, 4, 0, 2)            # slots 0..9 = frames 0..2 whole + 1 pair of frame 3
    assert touched[1] == (3, 7, 1, 1)            # slots 10..19: frame 3 minus its first pair ... frame 6 minus its last
    for (q0, q1), (f0, f1, h, t) in zip(parts, touched):
        assert 3 * (f1 - f0) - h - t == q1 - q0
    we = workload.for_config("E_stress")
    assert [q1 - q0 for q0, q1 in workload.partition_slots(we, 8)] == [13, 13, 13, 13, 12, 12, 12, 12]
    wd = workload.for_config("D_dl3dv")
    assert [q1 - q0 for q0, q1 in workload.partition_slots(wd, 8)] == [50] * 8
    assert [workload.frames_of_slots(wd, q0, q1) for q0, q1 in workload.partition_slots(wd, 8)][3] == (75, 100, 0, 0)
