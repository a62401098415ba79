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
