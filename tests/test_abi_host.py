"""CPU-only checks: the C-ABI library loads and exports every symbol include/*.h declares (no compute
calls without a GPU), argument validation answers with error codes, and the host-side small-matrix
logic equals the reference's NumPy / torch calls."""
import ctypes
import glob
import os
import re

import numpy as np
import pytest
import torch

from oracle import cpu_oracle as orc
from syn3r_b200 import _native as nat
from syn3r_b200 import synth
from syn3r_b200 import warp as w

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    names = set()
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        src = re.sub(r"/\*.*?\*/", "", open(h).read(), flags=re.S)
        names |= set(re.findall(r"\b(syn3r_[a-z0-9_]+)\s*\(", src))
    return names


def test_library_exports_every_declared_symbol():
    L = nat.lib()
    declared = _declared_symbols()
    assert declared == set(nat.SIGNATURES), declared ^ set(nat.SIGNATURES)
    for name in declared:
        assert hasattr(L, name), name
    assert L.syn3r_version() == nat.ABI_VERSION


def test_workspace_query_is_pure_host_arithmetic():
    L = nat.lib()
    hw = 576 * 1024
    leg = nat.FWD_LEGACY_GLOBAL
    assert L.syn3r_fwd_workspace_bytes(576, 1024, 3, 75, 75, leg) >= 75 * hw * 16
    assert L.syn3r_fwd_workspace_bytes(576, 1024, 3, 75, 75, leg | nat.FWD_WINNER) >= 75 * hw * 24
    # tile path: no per-pixel accumulators in global memory, only block statistics, cell boxes and the per-(pair, tile)
    # block lists (75 pairs x 96 tiles x 192 entries = 5.5 MB): 1 % of the 708 MB the global accumulators took
    assert 0 < L.syn3r_fwd_workspace_bytes(576, 1024, 3, 75, 75, nat.FWD_WINNER) < 10 << 20
    assert L.syn3r_fwd_workspace_bytes(0, 1024, 3, 75, 75, 0) == 0
    assert L.syn3r_inv_workspace_bytes(75) % 256 == 0


def test_argument_errors_return_codes_not_exceptions():
    L = nat.lib()
    rc = L.syn3r_forward_warp_batched(None, None, None, None, None, None, None, 1, 1, 1, 8, 8, 0, None, None, None,
                                      None, None, None, None, 0, None)
    assert rc == 1 and b"null" in L.syn3r_last_error()
    with pytest.raises(nat.Syn3rNativeError):
        nat.check(rc, "syn3r_forward_warp_batched")
    rc = L.syn3r_consistency_batched(None, None, None, None, None, None, None, None, 1, 4, 4, None, None)
    assert rc == 1
    outs = nat.InvOutputs()
    rc = L.syn3r_inverse_warp_batched(None, None, 1, None, None, None, None, None, None, None, 1, 4, 4, 20.0, 0.3, 0.1,
                                      ctypes.byref(outs), None, 0, None)
    assert rc == 1


def test_product_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "syn3r_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "cpu_oracle" not in src, f
                assert "libsyn3r_oracle" not in src, f


def test_forward_pair_params_equal_reference_host_calls():
    tr = synth.make_trajectory(3, 4, 8, 8, seed=1)
    ps, pt = synth.all_pairs(3, 4)
    Kinv, par = w.forward_pair_params(tr.pose_src, tr.pose_tgt, tr.K, None, ps, pt)
    assert Kinv.shape == (3, 9) and par.shape == (12, 24) and par.dtype == np.float64
    for p, (s, t) in enumerate(zip(ps, pt)):
        M, K1inv, K2 = orc.host_matrices(tr.pose_src[s], tr.pose_tgt[t], tr.K, None)
        assert np.array_equal(par[p, :12], M[:3].ravel())           # f32 product promoted exactly
        assert np.array_equal(par[p, 12:21], K2.ravel())
        assert np.array_equal(Kinv[s], K1inv.ravel())
    # f64 poses keep f64 precision (demo regime, model/SVD_1pass.py:886-915)
    _, par64 = w.forward_pair_params(tr.pose_src.astype(np.float64), tr.pose_tgt.astype(np.float64), tr.K, None,
                                     ps[:1], pt[:1])
    M64 = np.matmul(tr.pose_tgt[0].astype(np.float64), np.linalg.inv(tr.pose_src[0].astype(np.float64)))
    assert np.array_equal(par64[0, :12], M64[:3].ravel())


def test_inverse_pair_params_equal_torch_cpu():
    tr = synth.make_trajectory(2, 3, 8, 8, seed=2)
    ps, pt = synth.all_pairs(2, 3)
    PP = w.inverse_pair_params(tr.pose_src, tr.pose_tgt, ps, pt)
    for p, (s, t) in enumerate(zip(ps, pt)):
        p1, p2 = torch.from_numpy(tr.pose_src[s]), torch.from_numpy(tr.pose_tgt[t])
        assert np.array_equal(PP[p, 0], (p1 @ torch.inverse(p2)).numpy().ravel())
        assert np.array_equal(PP[p, 1], (p2 @ torch.inverse(p1)).numpy().ravel())


def test_pack_rgbd_and_pair_lists():
    tr = synth.make_trajectory(2, 25, 6, 10, seed=3)
    r = w.pack_rgbd(tr.rgb, tr.depth)
    assert r.shape == (2, 6, 10, 4) and r.dtype == np.float32
    assert np.array_equal(r[..., :3], tr.rgb) and np.array_equal(r[..., 3], tr.depth)
    ps, pt = synth.reference_policy_pairs(25, 2)
    assert list(ps[:12]) == [0] * 12 and list(ps[12:]) == [1] * 13 and list(pt) == list(range(25))


def test_product_fails_loudly_without_gpu():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import syn3r_b200.solver_utils.forward_warp as fw
    tr = synth.make_trajectory(1, 1, 8, 8, seed=4)
    with pytest.raises(nat.Syn3rNativeError):
        fw.forward_warp(tr.rgb[0], None, tr.depth[0], tr.pose_src[0], tr.pose_tgt[0], tr.K, None)


def test_step_host_slot_ranges_cover_every_slot_once():
    from syn3r_b200.trajectory import TrajectoryWarp
    for n_slots in (1, 2, 3, 4, 5, 10, 25, 75, 100, 1000):
        for chunks in (1, 2, 3, 4, 5, 8):
            for ramp in (True, False):
                b = TrajectoryWarp.slot_ranges(n_slots, chunks, ramp)
                assert b[0] == 0 and b[-1] == n_slots and len(b) == min(chunks, n_slots) + 1
                assert all(b[i] < b[i + 1] for i in range(len(b) - 1)), (n_slots, chunks, ramp, b)
    # the ramp starts small: the first device-to-host copy of config B's 75 pair warps waits for 5 of them, not 18
    assert TrajectoryWarp.slot_ranges(75, 4) == [0, 5, 15, 35, 75]
