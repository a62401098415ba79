"""syn3r_mean_uncertainty_static: the allocation-free variant of syn3r_mean_uncertainty_batched (scratch in a file-scope
device array inside the library; include/syn3r_warp.h).  Called through the C ABI and compared with the existing entry
point on two calls in a row (the array keeps the first call's contents, so the second call - a different batch size
and different data - shows that nothing stale is read).  Collected last on purpose
(file name)."""
import ctypes

import pytest
import torch

from syn3r_b200 import _native as nat

pytestmark = pytest.mark.gpu


def _batched(L, soft, stream):
    P, H, W = soft.shape
    out = torch.full((P,), -1.0, dtype=torch.float32, device=soft.device)
    ws = torch.empty(int(L.syn3r_mean_uncertainty_workspace_bytes(P)), dtype=torch.uint8, device=soft.device)
    nat.check(L.syn3r_mean_uncertainty_batched(soft.data_ptr(), P, H, W, out.data_ptr(), ws.data_ptr(), ws.numel(), stream),
              "syn3r_mean_uncertainty_batched")
    return out


def _static(L, soft, out, stream):
    P, H, W = soft.shape
    nat.check(L.syn3r_mean_uncertainty_static(soft.data_ptr(), P, H, W, out.data_ptr(), stream), "syn3r_mean_uncertainty_static")


def test_static_scratch_variant_matches_the_workspace_variant_twice():
    L = nat.lib()
    dev = torch.device("cuda", 0)
    with torch.cuda.device(dev):
        stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
        g = torch.Generator(device="cpu").manual_seed(11)
        # inputs, outputs and the comparison values exist before the calls under test: those allocate nothing
        soft_a = torch.rand((37, 72, 128), generator=g).to(dev)
        soft_b = (torch.rand((5, 96, 100), generator=g) ** 3).to(dev)
        out_a = torch.full((37,), -1.0, dtype=torch.float32, device=dev)
        out_b = torch.full((5,), -1.0, dtype=torch.float32, device=dev)
        want_a, want_b = _batched(L, soft_a, stream), _batched(L, soft_b, stream)
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated(dev)
        _static(L, soft_a, out_a, stream)                   # first call
        _static(L, soft_b, out_b, stream)                   # second call: smaller batch over the first call's leftovers
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated(dev) == before
        # same kernels, same fixed-order reduction: the bar of tests/test_gpu_scoring.py is bit equality between runs
        assert torch.equal(out_a, want_a)
        assert torch.equal(out_b, want_b)
        ref = (1.0 - soft_b.double()).mean(dim=(1, 2))
        assert float((out_b.double() - ref).abs().max()) <= 2e-6
        # and once more with the first batch: results do not depend on what the array held
        out_a.fill_(-1.0)
        _static(L, soft_a, out_a, stream)
        torch.cuda.synchronize()
        assert torch.equal(out_a, want_a)


def test_static_scratch_variant_rejects_batches_beyond_its_maximum():
    L = nat.lib()
    max_p = int(L.syn3r_mean_uncertainty_static_max_p())
    assert max_p == 4096
    dev = torch.device("cuda", 0)
    with torch.cuda.device(dev):
        soft = torch.zeros((1, 8, 8), dtype=torch.float32, device=dev)
        out = torch.zeros((1,), dtype=torch.float32, device=dev)
        rc = L.syn3r_mean_uncertainty_static(soft.data_ptr(), max_p + 1, 8, 8, out.data_ptr(), None)
        assert rc == 2                                      # SYN3R_EWORKSPACE, nothing launched
        msg = L.syn3r_last_error().decode()
        assert "exceeds the static scratch" in msg and str(max_p) in msg
        assert L.syn3r_mean_uncertainty_static(soft.data_ptr(), 0, 8, 8, out.data_ptr(), None) == 0   # empty batch: no-op
        with pytest.raises(nat.Syn3rNativeError):
            nat.check(L.syn3r_mean_uncertainty_static(None, 1, 8, 8, out.data_ptr(), None), "null input")
