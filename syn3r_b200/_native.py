"""ctypes binding of libsyn3r_warp.so (include/syn3r_warp.h).

There is NO fallback: if the CUDA library is missing, fails to load, or a call returns a non-zero
code, a `Syn3rNativeError` is raised.  The product never imports `oracle/`.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
# SYN3R_WARP_LIB: a differently configured build of the same library (A/B measurements only)
LIB_PATH = os.environ.get("SYN3R_WARP_LIB") or os.path.join(_HERE, "_lib", "libsyn3r_warp.so")
CSRC = os.path.join(_HERE, "csrc")

FWD_GATE_0P6 = 1 << 0
FWD_WINNER = 1 << 1
FWD_IS_IMAGE = 1 << 2
FWD_EXACT_WEIGHT = 1 << 3
FWD_LEGACY_GLOBAL = 1 << 4
FWD_FP32_TILE = 1 << 5
FWD_PHASE_ZMAX, FWD_PHASE_SPLAT, FWD_PHASE_RESOLVE = 1 << 8, 1 << 9, 1 << 10

ABI_VERSION = 17

c_void_p, c_int, c_size_t, c_float = ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t, ctypes.c_float


class Syn3rNativeError(RuntimeError):
    pass


class InvOutputs(ctypes.Structure):
    """struct syn3r_inv_outputs (host struct of device pointers)."""
    _fields_ = [(n, c_void_p) for n in (
        "warped_img", "warped_depth", "warped_masked_img", "soft_mask_reproj", "reproj_err",
        "mask_warp", "mask_depth", "mask", "mask_inv", "mask_depth_strict", "warped_bg", "mask_reproj")]


# name -> (restype, argtypes); every symbol include/syn3r_warp.h declares
SIGNATURES = {
    "syn3r_version": (c_int, []),
    "syn3r_last_error": (ctypes.c_char_p, []),
    "syn3r_selftest_ddiv2": (c_int, [ctypes.c_ulonglong, ctypes.c_ulonglong, c_int, c_void_p, c_void_p]),
    "syn3r_selftest_div_const": (c_int, [c_float, ctypes.c_ulonglong, ctypes.c_ulonglong, c_int, c_void_p,
                                         ctypes.POINTER(c_float), c_void_p]),
    "syn3r_device_info": (c_int, [ctypes.POINTER(c_int)] * 4),
    "syn3r_fwd_workspace_bytes": (c_size_t, [c_int] * 6),
    "syn3r_forward_warp_batched": (c_int, [c_void_p] * 7 + [c_int] * 6 + [c_void_p] * 7 + [c_size_t, c_void_p]),
    "syn3r_compute_transformed_points": (c_int, [c_void_p] * 3 + [c_int] * 2 + [c_void_p] * 3),
    "syn3r_bilinear_splat": (c_int, [c_void_p] * 5 + [c_int] * 4 + [c_void_p] * 5 + [c_size_t, c_void_p]),
    "syn3r_inv_workspace_bytes": (c_size_t, [c_int]),
    "syn3r_inverse_warp_batched": (c_int, [c_void_p] * 2 + [c_int] + [c_void_p] * 7 + [c_int] * 3 + [c_float] * 3 +
                                   [ctypes.POINTER(InvOutputs), c_void_p, c_size_t, c_void_p]),
    "syn3r_consistency_batched": (c_int, [c_void_p] * 8 + [c_int] * 3 + [c_void_p] * 2),
    "syn3r_points_from_depth": (c_int, [c_void_p] * 2 + [c_int] * 2 + [c_void_p] * 2),
    "syn3r_transform_points": (c_int, [c_void_p] * 2 + [c_int] * 2 + [c_void_p] * 2),
    "syn3r_pack_mask_bits": (c_int, [c_void_p, c_size_t, c_void_p, c_void_p]),
    "syn3r_condition_post_batched": (c_int, [c_void_p] * 4 + [c_int] * 3 + [c_void_p] * 7),
    "syn3r_fuse_uncertainty_batched": (c_int, [c_void_p] * 3 + [c_int] * 3 + [c_void_p] * 4),
    "syn3r_nearby_uncertainty_batched": (c_int, [c_void_p] * 4 + [c_int] * 3 + [c_float] + [c_void_p] * 3),
    "syn3r_flow_fb_consistency_batched": (c_int, [c_void_p] * 2 + [c_int] * 3 + [c_float] * 2 + [c_void_p] * 5),
    "syn3r_latent_guided_replace": (c_int, [c_void_p] * 4 + [c_int] * 10 + [c_void_p] * 3),
    "syn3r_latent_top_masks": (c_int, [c_void_p] * 4 + [c_int] * 10 + [c_float] + [c_void_p] * 4),
    "syn3r_mean_uncertainty_workspace_bytes": (c_size_t, [c_int]),
    "syn3r_mean_uncertainty_batched": (c_int, [c_void_p] + [c_int] * 3 + [c_void_p] * 2 + [c_size_t, c_void_p]),
    "syn3r_set_reproj_packed": (c_int, [c_int]),
    "syn3r_set_fx_overlap": (c_int, [c_int]),
    "syn3r_mean_uncertainty_static_max_p": (c_int, []),
    "syn3r_mean_uncertainty_static": (c_int, [c_void_p] + [c_int] * 3 + [c_void_p] * 2),
}

_lib = None


def build(force: bool = False, verbose: bool = False) -> str:
    """Compile the library in-tree with nvcc for sm_100a (Makefile in csrc/)."""
    cmd = ["make", "-C", CSRC, "-s"] + (["-B"] if force else [])
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise Syn3rNativeError("building libsyn3r_warp.so failed:\n" + r.stdout + r.stderr)
    if verbose:
        print(r.stdout + r.stderr)
    return LIB_PATH


def lib() -> ctypes.CDLL:
    """Load the CUDA library (once).  Raises if it is absent — there is no CPU path."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise Syn3rNativeError(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a).  syn3r_b200 has no CPU fallback.")
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError if a declared symbol is not exported
            fn.restype, fn.argtypes = res, args
        v = L.syn3r_version()
        if v != ABI_VERSION:
            raise Syn3rNativeError(f"libsyn3r_warp.so ABI {v} != expected {ABI_VERSION}; rebuild")
        _lib = L
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != 0:
        msg = lib().syn3r_last_error()
        raise Syn3rNativeError(f"{what or 'libsyn3r_warp'} failed (code {rc}): {msg.decode() if msg else ''}")


def ptr(t):
    """data pointer of a contiguous CUDA tensor (or None)."""
    if t is None:
        return None
    assert t.is_cuda and t.is_contiguous(), "native calls take contiguous CUDA tensors"
    return t.data_ptr()


def host_f32(a):
    """9- or 16-element float32 host array -> ctypes pointer (kept alive by the returned object)."""
    import numpy as np
    arr = np.ascontiguousarray(a, dtype=np.float32).ravel()
    return arr, arr.ctypes.data_as(c_void_p)


def device_info():
    v = [c_int() for _ in range(4)]
    check(lib().syn3r_device_info(*[ctypes.byref(x) for x in v]), "syn3r_device_info")
    return dict(sm_count=v[0].value, smem_optin=v[1].value, cc=(v[2].value, v[3].value))
