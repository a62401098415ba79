"""Execute the reference's OWN source text for code that cannot be imported (fixture generation only).

`model/diffusionGS.py` needs FSGS / open3d / trimesh / diffusers at import time, so its methods are not importable
in the build container (SURVEY §8c caveat 3).  Their source text, however, is on disk: this helper slices a line range
out of `/root/reference`, dedents it and `exec`s it in a namespace that supplies the free variables (`np`, `cv2`,
`PIL`, `torch`, `os`, the LIVE `forward_warp` / `inverse_warp`, a stub `self`, ...).  The fixtures built on it are
therefore outputs of the reference's own statements, not of a retyped copy; every fixture stores the line range and
the SHA-256 of the sliced text (`ref_source`), and any textual patch that was applied to run it at a reduced size
(`ref_patches`; an unpatched full-size fixture always accompanies the patched small ones).

Nothing from the reference is written into this repository: the text lives only in memory at generation time.
Used by make_golden_post.py, make_golden_fusion.py, make_golden_nearby.py, make_golden_latent.py,
make_golden_warp_images.py.  Not importable on the GPU box (no /root/reference there) and never needed there.
"""
import hashlib
import os
import sys
import textwrap
import types

import numpy as np
import torch

REF = os.environ.get("SYN3R_REFERENCE", "/root/reference")
COMMIT = "DecaYale/SYN3R@402f4e9"


def source(rel, lo, hi):
    """Lines lo..hi (1-based, inclusive) of REF/rel, dedented.  Returns (text, tag) where tag names the slice and its
    SHA-256 so that a fixture records exactly which text produced it."""
    with open(os.path.join(REF, rel)) as f:
        lines = f.readlines()
    text = textwrap.dedent("".join(lines[lo - 1:hi]))
    tag = f"{COMMIT} {rel}:{lo}-{hi} sha256={hashlib.sha256(text.encode()).hexdigest()}"
    return text, tag


def patched(text, patches):
    """Apply literal (old, new) replacements; every one must hit at least once.  Returns (text, description)."""
    desc = []
    for old, new in patches:
        n = text.count(old)
        assert n > 0, f"patch target not found in the reference text: {old!r}"
        text = text.replace(old, new)
        desc.append(f"{old!r} -> {new!r} (x{n})")
    return text, desc


def exec_function(rel, lo, hi, name, namespace, patches=()):
    """The `def name(...)` found in REF/rel:lo-hi as a plain function object (methods take `self` explicitly)."""
    text, tag = source(rel, lo, hi)
    text, desc = patched(text, patches)
    ns = dict(namespace)
    exec(compile(text, f"{REF}/{rel}:{lo}-{hi}", "exec"), ns)
    return ns[name], tag, desc


def exec_block(rel, lo, hi, namespace, patches=()):
    """Run statements REF/rel:lo-hi in `namespace` (a dict, updated in place).  Returns (tag, patch description)."""
    text, tag = source(rel, lo, hi)
    text, desc = patched(text, patches)
    exec(compile(text, f"{REF}/{rel}:{lo}-{hi}", "exec"), namespace)
    return tag, desc


class TorchOnCpu:
    """Stands in for the `torch` module in an exec namespace: `device='cuda'` becomes 'cpu' in factory calls (the
    reference hard-codes the device, diffusionGS.py:1438-1441, :1338-1343); everything else is the real torch."""

    def __getattr__(self, k):
        v = getattr(torch, k)
        if k in ("tensor", "zeros", "ones", "zeros_like", "ones_like", "arange", "as_tensor", "from_numpy"):
            def f(*a, **kw):
                if isinstance(kw.get("device"), str) and kw["device"].startswith("cuda"):
                    kw["device"] = "cpu"
                return v(*a, **kw)
            return f
        return v


def live_warp_modules():
    """The live `solver_utils.forward_warp` / `consistency` modules (with make_golden.py's Tensor.to shim)."""
    here = os.path.dirname(os.path.abspath(__file__))
    if here not in sys.path:
        sys.path.insert(0, here)
    import make_golden as mg
    return mg.ref_fw, mg.ref_cc


def live_scheduler():
    """The vendored, patched `EulerDiscreteScheduler` class itself.  The vendored diffusers imports
    `huggingface_hub.cached_download` (gone from the installed hub) and the scheduler file imports matplotlib (absent,
    used by a plotting helper only): both are stubbed in `sys.modules`, outside the reference tree."""
    import huggingface_hub
    if not hasattr(huggingface_hub, "cached_download"):
        def _no_download(*a, **k):
            raise RuntimeError("huggingface_hub.cached_download is a stub (fixture generation is offline)")
        huggingface_hub.cached_download = _no_download
    if "matplotlib" not in sys.modules:
        mpl, plt = types.ModuleType("matplotlib"), types.ModuleType("matplotlib.pyplot")
        mpl.pyplot = plt
        sys.modules["matplotlib"], sys.modules["matplotlib.pyplot"] = mpl, plt
    p = os.path.join(REF, "thirdparty", "diffusers", "src")
    if p not in sys.path:
        sys.path.insert(0, p)
    from diffusers.schedulers.scheduling_euler_discrete import EulerDiscreteScheduler
    return EulerDiscreteScheduler


def versions(*tags):
    import cv2
    import PIL
    return np.array([f"numpy {np.__version__}", f"torch {torch.__version__}", f"cv2 {cv2.__version__}",
                     f"PIL {PIL.__version__}"] + list(tags))
