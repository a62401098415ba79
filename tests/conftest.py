import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_golden(name):
    import numpy as np
    return dict(np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False))


@pytest.fixture(scope="session")
def golden():
    return load_golden


def expand_fusion(g):
    """fusion_*.npz in one shape.  The full-size fixture (fusion_3: the reference's statements unpatched at 576x1024) stores
    8-bit inputs, the `np.where` selection as bits and every 36th row of the float64 uncertainty; the small ones store
    everything.  Returns (cond_ori, gs, soft_ori, out_masks, out_cond_image, out_unc [N,h,W], rows) with `rows` the
    slice of image rows out_unc covers."""
    import numpy as np
    if "in_cond_u8" not in g:
        return (g["in_cond_ori"], g["in_gs"], g["in_soft_ori"], g["out_masks"], g["out_cond_image"],
                g["out_geo_inten_uncertainty"][..., 0], slice(None))
    f255 = np.float32(255)
    cond, gs = (g["in_cond_u8"] / f255).astype(np.float32), (g["in_gs_u8"] / f255).astype(np.float32)
    soft = (g["in_soft_u8"] / f255).astype(np.float32)
    N, H, W, _ = cond.shape
    sel = np.unpackbits(g["out_sel_bits"])[:N * H * W].reshape(N, H, W, 1).astype(bool)
    return (cond, gs, soft, g["out_masks"], np.clip(np.where(sel, gs, cond), 0, 1), g["out_geo_inten_uncertainty_rows9"][..., 0],
            slice(1, None, 36))
