"""CPU: the guided-latent-replacement oracle against the fixtures (outputs of the reference's live patched scheduler)."""
import glob
import os

import numpy as np
import pytest

from oracle import latent_oracle as lo  # checker only

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
LATENT = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "latent_[0-9].npz")))


def test_latent_fixture_set_present():
    assert len(LATENT) == 5      # latent_[0-4].npz


@pytest.mark.parametrize("name", LATENT)
def test_guided_replace_oracle_vs_golden(golden, name):
    g = golden(name)
    o = lo.guided_replace(g["pred"], g["temp_cond_latents"], g["mask"], g["lambda_ts"], int(g["step_index"]))
    assert o["out"].dtype == np.float32
    assert np.array_equal(o["out"], g["out"], equal_nan=True)                      # bit-exact: selection + exact float32 ops
    assert np.array_equal(o["cutoff_index"], g["aux_cutoff_index"])
    assert np.array_equal(o["cutoff_value"], g["aux_cutoff_value"], equal_nan=True)
    # first and last frame are the conditioning latents (:1496-1499)
    assert np.array_equal(o["out"][:, 0], np.broadcast_to(g["temp_cond_latents"][1, 0], o["out"][:, 0].shape), equal_nan=True)


TOP = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "latent_top_*.npz")))


def test_latent_top_fixture_set_present():
    assert len(TOP) == 3


@pytest.mark.parametrize("name", TOP)
def test_top_masks_oracle_vs_golden(golden, name):
    """The selection-only form (step_single :548-589) with the caller's weight_clamp."""
    g = golden(name)
    o = lo.top_masks(g["pred"], g["temp_cond_latents"], g["mask"], g["lambda_ts"], int(g["step_index"]), float(g["weight_clamp"]))
    assert o["top_masks"].dtype == np.bool_ and np.array_equal(o["top_masks"], g["out_top_masks"])
    assert np.array_equal(o["cutoff_index"], g["aux_cutoff_index"])
    assert np.array_equal(o["cutoff_value"], g["aux_cutoff_value"], equal_nan=True)


INTERP = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "latent_interp_*.npz")))


def test_latent_interp_fixture_set_present():
    assert len(INTERP) == 3


@pytest.mark.parametrize("name", INTERP)
def test_step_interp_top_masks_oracle_vs_golden(golden, name):
    """`step_interp` (:735-778): frames 1..23, clamp 0.4, a frame of ones on both ends."""
    g = golden(name)
    o = lo.top_masks(g["pred"], g["temp_cond_latents"], g["mask"], g["lambda_ts"], int(g["step_index"]), float(g["weight_clamp"]),
                     trailing_ones=True)
    assert o["top_masks"].shape == g["out_top_masks"].shape and np.array_equal(o["top_masks"], g["out_top_masks"])
    assert np.array_equal(o["cutoff_index"], g["aux_cutoff_index"])
    assert np.array_equal(o["cutoff_value"], g["aux_cutoff_value"], equal_nan=True)


@pytest.mark.parametrize("name", LATENT + TOP + INTERP)
def test_latent_fixtures_come_from_the_live_scheduler(golden, name):
    v = " ".join(str(x) for x in golden(name)["versions"])
    assert "live EulerDiscreteScheduler" in v and "sha256=" in v
