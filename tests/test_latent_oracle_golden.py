"""CPU: the guided-latent-replacement oracle against the fixtures (scheduler statements re-executed with torch)."""
import glob
import os

import numpy as np
import pytest

from oracle import latent_oracle as lo  # checker only

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
LATENT = sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, "latent_*.npz")))


def test_latent_fixture_set_present():
    assert len(LATENT) == 5


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
