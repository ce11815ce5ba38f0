"""The committed model-level golden vectors (tests/golden/model_*.json) are what oracle/_ref
produces today: guards against the fixtures and the oracle drifting apart.  CPU only."""
import os

import numpy as np
import pytest

from golden_util import golden_files, load


def test_fixtures_present():
    names = [os.path.basename(p) for p in golden_files()]
    for want in ("model_linescan_nac.json", "model_pushframe_wac.json", "model_frame_radial.json",
                 "model_sar_orbital.json"):
        assert want in names


@pytest.mark.parametrize("path", golden_files(), ids=os.path.basename)
def test_oracle_reproduces_golden(ref, path):
    g = load(path)
    rm = ref.RefModel(g["state"])
    assert list(rm.image_size()) == g["image_size"]
    for c in g["cases"]:
        x, y, z, a, st = rm.image_to_ground(g["line"], g["samp"], g["height"], c["precision"])
        for got, key in ((x, "x"), (y, "y"), (z, "z"), (a, "achieved")):
            assert np.array_equal(got, c["i2g"][key], equal_nan=True)
        assert np.array_equal(st, c["i2g"]["status"])
        l, s, a2, st2 = rm.ground_to_image(c["g2i_in"]["x"], c["g2i_in"]["y"], c["g2i_in"]["z"], c["precision"])
        for got, key in ((l, "line"), (s, "samp"), (a2, "achieved")):
            assert np.array_equal(got, c["g2i"][key], equal_nan=True)
        assert np.array_equal(st2, c["g2i"]["status"])
