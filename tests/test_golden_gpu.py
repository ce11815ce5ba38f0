"""The CUDA path against the committed golden vectors of the reference (tests/golden/): needs
no oracle library at run time.  Bar: <= 1e-6 px, <= 1e-4 m, observable status bits equal."""
import os

import numpy as np
import pytest

from golden_util import golden_files, load

pytestmark = pytest.mark.gpu


def _same(got, want, tol):
    f = np.isfinite(want)
    return np.array_equal(np.isfinite(got), f) and np.abs(got[f] - want[f]).max(initial=0.0) <= tol


@pytest.mark.parametrize("path", golden_files(), ids=os.path.basename)
def test_engine_matches_golden(path):
    import usgscsm_b200 as ub
    from usgscsm_b200 import _capi
    g = load(path)
    gm = ub.UsgsAstroPlugin(device=0).constructModelFromState(g["state"])
    sz = gm.getImageSize()
    assert [sz.line, sz.samp] == g["image_size"]
    if "ls_derived" in g:
        d = g["ls_derived"]
        np.testing.assert_allclose(gm.lsDerived()[:4], d[:4], rtol=1e-12)
        gm.lsSetProjective(d[4] != 0, d[5:19])
    for c in g["cases"]:
        x, y, z, a, st = gm.imageToGroundBatch(g["line"], g["samp"], g["height"], c["precision"])
        for got, key in ((x, "x"), (y, "y"), (z, "z"), (a, "achieved")):
            assert _same(got, c["i2g"][key], 1e-4), key
        assert np.array_equal(st & _capi.ST_REFERENCE_MASK, c["i2g"]["status"])
        l, s, a2, st2 = gm.groundToImageBatch(c["g2i_in"]["x"], c["g2i_in"]["y"], c["g2i_in"]["z"],
                                              c["precision"])
        for got, key in ((l, "line"), (s, "samp"), (a2, "achieved")):
            assert _same(got, c["g2i"][key], 1e-6), key
        assert np.array_equal(st2 & _capi.ST_REFERENCE_MASK, c["g2i"]["status"])
