"""Parity of the batched getImageTime / getSensorPosition / getSensorVelocity with the reference's
per-point calls (LS:1543-1653, PF:1258-1320, FR:665-765, SAR:1113-1238), all four models, with
and without parameter adjustments."""
import numpy as np
import pytest

from parity_util import models, pixels

pytestmark = pytest.mark.gpu


def _cases():
    from usgscsm_b200 import synth
    return {
        "linescan": synth.linescan_state(n_lines=2048, n_samples=512, seed=61),
        "linescan_adjusted": synth.linescan_state(n_lines=1024, n_samples=512, seed=62,
                                                  params=(3.0, -2.0, 1.0, 0.01, -0.01, 0.02, 0, 0.3, 0, 0, 0, 0, 0, 0, 0, 0.05)),
        "linescan_segments": synth.linescan_state(n_lines=1024, n_samples=256, seed=63, n_time_segments=3),
        "pushframe": synth.pushframe_state(n_framelets=32, framelet_height=16, n_samples=512),
        "frame": synth.frame_state(n_lines=1024, n_samples=1024),
        "sar": synth.sar_state(),
    }


@pytest.mark.parametrize("kind", ["linescan", "linescan_adjusted", "linescan_segments", "pushframe", "frame", "sar"])
def test_image_sensor_state(ref, kind):
    import torch
    from usgscsm_b200 import _capi
    state = _cases()[kind]
    gm, rm = models(ref, state)
    nl, ns = rm.image_size()
    rng = np.random.default_rng(9)
    n = 20000
    line, samp = pixels(rng, n, nl, ns)
    line[:40] -= 25.0          # outside the image: extrapolated by the reference, so here too
    samp[40:80] += ns          # frame model: getSensorVelocity throws BOUNDS
    line[80:90] = np.nan
    t, p, v, st = gm.imageSensorStateBatch(line, samp)
    exp, est = rm.image_sensor_state(line, samp, threads=0)
    assert np.array_equal(st & _capi.ST_REFERENCE_MASK, est)
    if kind == "frame":
        assert (est[40:80] & _capi.ST_ERROR_OTHER).all() and est[100] == 0
    got = np.stack([t, *p, *v], axis=1)
    ok = (est & (_capi.ST_ERROR_OTHER | _capi.ST_ERROR_ALGORITHM | _capi.ST_ERROR_UNKNOWN)) == 0
    # the reference stops at the throwing call; before it the values must still agree
    assert np.array_equal(got[ok], exp[ok], equal_nan=True)          # bit for bit
    assert np.array_equal(got[~ok][:, :4], exp[~ok][:, :4], equal_nan=True)
    assert np.isnan(got[~ok][:, 4:]).all()
    # device arrays
    d = gm.imageSensorStateBatch(torch.as_tensor(line, device="cuda"), torch.as_tensor(samp, device="cuda"))
    torch.cuda.synchronize()
    assert np.array_equal(d[0].cpu().numpy(), t, equal_nan=True)
    assert np.array_equal(d[2][1].cpu().numpy(), v[1], equal_nan=True)
    # scalar mirrors
    i = 200
    assert gm.getImageTime((line[i], samp[i])) == exp[i, 0]
    assert gm.getSensorPosition((line[i], samp[i])) == tuple(exp[i, 1:4])
    assert gm.getSensorVelocity((line[i], samp[i])) == tuple(exp[i, 4:7])
