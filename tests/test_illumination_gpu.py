"""Parity of the batched getIlluminationDirection with the reference's per-point call
(LS:1878-1899 + getSunPosition LS:3165-3193, PF:1477-1498, SAR:1095-1102 + SAR:1974-2004,
FR:643-652): one sun position with and without a velocity, and an interpolated sun table."""
import json

import numpy as np
import pytest

from parity_util import models, pixels

pytestmark = pytest.mark.gpu


def _state(kind, sun):
    from usgscsm_b200 import synth
    base = {"linescan": lambda: synth.linescan_state(n_lines=2048, n_samples=512, seed=71),
            "pushframe": lambda: synth.pushframe_state(n_framelets=32, framelet_height=16, n_samples=512),
            "frame": lambda: synth.frame_state(n_lines=1024, n_samples=1024),
            "sar": lambda: synth.sar_state()}[kind]()
    head, body = base.split("\n", 1)       # "<model name>\n{json}"
    d = json.loads(body)
    rng = np.random.default_rng(3)
    if sun == "moving":
        d["m_sunPosition"] = [1.4e11, 3.0e10, -2.0e10]
        d["m_sunVelocity"] = [1.0e4, -2.5e4, 3.0e3]
    elif sun == "fixed":
        d["m_sunPosition"] = [1.4e11, 3.0e10, -2.0e10]
        d["m_sunVelocity"] = []
    elif sun == "table":
        nodes = 12
        t = np.arange(nodes)[:, None]
        pos = np.array([1.4e11, 3.0e10, -2.0e10])[None, :] + t * np.array([2.0e6, -5.0e6, 1.0e6])[None, :] \
            + rng.normal(size=(nodes, 3)) * 1e3
        d["m_sunPosition"] = [float(v) for v in pos.reshape(-1)]
        d["m_sunVelocity"] = [0.0] * (3 * nodes)
    return head + "\n" + json.dumps(d)


CASES = [(k, s) for k in ("linescan", "pushframe", "sar") for s in ("moving", "fixed", "table")] + \
        [("frame", "fixed"), ("frame", "moving")]


@pytest.mark.parametrize("kind,sun", CASES)
def test_illumination_direction(ref, kind, sun):
    import torch
    from usgscsm_b200 import _capi, synth
    gm, rm = models(ref, _state(kind, sun))
    if kind == "linescan":
        rm.ls_set_projective(True, gm.lsDerived()[5:19])
    nl, ns = rm.image_size()
    rng = np.random.default_rng(8)
    n = 400 if kind == "sar" else 4000
    line, samp = pixels(rng, n, nl, ns)
    x, y, z, _, _ = gm.imageToGroundBatch(line, samp, synth.synthetic_dem(line, samp))
    x[:5] = np.nan
    dx, dy, dz, st = gm.getIlluminationDirectionBatch(x, y, z)
    exp, est = rm.illumination_direction(x, y, z, threads=0)
    assert np.array_equal(st & _capi.ST_REFERENCE_MASK, est)
    got = np.stack([dx, dy, dz], axis=1)
    ok = np.isfinite(exp).all(axis=1)
    assert ok.sum() >= n - 5 - n // 20
    assert np.array_equal(np.isfinite(got).all(axis=1), ok)
    if kind == "frame":
        assert np.array_equal(got[ok], exp[ok])                 # ground - sun, bit for bit
    else:
        assert np.abs(got[ok] - exp[ok]).max() <= 1e-15          # unit vectors
        assert np.abs(np.linalg.norm(got[ok], axis=1) - 1.0).max() <= 1e-15
    d = gm.getIlluminationDirectionBatch(*(torch.as_tensor(a, device="cuda") for a in (x, y, z)))
    torch.cuda.synchronize()
    assert np.array_equal(d[0].cpu().numpy(), dx, equal_nan=True)
    i = 100
    assert gm.getIlluminationDirection((x[i], y[i], z[i])) == (dx[i], dy[i], dz[i])
