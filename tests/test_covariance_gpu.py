"""Parity of the batched error-propagation overloads -- groundToImage(csm::EcefCoordCovar) and
imageToGround(csm::ImageCoordCovar, height, heightVariance) -- with the reference's per-point
calls (LS:910-964 / 1075-1157, PF:784-840 / 909-990, SAR:709-758 / 864-940), with a non-trivial
parameter covariance set through setParameterCovariance on both sides."""
import numpy as np
import pytest

from parity_util import M_TOL, PX_TOL, models, pixels

pytestmark = pytest.mark.gpu


def _cases():
    from usgscsm_b200 import synth
    return {
        "linescan": synth.linescan_state(n_lines=2048, n_samples=512, seed=51),
        "linescan_adjusted": synth.linescan_state(n_lines=1024, n_samples=512, seed=52,
                                                  params=(3.0, 0, 0, 0, -0.01, 0, 0, 0.3, 0, 0, 0, 0, 0, 0, 0, 0.05)),
        "pushframe": synth.pushframe_state(n_framelets=32, framelet_height=16, n_samples=512),
        "sar": synth.sar_state(),
    }


def _pair(ref, kind, rng):
    gm, rm = models(ref, _cases()[kind])
    if kind.startswith("linescan"):
        rm.ls_set_projective(True, gm.lsDerived()[5:19])
    # a symmetric positive semi-definite parameter covariance with off-diagonal terms
    npar = gm.getNumParameters()
    a = rng.normal(size=(npar, npar)) * 1e-3
    cov = a @ a.T
    for i in range(npar):
        for j in range(npar):
            gm.setParameterCovariance(i, j, cov[i, j])
            rm.set_parameter_covariance(i, j, cov[i, j])
    assert gm.getParameterCovariance(1, 2) == cov[1, 2]
    return gm, rm


def _spd(rng, n, k, scale):
    a = rng.normal(size=(n, k, k)) * scale
    return (a @ a.transpose(0, 2, 1)).reshape(n, k * k)


def _close(got, exp, ok):
    scale = np.maximum(1.0, np.abs(exp[ok]).max(axis=0))
    return (np.abs(got[ok] - exp[ok]) <= 1e-9 * scale).all()


@pytest.mark.parametrize("kind", ["linescan", "linescan_adjusted", "pushframe", "sar"])
def test_ground_to_image_covar(ref, kind):
    import torch
    from usgscsm_b200 import _capi, synth
    rng = np.random.default_rng(5)
    gm, rm = _pair(ref, kind, rng)
    nl, ns = rm.image_size()
    n = 300 if kind == "sar" else 1500
    line, samp = pixels(rng, n, nl, ns)
    x, y, z, _, _ = gm.imageToGroundBatch(line, samp, synth.synthetic_dem(line, samp))
    x[:5] = np.nan
    gcov = _spd(rng, n, 3, 3.0)
    gl, gs, gc, ga, gst = gm.groundToImageCovarBatch(x, y, z, gcov, 1e-6)
    el, es, ec, ea, est = rm.ground_to_image_covar(x, y, z, gcov, 1e-6, threads=0)
    assert np.array_equal(gst & _capi.ST_REFERENCE_MASK, est)
    ok = np.isfinite(ec).all(axis=1)
    assert ok.sum() >= n * 0.9 and np.array_equal(np.isfinite(gc).all(axis=1), ok)
    assert np.abs(gl[ok] - el[ok]).max() <= PX_TOL and np.abs(gs[ok] - es[ok]).max() <= PX_TOL
    assert _close(gc, ec, ok)
    assert (gc[ok][:, 0] > 0).all() and (gc[ok][:, 3] > 0).all()      # variances
    fa = np.isfinite(ea)
    assert np.array_equal(np.isfinite(ga), fa)
    assert np.abs(ga[fa] - ea[fa]).max(initial=0.0) <= PX_TOL
    d = gm.groundToImageCovarBatch(*(torch.as_tensor(a, device="cuda") for a in (x, y, z, gcov)), 1e-6)
    torch.cuda.synchronize()
    assert np.array_equal(d[2].cpu().numpy(), gc, equal_nan=True)
    assert np.array_equal(d[0].cpu().numpy(), gl, equal_nan=True)


@pytest.mark.parametrize("kind", ["linescan", "linescan_adjusted", "pushframe", "sar"])
def test_image_to_ground_covar(ref, kind):
    import torch
    from usgscsm_b200 import _capi, synth
    rng = np.random.default_rng(6)
    gm, rm = _pair(ref, kind, rng)
    nl, ns = rm.image_size()
    n = 300 if kind == "sar" else 1500
    line, samp = pixels(rng, n, nl, ns)
    h = synth.synthetic_dem(line, samp)
    hv = rng.uniform(1.0, 100.0, n)
    icov = _spd(rng, n, 2, 0.5)
    gx, gy, gz, gc, ga, gst = gm.imageToGroundCovarBatch(line, samp, icov, h, hv, 1e-6)
    ex, ey, ez, ec, ea, est = rm.image_to_ground_covar(line, samp, icov, h, hv, 1e-6, threads=0)
    assert np.array_equal(gst & _capi.ST_REFERENCE_MASK, est)
    ok = np.isfinite(ec).all(axis=1)
    assert ok.sum() >= n * 0.9 and np.array_equal(np.isfinite(gc).all(axis=1), ok)
    for g, e in ((gx, ex), (gy, ey), (gz, ez)):
        assert np.abs(g[ok] - e[ok]).max() <= M_TOL
    assert _close(gc, ec, ok)
    fa = np.isfinite(ea)
    assert np.array_equal(np.isfinite(ga), fa)
    assert np.abs(ga[fa] - ea[fa]).max(initial=0.0) <= M_TOL
    # one height / variance for all points; device arrays give the same bits
    s = gm.imageToGroundCovarBatch(line, samp, icov, 25.0, 4.0, 1e-6)
    e = rm.image_to_ground_covar(line, samp, icov, np.full(n, 25.0), np.full(n, 4.0), 1e-6, threads=0)
    ok2 = np.isfinite(e[3]).all(axis=1)
    assert _close(s[3], e[3], ok2)
    d = gm.imageToGroundCovarBatch(*(torch.as_tensor(a, device="cuda") for a in (line, samp, icov, h, hv)), 1e-6)
    torch.cuda.synchronize()
    assert np.array_equal(d[3].cpu().numpy(), gc, equal_nan=True)
    assert np.array_equal(d[0].cpu().numpy(), gx, equal_nan=True)


def test_frame_model_is_refused():
    """The reference's frame overloads are stubs (FR:234-278, 374-395)."""
    import usgscsm_b200 as ub
    from usgscsm_b200 import _capi, synth
    gm = ub.UsgsAstroPlugin().constructModelFromState(synth.frame_state())
    with pytest.raises(_capi.EngineError) as e:
        gm.groundToImageCovarBatch(np.zeros(2), np.zeros(2), np.zeros(2), np.zeros((2, 9)))
    assert e.value.code == _capi.EUNSUPPORTED


@pytest.mark.parametrize("kind", ["linescan", "pushframe"])
def test_default_covariance_from_update_state(ref, kind):
    """a state with m_gsd == 1.0 && m_flyingHeight == 1000.0 (every ISD-fresh state,
    every synth state) makes updateState REPLACE m_covariance by a gsd-derived diagonal
    (LS:697-735, PF:660-690).  No setParameterCovariance here: the error-propagation outputs,
    getParameterCovariance and getModelState's m_covariance must match the reference's."""
    import json
    from usgscsm_b200 import synth
    state = _cases()[kind]
    gm, rm = models(ref, state)
    if kind == "linescan":
        rm.ls_set_projective(True, gm.lsDerived()[5:19])
    want = np.array(json.loads(rm.state().split("\n", 1)[1])["m_covariance"])
    got = np.array(json.loads(gm.getModelState().split("\n", 1)[1])["m_covariance"])
    assert want.shape == got.shape == (256,)
    assert np.count_nonzero(want) == 16 and want[0] != 1.0        # not the state's identity
    np.testing.assert_allclose(got, want, rtol=1e-12, atol=0.0)
    assert gm.getParameterCovariance(3, 3) == pytest.approx(want[3 * 16 + 3], rel=1e-12)
    rng = np.random.default_rng(8)
    nl, ns = rm.image_size()
    line, samp = pixels(rng, 800, nl, ns)
    x, y, z, _, _ = gm.imageToGroundBatch(line, samp, synth.synthetic_dem(line, samp))
    gcov = _spd(rng, 800, 3, 2.0)
    l, s, ic, a, st = gm.groundToImageCovarBatch(x, y, z, gcov)
    rl, rs, ric, ra, rst = rm.ground_to_image_covar(x, y, z, gcov)
    ok = np.isfinite(rl)
    assert ok.all() and np.abs(l - rl).max() <= PX_TOL and np.abs(s - rs).max() <= PX_TOL
    assert _close(ic.reshape(-1, 4), ric, ok)
    # the sensor term is not negligible: it must show in the image covariance
    gm.setParameterCovariance(0, 0, 0.0)
    l2, s2, ic2, _, _ = gm.groundToImageCovarBatch(x, y, z, gcov)
    assert np.abs(ic2.reshape(-1, 4) - ic.reshape(-1, 4)).max() > 1e-6
