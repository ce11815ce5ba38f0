"""Parity of the batched imageToRemoteImagingLocus, computeGroundPartials and setEllipsoid with
the reference's per-point calls (LS:1254-1325, PF:1060-1116, FR:488-543 / 955-1003,
SAR:1009-1030 / 1384-1406; setEllipsoid LS:2116-2119), all four sensor models."""
import numpy as np
import pytest

from parity_util import M_TOL, PX_TOL, check_g2i, check_i2g, models, pixels

pytestmark = pytest.mark.gpu


def _cases():
    from usgscsm_b200 import synth
    return {
        "linescan": synth.linescan_state(n_lines=2048, n_samples=512, seed=31),
        "linescan_adjusted": synth.linescan_state(n_lines=1024, n_samples=512, seed=32,
                                                  params=(3.0, 0, 0, 0, -0.01, 0, 0, 0.3, 0, 0, 0, 0, 0, 0, 0, 0.05)),
        "pushframe": synth.pushframe_state(n_framelets=32, framelet_height=16, n_samples=512),
        "frame": synth.frame_state(n_lines=1024, n_samples=1024),
        "frame_transverse": synth.frame_state(distortion_type=1, dist_coeffs=(
            0.0, 1.0, 0.0, 1e-6, 2e-6, -1e-6, 1e-8, 0.0, 2e-8, 0.0,
            0.0, 0.0, 1.0, -1e-6, 1e-6, 2e-6, 0.0, 1e-8, 0.0, -2e-8)),
        "sar": synth.sar_state(),
    }


def _pair(ref, kind):
    state = _cases()[kind]
    gm, rm = models(ref, state)
    if kind.startswith("linescan"):
        rm.ls_set_projective(True, gm.lsDerived()[5:19])
    return gm, rm


KINDS = ["linescan", "linescan_adjusted", "pushframe", "frame", "frame_transverse", "sar"]


@pytest.mark.parametrize("kind", KINDS)
def test_remote_imaging_locus(ref, kind):
    gm, rm = _pair(ref, kind)
    nl, ns = rm.image_size()
    rng = np.random.default_rng(12)
    line, samp = pixels(rng, 20000, nl, ns)
    # a few pixels outside the image: the reference extrapolates, so must the kernel
    line[:50] -= 30.0
    samp[50:100] += ns
    got = gm.imageToRemoteImagingLocusBatch(line, samp, 1e-5)
    exp = rm.remote_locus(line, samp, 1e-5)
    from usgscsm_b200 import _capi
    assert np.array_equal(got[7] & _capi.ST_REFERENCE_MASK, exp[7])
    ok = (exp[7] & _capi.ST_NONFINITE) == 0
    for k in range(3):   # locus point, metres
        assert np.abs(got[k][ok] - exp[k][ok]).max() <= M_TOL
    for k in range(3, 6):  # direction (unit-scale)
        assert np.abs(got[k][ok] - exp[k][ok]).max() <= 1e-12
    fa = np.isfinite(exp[6])
    assert np.array_equal(np.isfinite(got[6]), fa)
    assert np.abs(got[6][fa] - exp[6][fa]).max(initial=0.0) <= M_TOL
    # device-pointer flavour gives the same bits as the host flavour
    import torch
    tl, ts = torch.as_tensor(line, device="cuda"), torch.as_tensor(samp, device="cuda")
    dev = gm.imageToRemoteImagingLocusBatch(tl, ts, 1e-5)
    torch.cuda.synchronize()
    for k in range(7):
        assert np.array_equal(dev[k].cpu().numpy(), got[k], equal_nan=True)


@pytest.mark.parametrize("kind", KINDS)
def test_proximate_imaging_locus(ref, kind):
    from usgscsm_b200 import _capi, synth
    gm, rm = _pair(ref, kind)
    nl, ns = rm.image_size()
    rng = np.random.default_rng(15)
    n = 12000
    line, samp = pixels(rng, n, nl, ns)
    x, y, z, _, _ = gm.imageToGroundBatch(line, samp, synth.synthetic_dem(line, samp))
    # ground points off the pixel's ray (the usual case for a caller), some far off, some NaN
    x += rng.normal(0, 30.0, n)
    y += rng.normal(0, 30.0, n)
    z += rng.normal(0, 30.0, n)
    x[:30] *= 1.5
    z[30:40] = np.nan
    got = gm.imageToProximateImagingLocusBatch(line, samp, x, y, z, 1e-5)
    exp = rm.proximate_locus(line, samp, x, y, z, 1e-5, threads=0)
    assert np.array_equal(got[7] & _capi.ST_REFERENCE_MASK, exp[7])
    ok = (exp[7] & _capi.ST_NONFINITE) == 0
    assert ok.sum() > n * 0.9
    for k in range(3):
        assert np.abs(got[k][ok] - exp[k][ok]).max() <= M_TOL
    for k in range(3, 6):
        assert np.abs(got[k][ok] - exp[k][ok]).max() <= 1e-9
    fa = np.isfinite(exp[6])
    assert np.array_equal(np.isfinite(got[6]), fa)
    assert np.abs(got[6][fa] - exp[6][fa]).max(initial=0.0) <= M_TOL
    import torch
    dev = gm.imageToProximateImagingLocusBatch(*(torch.as_tensor(a, device="cuda") for a in (line, samp, x, y, z)), 1e-5)
    torch.cuda.synchronize()
    for k in range(7):
        assert np.array_equal(dev[k].cpu().numpy(), got[k], equal_nan=True)


@pytest.mark.parametrize("kind", KINDS)
def test_ground_partials(ref, kind):
    from usgscsm_b200 import _capi, synth
    gm, rm = _pair(ref, kind)
    nl, ns = rm.image_size()
    rng = np.random.default_rng(13)
    n = 3000 if kind == "sar" else 12000
    line, samp = pixels(rng, n, nl, ns)
    x, y, z, _, _ = gm.imageToGroundBatch(line, samp, synth.synthetic_dem(line, samp))
    # some hopeless points too (far side of the body, NaN)
    x[:20] *= -1.0
    y[:20] *= -1.0
    z[:20] *= -1.0
    x[20:25] = np.nan
    p, st = gm.computeGroundPartialsBatch(x, y, z)
    ep, est = rm.ground_partials(x, y, z, threads=0)
    assert np.array_equal(st & _capi.ST_REFERENCE_MASK, est)
    ok = np.isfinite(ep).all(axis=1)
    assert np.array_equal(np.isfinite(p).all(axis=1), ok)
    assert ok.sum() > n * 0.9
    # partials are pixel differences (each within PX_TOL of the reference's) divided by the step
    scale = np.abs(ep[ok]).max()
    assert np.abs(p[ok] - ep[ok]).max() <= 1e-9 * max(1.0, scale)
    one = gm.computeGroundPartials((x[100], y[100], z[100]))
    assert one == [float(v) for v in p[100]]
    import torch
    dp, dst = gm.computeGroundPartialsBatch(*(torch.as_tensor(a, device="cuda") for a in (x, y, z)))
    torch.cuda.synchronize()
    assert np.array_equal(dp.cpu().numpy(), p, equal_nan=True)
    assert np.array_equal(dst.cpu().numpy().view(np.uint32), st)


@pytest.mark.parametrize("kind", ["linescan", "pushframe", "frame", "sar"])
def test_set_ellipsoid(ref, kind):
    from usgscsm_b200 import synth
    gm, rm = _pair(ref, kind)
    a, b = gm.getEllipsoid()
    gm.setEllipsoid(a + 1500.0, b - 2500.0)
    rm.set_ellipsoid(a + 1500.0, b - 2500.0)
    assert gm.getEllipsoid() == (a + 1500.0, b - 2500.0)
    nl, ns = rm.image_size()
    rng = np.random.default_rng(14)
    n = 4000
    line, samp = pixels(rng, n, nl, ns)
    x, y, z = check_i2g(gm, rm, line, samp, synth.synthetic_dem(line, samp))
    check_g2i(gm, rm, x, y, z)


@pytest.mark.parametrize("kind", ["linescan", "linescan_adjusted"])
@pytest.mark.parametrize("index", [0, 2, 4, 7, 11, 14, 15])
def test_sensor_partials_linescan(ref, kind, index):
    """computeSensorPartials LS:1336-1396: both overloads, every kind of adjustable parameter
    (position offsets, velocity, attitude bias / rate / acceleration, focal length)."""
    from usgscsm_b200 import _capi, synth
    gm, rm = _pair(ref, kind)
    nl, ns = rm.image_size()
    rng = np.random.default_rng(16 + index)
    n = 6000
    line, samp = pixels(rng, n, nl, ns)
    x, y, z, _, _ = gm.imageToGroundBatch(line, samp, synth.synthetic_dem(line, samp))
    x[:10] = np.nan
    dl, ds, ach, st = gm.computeSensorPartialsBatch(index, x, y, z, 1e-6)
    el, es, ea, est = rm.sensor_partials(index, x, y, z, 1e-6, threads=0)
    assert np.array_equal(st & _capi.ST_REFERENCE_MASK, est)
    ok = np.isfinite(el)
    assert np.array_equal(np.isfinite(dl), ok) and ok.sum() == n - 10
    scale = max(1.0, np.abs(el[ok]).max(), np.abs(es[ok]).max())
    assert np.abs(dl[ok] - el[ok]).max() <= 1e-9 * scale
    assert np.abs(ds[ok] - es[ok]).max() <= 1e-9 * scale
    assert np.abs(ach[ok] - ea[ok]).max() <= PX_TOL
    # the overload that is handed the image point
    dl2, ds2, _, st2 = gm.computeSensorPartialsBatch(index, x, y, z, 1e-6, imagePts=(line, samp))
    el2, es2, _, est2 = rm.sensor_partials(index, x, y, z, 1e-6, image_pts=(line, samp), threads=0)
    assert np.array_equal(st2 & _capi.ST_REFERENCE_MASK, est2)
    assert np.abs(dl2[ok] - el2[ok]).max() <= 1e-9 * scale
    assert np.abs(ds2[ok] - es2[ok]).max() <= 1e-9 * scale
    import torch
    d = gm.computeSensorPartialsBatch(index, *(torch.as_tensor(a, device="cuda") for a in (x, y, z)), 1e-6)
    torch.cuda.synchronize()
    assert np.array_equal(d[0].cpu().numpy(), dl, equal_nan=True)
    assert np.array_equal(d[1].cpu().numpy(), ds, equal_nan=True)


@pytest.mark.parametrize("kind,index", [("pushframe", 0), ("pushframe", 5), ("pushframe", 8), ("pushframe", 15),
                                        ("frame", 0), ("frame", 2), ("frame", 3), ("frame", 6),
                                        ("frame_transverse", 4), ("sar", 0), ("sar", 2), ("sar", 4)])
def test_sensor_partials_other_models(ref, kind, index):
    """PF:1121-1162 (step m_gsd, per-framelet orientation table rebuilt for the adjusted model),
    FR:814-873 (step 1.0 / 0.01), SAR:1308-1371 (step m_scaledPixelWidth)."""
    from usgscsm_b200 import _capi, synth
    gm, rm = _pair(ref, kind)
    nl, ns = rm.image_size()
    rng = np.random.default_rng(40 + index)
    n = 1500 if kind == "sar" else 5000
    line, samp = pixels(rng, n, nl, ns)
    x, y, z, _, _ = gm.imageToGroundBatch(line, samp, synth.synthetic_dem(line, samp))
    x[:5] = np.nan
    for image_pts in (None, (line, samp)):
        dl, ds, ach, st = gm.computeSensorPartialsBatch(index, x, y, z, 1e-6, imagePts=image_pts)
        el, es, ea, est = rm.sensor_partials(index, x, y, z, 1e-6, image_pts=image_pts, threads=0)
        assert np.array_equal(st & _capi.ST_REFERENCE_MASK, est)
        ok = np.isfinite(el) & np.isfinite(es)
        assert np.array_equal(np.isfinite(dl) & np.isfinite(ds), ok) and ok.sum() >= n * 0.9
        scale = max(1.0, np.abs(el[ok]).max(), np.abs(es[ok]).max())
        assert np.abs(dl[ok] - el[ok]).max() <= 1e-9 * scale
        assert np.abs(ds[ok] - es[ok]).max() <= 1e-9 * scale
        fa = np.isfinite(ea)
        assert np.array_equal(np.isfinite(ach), fa)
        assert np.abs(ach[fa] - ea[fa]).max(initial=0.0) <= PX_TOL


def test_sensor_partials_bad_index():
    import usgscsm_b200 as ub
    from usgscsm_b200 import _capi, synth
    gm = ub.UsgsAstroPlugin().constructModelFromState(synth.frame_state())
    with pytest.raises(_capi.EngineError) as e:
        gm.computeSensorPartialsBatch(7, np.zeros(3), np.zeros(3), np.zeros(3))
    assert e.value.code == _capi.EINVAL


@pytest.mark.parametrize("kind", KINDS)
def test_all_sensor_partials(ref, kind):
    """computeAllSensorPartials (csm/RasterGM.cpp:52-140 via LS:1413-1455, PF:1170-1201,
    FR:894-940; SAR: the RasterGM default): both overloads against the reference, bit-equal to
    the per-parameter call, index subsets, device flavour."""
    import torch
    from usgscsm_b200 import _capi, synth
    gm, rm = _pair(ref, kind)
    nl, ns = rm.image_size()
    npar = gm.getNumParameters()
    rng = np.random.default_rng(77)
    n = 600 if kind == "sar" else 3000
    line, samp = pixels(rng, n, nl, ns)
    x, y, z, _, _ = gm.imageToGroundBatch(line, samp, synth.synthetic_dem(line, samp))
    x[:7] = np.nan
    for image_pts in (None, (line, samp)):
        p, ach, st = gm.computeAllSensorPartialsBatch(x, y, z, 1e-6, imagePts=image_pts)
        ep, ea, est = rm.all_sensor_partials(npar, x, y, z, 1e-6, image_pts=image_pts, threads=0)
        assert p.shape == (n, npar, 2)
        assert np.array_equal(st & _capi.ST_REFERENCE_MASK, est)
        ok = np.isfinite(ep).all(axis=(1, 2))
        assert np.array_equal(np.isfinite(p).all(axis=(1, 2)), ok) and ok.sum() >= n * 0.9
        assert np.isnan(p[~ok]).all()          # the scalar call threw: no partial survives
        scale = np.maximum(1.0, np.abs(ep[ok]).max(axis=(0, 2)))[None, :, None]
        assert (np.abs(p[ok] - ep[ok]) <= 1e-9 * scale).all()
        fa = np.isfinite(ea)
        assert np.array_equal(np.isfinite(ach), fa)
        assert np.abs(ach[fa] - ea[fa]).max(initial=0.0) <= PX_TOL
        # slot k is what the single-parameter call gives
        for k in (0, npar // 2, npar - 1):
            dl, ds, _, _ = gm.computeSensorPartialsBatch(k, x, y, z, 1e-6, imagePts=image_pts)
            assert np.array_equal(p[ok, k, 0], dl[ok]) and np.array_equal(p[ok, k, 1], ds[ok])
    # a parameter subset (what getParameterSetIndices(pSet) would select), device arrays
    idx = [npar - 1, 1]
    sub, _, _ = gm.computeAllSensorPartialsBatch(x, y, z, 1e-6, indices=idx)
    assert np.array_equal(sub[ok], p_sub(gm, x, y, z, idx)[ok])
    d = gm.computeAllSensorPartialsBatch(*(torch.as_tensor(a, device="cuda") for a in (x, y, z)), 1e-6)
    torch.cuda.synchronize()
    full, ach0, st0 = gm.computeAllSensorPartialsBatch(x, y, z, 1e-6)
    assert np.array_equal(d[0].cpu().numpy(), full, equal_nan=True)
    assert np.array_equal(d[1].cpu().numpy(), ach0, equal_nan=True)
    assert np.array_equal(d[2].cpu().numpy().view(np.uint32), st0)


def p_sub(gm, x, y, z, idx):
    out = np.empty((x.size, len(idx), 2))
    for j, k in enumerate(idx):
        out[:, j, 0], out[:, j, 1], _, _ = gm.computeSensorPartialsBatch(k, x, y, z, 1e-6)
    return out


def test_all_sensor_partials_bad_arguments():
    import usgscsm_b200 as ub
    from usgscsm_b200 import _capi, synth
    gm = ub.UsgsAstroPlugin().constructModelFromState(synth.frame_state())
    with pytest.raises(_capi.EngineError) as e:
        gm.computeAllSensorPartialsBatch(np.zeros(3), np.zeros(3), np.zeros(3), indices=[0, 7])
    assert e.value.code == _capi.EINVAL
    p, ach, st = gm.computeAllSensorPartialsBatch(np.zeros(0), np.zeros(0), np.zeros(0))
    assert p.shape == (0, 7, 2)
