"""Parity of the line-scan CUDA kernels with the reference's per-point CPU calls.

Bar (BASELINE.json north_star): groundToImage <= 1e-6 px, imageToGround <= 1e-4 m, and the
status bits the reference makes observable (warnings, thrown errors, non-finite, out of
image) bit-exact.  Both sides get the same model state and the same host-fitted projective
seed (an input of the hot path).  Everything goes through the C-ABI.
"""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

PX_TOL = 1e-6   # pixels
M_TOL = 1e-4    # metres


def _models(ref, state, same_seed=True):
    import usgscsm_b200 as ub
    gm = ub.UsgsAstroPlugin(device=0).constructModelFromState(state)
    rm = ref.RefModel(state)
    if same_seed:
        d = gm.lsDerived()
        rm.ls_set_projective(d[4] != 0, d[5:19])
    return gm, rm


def _check_i2g(gm, rm, line, samp, h, prec=0.001):
    from usgscsm_b200 import _capi
    x, y, z, a, st = gm.imageToGroundBatch(line, samp, h, prec)
    rx, ry, rz, ra, rst = rm.image_to_ground(line, samp, h, prec)
    ok = np.isfinite(rx)
    assert np.array_equal(np.isfinite(x), ok)
    for u, v in ((x, rx), (y, ry), (z, rz)):
        assert np.abs(u[ok] - v[ok]).max(initial=0.0) <= M_TOL
    assert np.abs(a[ok] - ra[ok]).max(initial=0.0) <= M_TOL
    assert np.array_equal(st & _capi.ST_REFERENCE_MASK, rst)
    return x, y, z


def _check_g2i(gm, rm, x, y, z, prec=0.001):
    from usgscsm_b200 import _capi
    l, s, a, st = gm.groundToImageBatch(x, y, z, prec)
    rl, rs, ra, rst = rm.ground_to_image(x, y, z, prec)
    ok = np.isfinite(rl) & np.isfinite(rs)
    assert np.array_equal(np.isfinite(l) & np.isfinite(s), ok)
    assert np.abs(l[ok] - rl[ok]).max(initial=0.0) <= PX_TOL
    assert np.abs(s[ok] - rs[ok]).max(initial=0.0) <= PX_TOL
    assert np.abs(a[ok] - ra[ok]).max(initial=0.0) <= PX_TOL
    assert np.array_equal(st & _capi.ST_REFERENCE_MASK, rst)
    return l, s


def _pixels(rng, n, nl, ns):
    return rng.uniform(0, nl, n), rng.uniform(0, ns, n)


CASES = {
    "nac_sphere": dict(),
    "nac_ellipsoid": dict(major_axis=3396190.0, minor_axis=3376200.0, gm=4.2828e13, altitude=300000.0),
    "radial": dict(distortion_type=0, dist_coeffs=(1e-4, -2e-7, 3e-10)),
    "transverse": dict(distortion_type=1, dist_coeffs=(
        0.0, 1.0, 0.0, 1e-6, 2e-6, -1e-6, 1e-8, 0.0, 2e-8, 0.0,
        0.0, 0.0, 1.0, -1e-6, 1e-6, 2e-6, 0.0, 1e-8, 0.0, -2e-8)),
    "kaguya": dict(distortion_type=2, dist_coeffs=(
        -0.0725, -0.0009649900000000001, 0.00098441, 8.5773e-06, -3.7438e-06,
        0.0214, -0.0013796, 1.3502e-05, 2.7251e-06, -6.193800000000001e-06)),
    "dawnfc": dict(distortion_type=3, dist_coeffs=(8.4e-06,)),
    "cahvor": dict(distortion_type=5, dist_coeffs=(1e-5, -2e-8, 1e-11, 0.01, -0.02)),
    "lunarorbiter": dict(distortion_type=6, dist_coeffs=(0.0,)),
    "radtan": dict(distortion_type=7, dist_coeffs=(1e-3, -2e-4, 1e-4, -1e-4, 1e-5)),
    "kplo": dict(distortion_type=8, dist_coeffs=(1.81e-5,)),
    # LRO NAC coefficients outside |k1| <= 1e-4: the forward iteration keeps its divergence test
    "lro_strong": dict(distortion_type=4, dist_coeffs=(3.0e-4,)),
    "lro_negative": dict(distortion_type=4, dist_coeffs=(-2.5e-4,)),
    "platform0": dict(platform_flag=0),
    # HiRISE-like attitude sampling: 4 x 41 k quaternion doubles = 1.3 MB, not staged into shared
    # memory (kMaxStagedTableBytes) -- the kernels read the tables through L1/L2
    "unstaged_tables": dict(dt_quat=0.0001, dt_ephem=0.01),
    "ellipsoid": dict(major_axis=3396190.0, minor_axis=3376200.0, altitude=280000.0, gm=4.2828e13),
    "segments": dict(n_time_segments=3),
    "summing2": dict(summing=2, n_samples=2532, n_lines=8192, starting_sample=10.0),
    "adjusted": dict(params=(3.0, -2.0, 1.5, 0.02, -0.01, 0.005, 0.5, -0.3, 0.2, 0.05, 0.04, -0.03,
                             0.01, -0.02, 0.015, 0.05)),
    "focal_only": dict(params=(0,) * 15 + (0.1,)),
    "position_only": dict(params=(5.0, 0, 0) + (0,) * 13),
}


@pytest.mark.parametrize("case", sorted(CASES))
def test_round_trip_parity(ref, case):
    from usgscsm_b200 import synth
    kw = dict(n_lines=8192, n_samples=2048, seed=11)
    kw.update(CASES[case])
    state = synth.linescan_state(**kw)
    gm, rm = _models(ref, state)
    nl, ns = rm.image_size()
    rng = np.random.default_rng(5)
    n = 20000
    line, samp = _pixels(rng, n, nl, ns)
    h = synth.synthetic_dem(line, samp)
    x, y, z = _check_i2g(gm, rm, line, samp, h)
    _check_g2i(gm, rm, x, y, z)
    _check_g2i(gm, rm, x, y, z, prec=1e-8)


def test_own_projective_fit(ref):
    """Each side fits its own seed (Eigen-free QR on both): results still within tolerance."""
    from usgscsm_b200 import synth
    state = synth.linescan_state(n_lines=8192, n_samples=2048, seed=3)
    gm, rm = _models(ref, state, same_seed=False)
    d, rd = gm.lsDerived(), rm.ls_derived()
    assert d[4] == rd[4] == 1.0
    np.testing.assert_allclose(d[:4], rd[:4], rtol=1e-12)
    np.testing.assert_allclose(d[19:22], rd[19:22], atol=1e-6)
    rng = np.random.default_rng(6)
    line, samp = _pixels(rng, 5000, 8192, 2048)
    x, y, z = _check_i2g(gm, rm, line, samp, np.zeros(1))
    l, s, a, st = gm.groundToImageBatch(x, y, z, 1e-8)
    rl, rs, ra, rst = rm.ground_to_image(x, y, z, 1e-8)
    assert np.abs(l - rl).max() <= PX_TOL and np.abs(s - rs).max() <= PX_TOL


def test_no_seed_and_failures(ref):
    """Image-centre seed (no projective fit), rays that miss the body, ground points far
    outside the image and on the far side: warnings / flags must agree bit for bit."""
    from usgscsm_b200 import synth
    state = synth.linescan_state(n_lines=8192, n_samples=2048, seed=13)
    gm, rm = _models(ref, state)
    gm.lsSetProjective(False)
    rm.ls_set_projective(False)
    rng = np.random.default_rng(8)
    line, samp = _pixels(rng, 4000, 8192, 2048)
    x, y, z = _check_i2g(gm, rm, line, samp, np.zeros(1))
    _check_g2i(gm, rm, x, y, z)
    # off-body rays: a huge negative height makes the discriminant negative
    _check_i2g(gm, rm, line[:500], samp[:500], np.full(500, -1.7e6))
    # pixels far outside the image (Lagrange extrapolation at reduced order)
    _check_i2g(gm, rm, line[:500] * 40 - 1e5, samp[:500] * 3 - 2000, np.zeros(1))
    # ground points: scaled, mirrored, far side, origin
    pts = np.stack([x[:600], y[:600], z[:600]])
    pts[:, 100:200] *= 1.01
    pts[:, 200:300] *= -1.0
    pts[:, 300:400] = rng.normal(size=(3, 100)) * 1.7374e6
    pts[:, 400:410] = 0.0
    pts[:, 410:420] = np.nan
    _check_g2i(gm, rm, pts[0], pts[1], pts[2])


def test_grid_matches_explicit(ref):
    """The dense-grid flavour generates the pixel coordinates the explicit flavour is given."""
    import torch
    from usgscsm_b200 import synth
    state = synth.linescan_state(n_lines=4096, n_samples=1024, seed=2)
    gm, rm = _models(ref, state)
    nl, ns = 96, 80
    r = np.repeat(np.arange(nl) * 40.0 + 0.5, ns)
    c = np.tile(np.arange(ns) * 12.0 + 0.5, nl)
    h = synth.synthetic_dem(r, c)
    x, y, z, a, st = gm.imageToGroundGrid(0.5, 0.5, 40.0, 12.0, nl, ns,
                                          torch.as_tensor(h, device="cuda:0"), achieved=True)
    ex, ey, ez, ea, est = gm.imageToGroundBatch(r, c, h)
    assert np.array_equal(x.cpu().numpy(), ex) and np.array_equal(y.cpu().numpy(), ey)
    assert np.array_equal(z.cpu().numpy(), ez) and np.array_equal(a.cpu().numpy(), ea)
    assert np.array_equal(st.cpu().numpy().view(np.uint32), est)


def test_empty_and_single(ref):
    from usgscsm_b200 import synth
    state = synth.linescan_state(n_lines=4096, n_samples=1024, seed=2)
    gm, rm = _models(ref, state)
    out = gm.imageToGroundBatch(np.empty(0), np.empty(0), np.zeros(1))
    assert all(o.size == 0 for o in out)
    out = gm.groundToImageBatch(np.empty(0), np.empty(0), np.empty(0))
    assert all(o.size == 0 for o in out)
    g = gm.imageToGround((100.5, 200.5), 10.0)
    rx, ry, rz, _, _ = rm.image_to_ground([100.5], [200.5], [10.0])
    assert abs(g.x - rx[0]) <= M_TOL and abs(g.y - ry[0]) <= M_TOL and abs(g.z - rz[0]) <= M_TOL
    ach, warn = [], []
    ip = gm.groundToImage(g, 1e-9, ach, warn)
    assert abs(ip.line - 100.5) < 1e-3 and abs(ip.samp - 200.5) < 1e-3 and len(ach) == 1


def test_chunked_pipeline_large(ref):
    """More points than one pipeline chunk (4 Mi): the host flavour must stitch chunks exactly;
    checked by the round trip and against the oracle on a strided subsample."""
    from usgscsm_b200 import synth
    state = synth.linescan_state(seed=20261104)  # the full 52224 x 5064 NAC-style image
    gm, rm = _models(ref, state)
    n = 9_000_001
    rng = np.random.default_rng(1)
    line, samp = _pixels(rng, n, 52224, 5064)
    h = synth.synthetic_dem(line, samp)
    x, y, z, a, st = gm.imageToGroundBatch(line, samp, h)
    l, s, a2, st2 = gm.groundToImageBatch(x, y, z)
    from usgscsm_b200 import _capi
    keep = np.uint32(0xffffffff & ~_capi.ST_EVALS_MASK)
    assert not (st & keep).any() and not (st2 & keep).any()
    assert np.abs(l - line).max() < 5e-3 and np.abs(s - samp).max() < 1e-4
    sub = slice(0, n, 997)
    rx, ry, rz, _, _ = rm.image_to_ground(line[sub], samp[sub], h[sub])
    assert max(np.abs(rx - x[sub]).max(), np.abs(ry - y[sub]).max(), np.abs(rz - z[sub]).max()) <= M_TOL
    rl, rs, _, _ = rm.ground_to_image(x[sub], y[sub], z[sub])
    assert max(np.abs(rl - l[sub]).max(), np.abs(rs - s[sub]).max()) <= PX_TOL


def test_table_ends_reduced_order(ref):
    """Times near the ends of the ephemeris / attitude tables: lagrangeInterp drops to order
    6 / 4 / 2 (src/Utilities.cpp:437-452) and beyond the ends extrapolates from the clamped
    index.  The padding nodes of the synthetic state are cut off so that the first and last
    image lines sit on the table ends."""
    import json
    from usgscsm_b200 import synth
    d = synth.linescan_state(n_lines=4096, n_samples=1024, seed=17, as_dict=True)
    for key, vlen, t0, dt in (("m_positions", 3, "m_t0Ephem", "m_dtEphem"),
                              ("m_velocities", 3, "m_t0Ephem", None),
                              ("m_quaternions", 4, "m_t0Quat", "m_dtQuat")):
        cut = 4 if key != "m_quaternions" else 5   # attitude table ends INSIDE the image
        d[key] = d[key][cut * vlen:len(d[key]) - cut * vlen]
        if dt is not None:
            d[t0] = d[t0] + cut * d[dt]
    d["m_numPositions"] = len(d["m_positions"])
    d["m_numQuaternions"] = len(d["m_quaternions"])
    state = d["m_modelName"] + "\n" + json.dumps(d)
    gm, rm = _models(ref, state)
    nl, ns = rm.image_size()
    rng = np.random.default_rng(6)
    # dense near both ends, plus a sprinkling everywhere and a little outside the image
    line = np.concatenate([rng.uniform(-40, 400, 6000), rng.uniform(nl - 400, nl + 40, 6000),
                           rng.uniform(0, nl, 3000)])
    samp = rng.uniform(0, ns, line.size)
    h = synth.synthetic_dem(line, samp)
    x, y, z = _check_i2g(gm, rm, line, samp, h)
    _check_g2i(gm, rm, x, y, z)
    _check_g2i(gm, rm, x, y, z, prec=1e-8)
    got = gm.imageToRemoteImagingLocusBatch(line, samp)
    exp = rm.remote_locus(line, samp)
    for k in range(6):
        assert np.abs(got[k] - exp[k]).max() <= (1e-4 if k < 3 else 1e-12)


@pytest.mark.parametrize("kind", ["linescan", "pushframe", "frame", "sar"])
def test_host_grid_equals_explicit(kind):
    """usgscsm_b200_image_to_ground_grid (host arrays, chunked by whole rows, several chunks)
    against the explicit-pixel host call, per-point and single heights."""
    import usgscsm_b200 as ub
    from usgscsm_b200 import synth
    state = {"linescan": lambda: synth.linescan_state(n_lines=3000, n_samples=1999, seed=8),
             "pushframe": lambda: synth.pushframe_state(n_framelets=180, framelet_height=16, n_samples=2001),
             "frame": lambda: synth.frame_state(n_lines=2500, n_samples=2047),
             "sar": lambda: synth.sar_state(n_lines=2600, n_samples=2000)}[kind]()
    gm = ub.UsgsAstroPlugin().constructModelFromState(state)
    sz = gm.getImageSize()
    nl, ns = int(sz.line) - 3, int(sz.samp) - 5           # 5.9 M points: two pipeline chunks
    rng = np.random.default_rng(3)
    h = rng.uniform(-300, 300, nl * ns)
    line = np.repeat(1.25 + np.arange(nl), ns)
    samp = np.tile(2.5 + np.arange(ns), nl)
    got = gm.imageToGroundGrid(1.25, 2.5, 1.0, 1.0, nl, ns, h, 1e-5, achieved=True)
    exp = gm.imageToGroundBatch(line, samp, h, 1e-5)
    for u, v in zip(got, exp):
        assert np.array_equal(u, v, equal_nan=True)
    got = gm.imageToGroundGrid(1.25, 2.5, 1.0, 1.0, nl, ns, np.array([40.0]), 1e-5)
    exp = gm.imageToGroundBatch(line, samp, 40.0, 1e-5)
    for k in (0, 1, 2, 4):
        assert np.array_equal(got[k], exp[k], equal_nan=True)
