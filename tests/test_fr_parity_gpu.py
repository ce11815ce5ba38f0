"""Parity of the framing-camera CUDA kernels with the reference's per-point CPU calls
(UsgsAstroFrameSensorModel::groundToImage FR:137-232 / imageToGround FR:280-372)."""
import numpy as np
import pytest

from parity_util import check_g2i, check_i2g, models, nasty_ground_points, pixels

pytestmark = pytest.mark.gpu

TRANSVERSE_MILD = (0.0, 1.0, 0.0, 1e-6, 2e-6, -1e-6, 1e-8, 0.0, 2e-8, 0.0,
                   0.0, 0.0, 1.0, -1e-6, 1e-6, 2e-6, 0.0, 1e-8, 0.0, -2e-8)

CASES = {
    "radial": dict(),
    "transverse": dict(distortion_type=1, dist_coeffs=TRANSVERSE_MILD),
    "kaguya": dict(distortion_type=2, dist_coeffs=(
        -0.0725, -0.0009649900000000001, 0.00098441, 8.5773e-06, -3.7438e-06,
        0.0214, -0.0013796, 1.3502e-05, 2.7251e-06, -6.193800000000001e-06)),
    "dawnfc": dict(distortion_type=3, dist_coeffs=(8.4e-06,)),
    "lronac": dict(distortion_type=4, dist_coeffs=(1.81e-5,)),
    "cahvor": dict(distortion_type=5, dist_coeffs=(1e-5, -2e-8, 1e-11, 0.01, -0.02)),
    "lunarorbiter": dict(distortion_type=6, dist_coeffs=(0.0,)),
    "radtan": dict(distortion_type=7, dist_coeffs=(1e-3, -2e-4, 1e-4, -1e-4, 1e-5)),
    "kplo": dict(distortion_type=8, dist_coeffs=(1.81e-5,)),
    "ellipsoid": dict(major_axis=3396190.0, minor_axis=3376200.0, altitude=300000.0),
    "summing2_offset": dict(summing=2, n_lines=2048, n_samples=2048, starting_sample=8.0, starting_line=4.0),
    "unnormalised_quaternion": dict(unnormalised=1.0 + 1e-7),
    "jitter": dict(n_lines=1024, n_samples=1024,
                   jitter=((1e-2, -3e-2, 0.5), (-2e-2, 1e-2, -0.3), 1e-3)),
    "jitter_strong": dict(n_lines=512, n_samples=512,
                          jitter=((5.0, 2.0), (-3.0, 1.0), 2e-3)),
}


@pytest.mark.parametrize("case", sorted(CASES))
def test_round_trip_parity(ref, case):
    from usgscsm_b200 import synth
    state = synth.frame_state(seed=11, **CASES[case])
    gm, rm = models(ref, state)
    nl, ns = rm.image_size()
    rng = np.random.default_rng(5)
    n = 20000
    line, samp = pixels(rng, n, nl, ns)
    h = synth.synthetic_dem(line, samp)
    x, y, z = check_i2g(gm, rm, line, samp, h)
    l, s = check_g2i(gm, rm, x, y, z)
    check_g2i(gm, rm, x, y, z, prec=1e-8)
    if case not in ("unnormalised_quaternion",):
        assert np.abs(l - line).max() < 5e-2 and np.abs(s - samp).max() < 5e-2


def test_failures_and_out_of_image(ref):
    from usgscsm_b200 import synth
    state = synth.frame_state(seed=13)
    gm, rm = models(ref, state)
    nl, ns = rm.image_size()
    rng = np.random.default_rng(8)
    line, samp = pixels(rng, 4000, nl, ns)
    x, y, z = check_i2g(gm, rm, line, samp, np.zeros(1))
    check_i2g(gm, rm, line[:500], samp[:500], np.full(500, -1.7e6))      # rays miss the body
    check_i2g(gm, rm, line[:500] * 40 - 1e5, samp[:500] * 30 - 2e4, np.zeros(1))
    check_i2g(gm, rm, np.array([np.nan, np.inf, -np.inf, 1e300]), np.full(4, 10.0), np.zeros(1))
    gx, gy, gz = nasty_ground_points(rng, x, y, z, 1.7374e6)
    check_g2i(gm, rm, gx, gy, gz)
    # a ground point in the camera's focal plane (denominator ~ 0) and at the camera itself
    p = np.array([gm.getParameterValue(i) for i in range(3)])
    check_g2i(gm, rm, [p[0], p[0] + 1.0], [p[1], p[1]], [p[2], p[2] + 1.0])


def test_jitter_out_of_range_lines(ref):
    """Rolling-shutter jitter: line indices are clamped to the line-time table (U:203)."""
    from usgscsm_b200 import synth
    state = synth.frame_state(seed=3, **CASES["jitter"])
    gm, rm = models(ref, state)
    line = np.array([-50.0, -0.6, 0.0, 0.49, 0.5, 1.5, 1023.4, 1023.5, 1024.0, 5000.0, 1e12, np.nan])
    samp = np.linspace(0.0, 1000.0, line.size)
    check_i2g(gm, rm, line, samp, np.zeros(1))


def test_set_parameter(ref):
    from usgscsm_b200 import synth
    state = synth.frame_state(seed=17)
    gm, rm = models(ref, state)
    nl, ns = rm.image_size()
    rng = np.random.default_rng(3)
    line, samp = pixels(rng, 3000, nl, ns)
    for idx, delta in ((0, 12.0), (2, -30.0), (3, 1e-4), (6, -2e-4)):
        v = gm.getParameterValue(idx) + delta
        gm.setParameterValue(idx, v)
        rm.set_parameter(idx, v)
        x, y, z = check_i2g(gm, rm, line, samp, np.zeros(1))
        check_g2i(gm, rm, x, y, z)


def test_grid_matches_explicit(ref):
    import torch
    from usgscsm_b200 import synth
    state = synth.frame_state(seed=2)
    gm, rm = models(ref, state)
    nl, ns = 96, 80
    r = np.repeat(np.arange(nl) * 40.0 + 0.5, ns)
    c = np.tile(np.arange(ns) * 12.0 + 0.5, nl)
    h = synth.synthetic_dem(r, c)
    x, y, z, a, st = gm.imageToGroundGrid(0.5, 0.5, 40.0, 12.0, nl, ns,
                                          torch.as_tensor(h, device="cuda:0"), achieved=False)
    ex, ey, ez, ea, est = gm.imageToGroundBatch(r, c, h)
    assert np.array_equal(x.cpu().numpy(), ex) and np.array_equal(y.cpu().numpy(), ey)
    assert np.array_equal(z.cpu().numpy(), ez)
    assert np.array_equal(st.cpu().numpy().view(np.uint32), est)
    check_i2g(gm, rm, r, c, h)
