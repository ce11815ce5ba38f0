"""Parity of the push-frame CUDA kernels with the reference's per-point CPU calls
(UsgsAstroPushFrameSensorModel::groundToImage PF:715-778 / imageToGround PF:842-870)."""
import numpy as np
import pytest

from parity_util import check_g2i, check_i2g, models, nasty_ground_points, pixels

pytestmark = pytest.mark.gpu

TRANSVERSE_MILD = (0.0, 1.0, 0.0, 1e-6, 2e-6, -1e-6, 1e-6, 0.0, 2e-6, 0.0,
                   0.0, 0.0, 1.0, -1e-6, 1e-6, 2e-6, 0.0, 1e-6, 0.0, -2e-6)

CASES = {
    "wac_radial": dict(),
    "line_sign_plus": dict(line_sign=1.0),
    "flipped": dict(flipped=True),
    "reversed": dict(reversed_order=True),
    "flipped_reversed": dict(flipped=True, reversed_order=True, line_sign=1.0),
    "summing2": dict(summing=2.0),
    "transverse": dict(distortion_type=1, dist_coeffs=TRANSVERSE_MILD),
    "ellipsoid": dict(major_axis=3396190.0, minor_axis=3376200.0, gm=4.2828e13, altitude=300000.0),
    "platform0": dict(platform_flag=0),
    "no_overlap": dict(overlap=0),
    "overlap4": dict(overlap=4),
    "adjusted": dict(params=(3.0, -2.0, 1.5, 0.02, -0.01, 0.005, 0.5, -0.3, 0.2, 0.05, 0.04, -0.03,
                             0.01, -0.02, 0.015, 0.0005)),
    "position_only": dict(params=(5.0, 0, 0) + (0,) * 13),
    # 5000 framelets: the per-framelet orientation table exceeds the staged size (L1/L2 path)
    "many_framelets": dict(n_framelets=5000, framelet_height=4, n_samples=512, dt_quat=1.0),
    "few_framelets": dict(n_framelets=2, framelet_height=64, n_samples=256),
    "one_framelet": dict(n_framelets=1, framelet_height=64, n_samples=256),
}


@pytest.mark.parametrize("case", sorted(CASES))
def test_round_trip_parity(ref, case):
    from usgscsm_b200 import synth
    state = synth.pushframe_state(seed=11, **CASES[case])
    gm, rm = models(ref, state)
    nl, ns = rm.image_size()
    sz = gm.getImageSize()
    assert (sz.line, sz.samp) == (nl, ns)
    rng = np.random.default_rng(5)
    n = 20000
    line, samp = pixels(rng, n, nl, ns)
    h = synth.synthetic_dem(line, samp)
    x, y, z = check_i2g(gm, rm, line, samp, h)
    check_g2i(gm, rm, x, y, z)
    check_g2i(gm, rm, x, y, z, prec=1e-8)
    # exact framelet boundaries and integer lines
    bl = np.arange(0, min(nl, 2000), dtype=np.float64)
    check_i2g(gm, rm, bl, np.full(bl.size, ns / 3.0), np.zeros(1))


def test_failures_and_out_of_image(ref):
    from usgscsm_b200 import synth
    state = synth.pushframe_state(seed=13)
    gm, rm = models(ref, state)
    nl, ns = rm.image_size()
    rng = np.random.default_rng(8)
    line, samp = pixels(rng, 4000, nl, ns)
    x, y, z = check_i2g(gm, rm, line, samp, np.zeros(1))
    # rays that miss the body, pixels far outside the image (framelet numbers outside the table,
    # negative lines: fmod and the int truncation keep the sign)
    check_i2g(gm, rm, line[:500], samp[:500], np.full(500, -1.7e6))
    check_i2g(gm, rm, line[:500] * 40 - 1e5, samp[:500] * 3 - 2000, np.zeros(1))
    check_i2g(gm, rm, -line[:500], samp[:500], np.zeros(1))
    check_i2g(gm, rm, np.array([np.nan, np.inf, -np.inf, 1e300, -1e300, 3e9, -3e9]),
              np.full(7, 10.0), np.zeros(1))
    gx, gy, gz = nasty_ground_points(rng, x, y, z, 1.7374e6)
    check_g2i(gm, rm, gx, gy, gz)
    # ground points ahead of / behind the image: the secant clips to framelet 0 / numFramelets
    check_g2i(gm, rm, x[:300] * 1.002, y[:300] * 0.998, z[:300])


def test_set_parameter_rebuilds_orientations(ref):
    from usgscsm_b200 import synth
    state = synth.pushframe_state(seed=17)
    gm, rm = models(ref, state)
    nl, ns = rm.image_size()
    rng = np.random.default_rng(3)
    line, samp = pixels(rng, 3000, nl, ns)
    for idx, val in ((0, 12.0), (7, 0.4), (15, 0.001), (4, -0.02)):
        gm.setParameterValue(idx, val)
        rm.set_parameter(idx, val)
        x, y, z = check_i2g(gm, rm, line, samp, np.zeros(1))
        check_g2i(gm, rm, x, y, z)


def test_grid_matches_explicit(ref):
    import torch
    from usgscsm_b200 import synth
    state = synth.pushframe_state(seed=2)
    gm, rm = models(ref, state)
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
    check_i2g(gm, rm, r, c, h)
