"""Parity of the SAR CUDA kernels with the reference's per-point CPU calls
(UsgsAstroSarSensorModel::groundToImage SAR:541-593 / imageToGround SAR:775-861)."""
import numpy as np
import pytest

from parity_util import check_g2i, check_i2g, models, nasty_ground_points, pixels

pytestmark = pytest.mark.gpu

CASES = {
    "orbital_right": dict(),
    "orbital_left": dict(look="left"),
    "ellipsoid": dict(major_axis=3396190.0, minor_axis=3376200.0, gm=4.2828e13),
    "adjusted": dict(params=(10, -5, 3, 0.1, -0.2, 0.05)),
    "one_scale_set": dict(n_scale_sets=1),
    "two_scale_sets": dict(n_scale_sets=2),
    "short_image": dict(n_lines=300, n_samples=200),
    "coarse_ephemeris": dict(dt_ephem=20.0),          # large table step: order reduces at the ends
    "single_node": dict(single_node=True, n_lines=2000),  # constant position: no Doppler bracket
}


@pytest.mark.parametrize("case", sorted(CASES))
def test_round_trip_parity(ref, case):
    from usgscsm_b200 import synth
    state = synth.sar_state(seed=11, **CASES[case])
    gm, rm = models(ref, state)
    nl, ns = rm.image_size()
    rng = np.random.default_rng(5)
    n = 20000
    line, samp = pixels(rng, n, nl, ns)
    h = rng.uniform(-1000.0, 1000.0, n)
    x, y, z = check_i2g(gm, rm, line, samp, h)
    check_g2i(gm, rm, x, y, z)
    check_g2i(gm, rm, x, y, z, prec=1e-8)
    check_i2g(gm, rm, line[:2000], samp[:2000], h[:2000], prec=1e-9)


def test_failures_and_out_of_image(ref):
    """Points outside the Doppler bracket (thrown std::invalid_argument -> UNKNOWN_ERROR), far
    side, origin, NaN; pixels outside the image; slant ranges that cannot be met."""
    from usgscsm_b200 import synth
    state = synth.sar_state(seed=13)
    gm, rm = models(ref, state)
    nl, ns = rm.image_size()
    rng = np.random.default_rng(8)
    line, samp = pixels(rng, 4000, nl, ns)
    x, y, z = check_i2g(gm, rm, line, samp, np.zeros(1))
    gx, gy, gz = nasty_ground_points(rng, x, y, z, 1.7374e6)
    check_g2i(gm, rm, gx, gy, gz)
    # lines well outside the image: i2g extrapolates the ephemeris, g2i leaves the bracket
    ol = np.concatenate([rng.uniform(-3 * nl, 0, 600), rng.uniform(nl, 4 * nl, 600)])
    osamp = rng.uniform(-500, ns + 500, ol.size)
    ox, oy, oz = check_i2g(gm, rm, ol, osamp, np.zeros(1))
    check_g2i(gm, rm, ox, oy, oz)
    check_i2g(gm, rm, line[:500], samp[:500] * 200 - 1e5, np.zeros(1))
    check_i2g(gm, rm, line[:500], samp[:500], np.full(500, -1.0e6))
    check_i2g(gm, rm, np.array([np.nan, np.inf, -np.inf, 1e300]), np.full(4, 10.0), np.zeros(1))


def test_set_parameter(ref):
    from usgscsm_b200 import synth
    state = synth.sar_state(seed=17, n_lines=8000)
    gm, rm = models(ref, state)
    nl, ns = rm.image_size()
    rng = np.random.default_rng(3)
    line, samp = pixels(rng, 3000, nl, ns)
    for idx, val in ((0, 12.0), (2, -30.0), (3, 0.5), (5, -0.2)):
        gm.setParameterValue(idx, val)
        rm.set_parameter(idx, val)
        x, y, z = check_i2g(gm, rm, line, samp, np.zeros(1))
        check_g2i(gm, rm, x, y, z)


def test_grid_matches_explicit(ref):
    import torch
    from usgscsm_b200 import synth
    state = synth.sar_state(seed=2)
    gm, rm = models(ref, state)
    nl, ns = 96, 80
    r = np.repeat(np.arange(nl) * 600.0 + 0.5, ns)
    c = np.tile(np.arange(ns) * 25.0 + 0.5, nl)
    h = synth.synthetic_dem(r, c)
    x, y, z, a, st = gm.imageToGroundGrid(0.5, 0.5, 600.0, 25.0, nl, ns,
                                          torch.as_tensor(h, device="cuda:0"), achieved=True)
    ex, ey, ez, ea, est = gm.imageToGroundBatch(r, c, h)
    assert np.array_equal(x.cpu().numpy(), ex) and np.array_equal(y.cpu().numpy(), ey)
    assert np.array_equal(z.cpu().numpy(), ez) and np.array_equal(a.cpu().numpy(), ea)
    assert np.array_equal(st.cpu().numpy().view(np.uint32), est)
    check_i2g(gm, rm, r, c, h)
