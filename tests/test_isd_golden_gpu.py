"""The reference's OWN asserted values (tests/ref_goldens.py <- /root/reference/tests/*.cpp), run
through the CUDA engine built from the reference's ISD fixtures by the product's ISD -> state
converter -- the path a user of the reference takes: constructModelFromISD, then the per-point
calls.  Tolerances are the reference's own.  Then parity at the bar (<= 1e-4 m, <= 1e-6 px,
observable flags exact) against the oracle on the SAME state, for every fixture, and config C1
of BASELINE.json: the usgscsm_cam_test loop over the full 16 x 16 grid of
constVelocityLineScan.json (bin/usgscsm_cam_test.cc:294-309)."""
import numpy as np
import pytest

import ref_goldens as G
from parity_util import check_g2i, check_i2g

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def engine():
    import usgscsm_b200 as ub
    plugin = ub.UsgsAstroPlugin(device=0)
    cache = {}

    def get(name):
        if name not in cache:
            cache[name] = plugin.constructModelFromISD(G.isd_image(name), None)
        return cache[name]
    return get


@pytest.mark.parametrize("case", G.I2G, ids=lambda c: "%s@%s" % (c[0], c[5]))
def test_image_to_ground_goldens(engine, case):
    name, (line, samp), h, exp, tol, _ = case
    x, y, z, _, _ = engine(name).imageToGroundBatch([line], [samp], h)
    assert abs(x[0] - exp[0]) <= tol and abs(y[0] - exp[1]) <= tol and abs(z[0] - exp[2]) <= tol


@pytest.mark.parametrize("case", G.G2I, ids=lambda c: "%s@%s" % (c[0], c[5]))
def test_ground_to_image_goldens(engine, case):
    name, (x, y, z), prec, exp, tol, _ = case
    l, s, ach, _ = engine(name).groundToImageBatch([x], [y], [z], prec)
    assert abs(l[0] - exp[0]) <= tol and abs(s[0] - exp[1]) <= tol


def test_round_trip_goldens(engine):
    for name, (line, samp, h, prec), tol, cite in G.ROUND_TRIP:
        m = engine(name)
        x, y, z, _, _ = m.imageToGroundBatch([line], [samp], h)
        l, s, ach, _ = m.groundToImageBatch(x, y, z, prec)
        assert abs(l[0] - line) <= tol and abs(s[0] - samp) <= tol, (name, line, samp, cite)
        if name != "orbitalPushFrame" and "Framer" not in name:
            assert abs(ach[0]) < 0.001


def test_pushframe_inversion(engine):
    name, nl, ns, tol, _ = G.PUSHFRAME_INVERSION
    m = engine(name)
    line = np.repeat(np.arange(nl) + 0.5, ns)
    samp = np.tile(np.arange(ns) + 0.5, nl)
    x, y, z, _, _ = m.imageToGroundBatch(line, samp, 0.0)
    l, s, _, _ = m.groundToImageBatch(x, y, z)
    assert np.abs(l - line).max() <= tol and np.abs(s - samp).max() <= tol


@pytest.mark.parametrize("name", G.ISD_FIXTURES)
def test_fixture_parity_on_the_same_state(engine, ref, name):
    """Engine from the ISD; the oracle from the engine's saved state (so both sides hold the same
    tables): full bar over a grid covering the image, at 0.001 and 1e-8."""
    gm = engine(name)
    rm = ref.RefModel(gm.getModelState())
    if gm.getModelName() == "USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL":
        d = gm.lsDerived()
        rm.ls_set_projective(d[4] != 0, d[5:19])
    sz = gm.getImageSize()
    nl, ns = int(sz.line), int(sz.samp)
    line = np.repeat(np.linspace(0.0, nl, 33), 31)
    samp = np.tile(np.linspace(0.0, ns, 31), 33)
    a, _ = gm.getEllipsoid()
    for h in (0.0, 1e-4 * a):
        x, y, z = check_i2g(gm, rm, line, samp, np.full(line.size, h))
        for prec in (0.001, 1e-8):
            check_g2i(gm, rm, x, y, z, prec=prec)


@pytest.mark.parametrize("name", G.ISD_FIXTURES)
def test_fixture_vs_oracle_from_isd(engine, ref, name):
    """Engine and oracle each convert the ISD themselves (product converter vs the reference's
    constructStateFromIsd on the restated ALE): the tables agree to ~1e-13 relative
    (tests/test_isd_state.py), so do the results."""
    gm = engine(name)
    rm = ref.RefModel.from_isd(G.isd_image(name))
    sz = gm.getImageSize()
    assert (int(sz.line), int(sz.samp)) == tuple(rm.image_size())
    nl, ns = int(sz.line), int(sz.samp)
    line = np.repeat(np.linspace(0.5, nl - 0.5, 17), 15)
    samp = np.tile(np.linspace(0.5, ns - 0.5, 15), 17)
    x, y, z, _, st = gm.imageToGroundBatch(line, samp, 0.0)
    rx, ry, rz, _, rst = rm.image_to_ground(line, samp, [0.0])
    a, _ = gm.getEllipsoid()
    tol = max(1e-4, 1e-10 * a)  # metres
    assert max(np.abs(x - rx).max(), np.abs(y - ry).max(), np.abs(z - rz).max()) <= tol
    l, s, _, _ = gm.groundToImageBatch(rx, ry, rz, 1e-8)
    rl, rs, _, _ = rm.ground_to_image(rx, ry, rz, 1e-8)
    ptol = 10.0 if "PushFrame" in name else 1e-6
    assert max(np.abs(l - rl).max(), np.abs(s - rs).max()) <= ptol


def test_config_c1_cam_test_loop(engine, ref):
    """BASELINE.json configs[0]: tests/data constant-velocity line-scan ISD, imageToGround ->
    groundToImage over the full pixel grid, as bin/usgscsm_cam_test.cc:294-309 does it
    (samples outer, lines inner, sample rate 1, height 0, precision 0.001)."""
    gm = engine("constVelocityLineScan")
    rm = ref.RefModel.from_isd(G.isd_image("constVelocityLineScan"))
    samp, line = [a.ravel().astype(float) for a in np.meshgrid(np.arange(16), np.arange(16), indexing="ij")]
    x, y, z, ach, st = gm.imageToGroundBatch(line, samp, 0.0, 0.001)
    l, s, ach2, st2 = gm.groundToImageBatch(x, y, z, 0.001)
    err = np.hypot(l - line, s - samp)
    assert err.max() < 1e-3
    rx, ry, rz, _, rst = rm.image_to_ground(line, samp, [0.0], 0.001)
    rl, rs, _, rst2 = rm.ground_to_image(rx, ry, rz, 0.001)
    rerr = np.hypot(rl - line, rs - samp)
    assert max(np.abs(x - rx).max(), np.abs(y - ry).max(), np.abs(z - rz).max()) <= 1e-4
    assert np.abs(np.sort(err) - np.sort(rerr)).max() <= 1e-6   # the report's min / median / max


def test_plugin_isd_methods(engine):
    import usgscsm_b200 as ub
    p = ub.UsgsAstroPlugin(device=0)
    img = G.isd_image("orbitalLineScan")
    assert p.canModelBeConstructedFromISD(img, "USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL")
    warn = []
    assert not p.canModelBeConstructedFromISD(G.isd_image("empty"), "USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL", warn)
    assert len(warn) == 1
    assert not p.canModelBeConstructedFromISD(G.isd_image("no_such_file"), "USGS_ASTRO_FRAME_SENSOR_MODEL")
    state = p.getStateFromISD(img)
    assert state == p.convertISDToModelState(img, "USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL")
    assert p.canISDBeConvertedToModelState(img, "USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL")
    # the converted state reloads to the same model
    m1, m2 = engine("orbitalLineScan"), p.constructModelFromState(state)
    a = m1.imageToGroundBatch([3.5, 9.25], [2.0, 11.5], 0.0)
    b = m2.imageToGroundBatch([3.5, 9.25], [2.0, 11.5], 0.0)
    for u, v in zip(a, b):
        assert np.array_equal(u, v)
