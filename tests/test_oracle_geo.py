"""Pins oracle/geo_restate.c (the PROJ arithmetic of UsgsAstroProjectedSensorModel, restated in C)
on vectors the reference tree itself holds, and checks the product's host-side PROJ-string
parser against it.  CPU only."""
import ctypes as C
import json
import os

import numpy as np
import pytest

from oracle import geo

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.skipif(not geo.available(), reason="oracle/_ref/libgeo_oracle.so not built")

GRS80 = (6378137.0, 6356752.314140347)   # PROJ/src/ellps.cpp: a, rf = 298.257222101


def test_cart_vectors_of_proj():
    """PROJ/test/gie/more_builtins.gie:828-882 (+proj=cart +ellps=GRS80, tolerance 0.001 mm)."""
    g = geo.Geo([0, 1, 0, 0, 0, 1], geo.PROJ_LONGLAT, proj_a=GRS80[0], proj_b=GRS80[1])
    fwd = [((0, 0, 0), (6378137, 0, 0)), ((0, 90, 0), (0, 0, 6356752.314140347)),
           ((0, -90, 0), (0, 0, -6356752.314140347)), ((90, 0, 0), (0, 6378137, 0)),
           ((-90, 0, 0), (0, -6378137, 0)), ((180, 0, 0), (-6378137, 0, 0)),
           ((-180, 0, 0), (-6378137, 0, 0)), ((0, 0, -6378137), (0, 0, 0)),
           ((0, 90, -6356752.314140347), (0, 0, 0))]
    for (lon, lat, h), exp in fwd:
        assert np.allclose(g.cart_forward_deg(lon, lat, h), exp, rtol=0, atol=1e-6)
    inv = [((6378137, 0, 0), (0, 0, 0)), ((0, 0, 6356752.314140347), (0, 90, 0)),
           ((0, 0, -6356752.314140347), (0, -90, 0)), ((0, 6378137, 0), (90, 0, 0)),
           ((0, -6378137, 0), (-90, 0, 0)), ((-6378137, 0, 0), (180, 0, 0)),
           ((0, 0, 0), (0, 90, -6356752.314140356)), ((0, 0, 1e-6), (0, 90, -6356752.314139356)),
           ((0, 0, -1e-6), (0, -90, -6356752.314139356))]
    for xyz, (lon, lat, h) in inv:
        o = g.cart_inverse_deg(*xyz)
        # gie compares geographic results as a distance on the ellipsoid: 1e-6 mm ~ 1e-14 deg
        assert abs(o[0] - lon) < 1e-11 and abs(o[1] - lat) < 1e-11 and abs(o[2] - h) < 1e-6


def test_eqc_vectors_of_proj():
    """PROJ/test/gie/builtins.gie:1359-1379 (+proj=eqc +a=6400000, tolerance 0.1 mm)."""
    g = geo.Geo([0, 1, 0, 0, 0, 1], geo.PROJ_EQC, proj_a=6400000.0)
    for sx in (1, -1):
        for sy in (1, -1):
            assert np.allclose(g.proj_forward_deg(2 * sx, 1 * sy),
                               (223402.144255274 * sx, 111701.072127637 * sy), rtol=0, atol=1e-4)
            assert np.allclose(g.proj_inverse_deg(200 * sx, 100 * sy),
                               (0.001790493 * sx, 0.000895247 * sy), rtol=0, atol=5e-10)


def test_projected_model_golden_values_of_the_reference():
    """tests/ProjectedCameraTests.cpp:32-58 on the fixture tests/data/constVelocityProj.json: the
    projected model's imageToGround is the geotransform + PROJ chain alone (the camera is not
    involved), ellipsoid of the cartesian step = the ISD radii (10 m)."""
    isd = json.load(open(os.path.join(ROOT, "tests", "golden", "ref_data", "constVelocityProj.json")))
    assert isd["projection"].startswith("+proj=sinu +lon_0=355.88485589576")
    r = isd["radii"]["semimajor"] * 1000.0
    g = geo.Geo(isd["geotransform"], geo.PROJ_SINU, lon_0=355.88485589576, proj_a=3396190.0,
                cart_a=r, cart_b=isd["radii"]["semiminor"] * 1000.0)
    x, y, z = g.image_to_ground([8.0, 0.0], [8.0, 4.0])
    assert abs(x[0] - 9.8989858043564229) < 1e-9 and abs(y[0] + 1.4118584049198812) < 1e-9
    assert abs(z[0] - 0.12936726636640034) < 1e-9
    assert abs(x[1] - 9.8988664296556497) < 1e-9 and abs(y[1] + 1.4125636827694117) < 1e-9
    assert abs(z[1] - 0.1307946862733107) < 1e-9
    # Inversion (ProjectedCameraTests.cpp:40-50): back to the pixel within 1e-3
    l, s, h = g.ground_to_image(x, y, z)
    assert np.allclose(l, [8.0, 0.0], atol=1e-7) and np.allclose(s, [8.0, 4.0], atol=1e-7)
    assert np.abs(h).max() < 1e-8


def test_pixel_to_meter_kat():
    """tests/UtilitiesTests.cpp:656-668."""
    gt = [-104607.78948592, 4.980137561815, 0.0, -570564.40018202, 0.0, -4.980137561815]
    g = geo.Geo(gt, geo.PROJ_EQC, proj_a=3396190.0)
    # pixelToMeter(13.6191, 4981.58) = (-570632.225173488, -79798.83581073358)
    my = 13.6191 * gt[5] + gt[3]
    mx = 4981.58 * gt[1] + gt[0]
    assert abs(my + 570632.225173488) < 1e-4 and abs(mx + 79798.83581073358) < 1e-4
    # the same pixel through the whole chain and back
    l, s, _ = g.ground_to_image(*g.image_to_ground([13.6191], [4981.58]))
    assert abs(l[0] - 13.6191) < 1e-4 and abs(s[0] - 4981.58) < 1e-4


def test_sinu_ellipsoid_round_trip():
    """tests/data/marsLineScanProj.json: +proj=sinu +a=3396190 +b=3376200 (the meridional-distance
    series, PROJ/src/mlfn.cpp): forward o inverse is the identity to the accuracy of PROJ's
    closed-form geodetic conversion (Bowring, cart.cpp:146-199: a few micrometres at +-5 km)."""
    isd = json.load(open(os.path.join(ROOT, "tests", "golden", "ref_data", "marsLineScanProj.json")))
    assert "+a=3396190.0 +b=3376200.0" in isd["projection"]
    g = geo.Geo(isd["geotransform"], geo.PROJ_SINU, proj_a=3396190.0, proj_b=3376200.0)
    rng = np.random.default_rng(1)
    line, samp = rng.uniform(-3000, 3000, 2000), rng.uniform(-3000, 3000, 2000)
    h = rng.uniform(-5000, 5000, 2000)
    x, y, z = g.image_to_ground(line, samp, h)
    l2, s2, h2 = g.ground_to_image(x, y, z)
    assert np.abs(l2 - line).max() < 1e-8 and np.abs(s2 - samp).max() < 1e-8   # 1000 m pixels
    assert np.abs(h2 - h).max() < 1e-5
    # a point on the equator / central meridian sits at (a, 0, 0)
    px = (0.0 - isd["geotransform"][3]) / isd["geotransform"][5]
    sx = (0.0 - isd["geotransform"][0]) / isd["geotransform"][1]
    x0, y0, z0 = g.image_to_ground([px], [sx])
    assert abs(x0[0] - 3396190.0) < 1e-6 and abs(y0[0]) < 1e-6 and abs(z0[0]) < 1e-6


def test_product_proj_string_parser():
    """usgscsm_b200_geogrid_from_proj_string (host only): the fixture strings parse to the
    parameters the oracle is set up with; unsupported projections are refused."""
    from usgscsm_b200 import _capi
    import usgscsm_b200 as ub
    gt = [-237682.97649184, 60.60249273122, 0.0, 44421.627171984, 0.0, -60.60249273122]
    g = ub.GeoGrid.fromProjString("+proj=sinu +lon_0=355.88485589576 +x_0=0 +y_0=0 +R=3396190 +units=m +no_defs",
                                  gt, 10.0, 10.0).c
    assert g.projection == _capi.PROJ_SINU and g.lon_0_deg == 355.88485589576
    assert g.proj_semi_major == 3396190.0 and g.proj_semi_minor == 3396190.0
    assert g.semi_major == 10.0 and list(g.geotransform) == gt
    g = ub.GeoGrid.fromProjString("+proj=sinu +lon_0=0.0 +x_0=0 +y_0=0 +a=3396190.0 +b=3376200.0 +units=m +no_defs",
                                  gt).c
    assert (g.proj_semi_major, g.proj_semi_minor) == (3396190.0, 3376200.0)
    g = ub.GeoGrid.fromProjString("+proj=eqc +lat_ts=30 +lat_0=0 +lon_0=180 +x_0=5 +y_0=-7 +R=1737400 +units=m",
                                  gt).c
    assert g.projection == _capi.PROJ_EQC and (g.lat_ts_deg, g.lon_0_deg, g.x_0, g.y_0) == (30.0, 180.0, 5.0, -7.0)
    g = ub.GeoGrid.fromProjString("+proj=longlat +a=1737400 +b=1737400 +no_defs", gt).c
    assert g.projection == _capi.PROJ_LONGLAT
    for bad in ("+proj=stere +lat_0=90 +R=1737400", "+proj=eqc +R=1737400 +units=km",
                "+proj=eqc +lat_ts=0", "+proj=sinu +R=3396190 +pm=paris"):
        with pytest.raises(_capi.EngineError) as e:
            ub.GeoGrid.fromProjString(bad, gt)
        assert e.value.code == _capi.EUNSUPPORTED
