"""Georeferenced-grid entry points on the GPU against the CPU checkers:
oracle/geo_restate.c (PROJ's arithmetic, pinned on the reference's own golden values in
tests/test_oracle_geo.py) for the coordinate chain, and the REFERENCE sensor model (oracle/_ref)
for groundToImage / imageToGround around it.

  - DEM window -> ground points: bit-exact for longlat / eqc grids (host-built trigonometry
    tables, + and * in the kernel), <= 1e-6 m for sinu (CUDA sin / cos per pixel);
  - raster pixel <-> body-fixed (UsgsAstroProjectedSensorModel imageToGround / groundToImage):
    <= 1e-4 m, <= 1e-6 px;
  - groundToImage over a DEM window, points generated in the kernel, all four models: the
    reference's per-point groundToImage on the oracle's ground points, <= 1e-6 px, status bits
    bit-exact, nodata samples flagged;
  - DEM-iterated intersection: the same fixed point written as a loop around the reference's
    imageToGround, <= 1e-4 m.
"""
import json
import os

import numpy as np
import pytest

from oracle import geo
from parity_util import PX_TOL, M_TOL, models

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
MOON = 1737400.0


def _grids():
    """(label, proj string, oracle kwargs, geotransform) of grids near lon 10, lat 5 deg."""
    deg = 1.0 / 4096.0
    m = MOON * np.pi / 180.0 * deg
    return {
        "longlat": ("+proj=longlat +a=1737400 +b=1737400 +no_defs", dict(projection=geo.PROJ_LONGLAT, proj_a=MOON),
                    [9.5, deg, 0.0, 5.5, 0.0, -deg]),
        "longlat_ellipsoid": ("+proj=longlat +a=3396190 +b=3376200 +no_defs",
                              dict(projection=geo.PROJ_LONGLAT, proj_a=3396190.0, proj_b=3376200.0),
                              [9.5, deg, 0.0, 5.5, 0.0, -deg]),
        "eqc": ("+proj=eqc +lat_ts=20 +lat_0=0 +lon_0=8 +x_0=100 +y_0=-50 +R=1737400 +units=m",
                dict(projection=geo.PROJ_EQC, lat_ts=20.0, lon_0=8.0, x_0=100.0, y_0=-50.0, proj_a=MOON),
                [40000.0, m, 0.0, 170000.0, 0.0, -m]),
        "sinu": ("+proj=sinu +lon_0=8 +x_0=0 +y_0=0 +R=1737400 +units=m +no_defs",
                 dict(projection=geo.PROJ_SINU, lon_0=8.0, proj_a=MOON), [40000.0, m, 0.0, 170000.0, 0.0, -m]),
        "sinu_ellipsoid": ("+proj=sinu +lon_0=8 +a=3396190 +b=3376200 +units=m",
                           dict(projection=geo.PROJ_SINU, lon_0=8.0, proj_a=3396190.0, proj_b=3376200.0),
                           [80000.0, 2 * m, 0.0, 330000.0, 0.0, -2 * m]),
    }


def _dem_samples(rng, nr, nc, dtype):
    r, c = np.meshgrid(np.arange(nr), np.arange(nc), indexing="ij")
    h = 400.0 * np.sin(r / 37.0) * np.cos(c / 23.0) + rng.uniform(-5, 5, (nr, nc))
    if dtype == np.int16:
        return np.round(h * 2).astype(np.int16), 0.5, -3.0   # raw, scale, offset
    return h.astype(dtype), 1.0, 0.0


@pytest.mark.parametrize("grid", ["longlat", "longlat_ellipsoid", "eqc", "sinu", "sinu_ellipsoid"])
@pytest.mark.parametrize("dtype", [np.float64, np.float32, np.int16])
def test_dem_ground_points(grid, dtype):
    import usgscsm_b200 as ub
    proj, okw, gt = _grids()[grid]
    rng = np.random.default_rng(11)
    nr, nc = 301, 517
    raw, scale, offset = _dem_samples(rng, nr, nc, dtype)
    raw[5, 7] = raw[200, 300] = -32768 if dtype == np.int16 else -1e30
    nodata = float(raw[5, 7])
    dem = ub.Dem(raw, ub.GeoGrid.fromProjString(proj, gt), nodata=nodata, scale=scale, offset=offset)
    og = geo.Geo(gt, **okw)
    for (r0, c0, wr, wc) in ((0, 0, nr, nc), (3, 5, 250, 400), (100, 200, 1, 1)):
        x, y, z, st = dem.groundPoints(r0, c0, wr, wc)
        rr, cc = np.meshgrid(np.arange(r0, r0 + wr), np.arange(c0, c0 + wc), indexing="ij")
        h = raw[r0:r0 + wr, c0:c0 + wc].astype(np.float64) * scale + offset
        ex, ey, ez = og.image_to_ground(rr + 0.5, cc + 0.5, h)
        hole = (raw[r0:r0 + wr, c0:c0 + wc] == raw[5, 7]).reshape(-1)
        assert np.array_equal((st & ub._capi.ST_DEM_NODATA) != 0, hole)
        assert np.all(np.isnan(x[hole])) and np.all(np.isfinite(x[~hole]))
        ok = ~hole
        if grid.startswith("sinu"):
            assert max(np.abs(x[ok] - ex[ok]).max(), np.abs(y[ok] - ey[ok]).max(), np.abs(z[ok] - ez[ok]).max()) <= 1e-6
        else:   # bit for bit
            assert np.array_equal(x[ok], ex[ok]) and np.array_equal(y[ok], ey[ok]) and np.array_equal(z[ok], ez[ok])


def test_dem_ground_points_device_and_chunked_host():
    """The device-pointer flavour and a host raster cut into several pipeline chunks (2100 x 2100 >
    the 4 Mi-point chunk) give the same bits as the one-chunk host call."""
    import torch
    import usgscsm_b200 as ub
    proj, okw, gt = _grids()["eqc"]
    rng = np.random.default_rng(12)
    raw = rng.uniform(-500, 500, (2100, 2100)).astype(np.float32)
    grid = ub.GeoGrid.fromProjString(proj, gt)
    x, y, z, st = ub.Dem(raw, grid).groundPoints(10, 20, 2080, 2070)
    og = geo.Geo(gt, **okw)
    rr, cc = np.meshgrid(np.arange(10, 2090), np.arange(20, 2090), indexing="ij")
    ex, ey, ez = og.image_to_ground(rr + 0.5, cc + 0.5, raw[10:2090, 20:2090].astype(np.float64))
    assert np.array_equal(x, ex) and np.array_equal(y, ey) and np.array_equal(z, ez) and not st.any()
    dx, dy, dz, dst = ub.Dem(torch.as_tensor(raw, device="cuda:0"), grid).groundPoints(10, 20, 2080, 2070)
    assert np.array_equal(dx.cpu().numpy(), x) and np.array_equal(dz.cpu().numpy(), z)


@pytest.mark.parametrize("grid", ["longlat_ellipsoid", "eqc", "sinu", "sinu_ellipsoid"])
def test_projected_model_pair(grid):
    """UsgsAstroProjectedSensorModel::imageToGround / groundToImage for arrays."""
    import usgscsm_b200 as ub
    proj, okw, gt = _grids()[grid]
    g, og = ub.GeoGrid.fromProjString(proj, gt), geo.Geo(gt, **okw)
    rng = np.random.default_rng(13)
    line, samp = rng.uniform(-2000, 6000, 5000), rng.uniform(-2000, 6000, 5000)
    h = rng.uniform(-3000, 3000, 5000)
    x, y, z, st = g.imageToGround(line, samp, h)
    ex, ey, ez = og.image_to_ground(line, samp, h)
    assert not st.any()
    assert max(np.abs(x - ex).max(), np.abs(y - ey).max(), np.abs(z - ez).max()) <= M_TOL * 1e-2
    l, s, hh, st2 = g.groundToImage(ex, ey, ez)
    el, es, eh = og.ground_to_image(ex, ey, ez)
    assert not st2.any()
    assert max(np.abs(l - el).max(), np.abs(s - es).max()) <= PX_TOL and np.abs(hh - eh).max() <= M_TOL
    # height omitted = 0, as the reference's projected model effectively does
    x0, _, _, _ = g.imageToGround(line[:100], samp[:100])
    assert np.abs(x0 - og.image_to_ground(line[:100], samp[:100])[0]).max() <= M_TOL * 1e-2


def test_reference_projected_fixture_through_the_engine():
    """tests/ProjectedCameraTests.cpp:32-58 (golden values of the reference) on the GPU."""
    import usgscsm_b200 as ub
    isd = json.load(open(os.path.join(ROOT, "tests", "golden", "ref_data", "constVelocityProj.json")))
    g = ub.GeoGrid.fromProjString(isd["projection"], isd["geotransform"],
                                  isd["radii"]["semimajor"] * 1000.0, isd["radii"]["semiminor"] * 1000.0)
    x, y, z, st = g.imageToGround([8.0, 0.0], [8.0, 4.0])
    assert abs(x[0] - 9.8989858043564229) < 1e-9 and abs(y[0] + 1.4118584049198812) < 1e-9
    assert abs(z[0] - 0.12936726636640034) < 1e-9
    assert abs(x[1] - 9.8988664296556497) < 1e-9 and abs(y[1] + 1.4125636827694117) < 1e-9
    assert abs(z[1] - 0.1307946862733107) < 1e-9
    l, s, h, _ = g.groundToImage(x, y, z)   # Inversion, :40-50
    assert np.allclose(l, [8.0, 0.0], atol=1e-3) and np.allclose(s, [8.0, 4.0], atol=1e-3)


def _footprint_dem(rm, og_kwargs_proj, nr, nc, rng, margin=0.15):
    """A longlat DEM raster covering the image footprint of reference model rm."""
    nl, ns = rm.image_size()
    cl = np.array([0.0, 0.0, nl, nl, nl / 2.0]); cs = np.array([0.0, ns, 0.0, ns, ns / 2.0])
    x, y, z, _, _ = rm.image_to_ground(cl, cs, np.zeros(5))
    lon, lat = np.degrees(np.arctan2(y, x)), np.degrees(np.arcsin(z / np.sqrt(x * x + y * y + z * z)))
    lo0, lo1, la0, la1 = lon.min(), lon.max(), lat.min(), lat.max()
    dlo, dla = (lo1 - lo0) * margin, (la1 - la0) * margin
    gt = [lo0 - dlo, (lo1 - lo0 + 2 * dlo) / nc, 0.0, la1 + dla, 0.0, -(la1 - la0 + 2 * dla) / nr]
    return gt


def _model_cases():
    from usgscsm_b200 import synth
    return {
        "linescan": synth.linescan_state(n_lines=4096, n_samples=1024, seed=5),
        "linescan_adjusted": synth.linescan_state(n_lines=2048, n_samples=512, seed=6),
        "pushframe": synth.pushframe_state(n_framelets=64, framelet_height=16, n_samples=1024),
        "frame": synth.frame_state(n_lines=1024, n_samples=1024),
        "sar": synth.sar_state(),
    }


@pytest.mark.parametrize("kind", ["linescan", "linescan_adjusted", "pushframe", "frame", "sar"])
@pytest.mark.parametrize("dtype", [np.float64, np.int16])
def test_ground_to_image_dem_grid(ref, kind, dtype):
    import torch
    import usgscsm_b200 as ub
    from usgscsm_b200 import _capi
    state = _model_cases()[kind]
    gm, rm = models(ref, state)
    if kind == "linescan_adjusted":
        for i, v in {0: 3.0, 4: -0.01, 7: 2e-5, 15: 0.02}.items():
            gm.setParameterValue(i, v)
            rm.set_parameter(i, v)
    if kind.startswith("linescan"):
        rm.ls_set_projective(True, gm.lsDerived()[5:19])
    a, b = gm.getEllipsoid()
    rng = np.random.default_rng(21)
    nr, nc = 257, 311
    gt = _footprint_dem(rm, None, nr, nc, rng)
    raw, scale, offset = _dem_samples(rng, nr, nc, dtype)
    hole_value = -32768 if dtype == np.int16 else -9999.0
    raw[::41, ::29] = hole_value
    proj = "+proj=longlat +a=%.17g +b=%.17g" % (a, b)
    dem = ub.Dem(raw, ub.GeoGrid.fromProjString(proj, gt), nodata=hole_value, scale=scale, offset=offset)
    og = geo.Geo(gt, projection=geo.PROJ_LONGLAT, proj_a=a, proj_b=b)
    for (r0, c0, wr, wc) in ((0, 0, nr, nc), (17, 9, 200, 256)):
        line, samp, ach, st = gm.groundToImageDemGrid(dem, r0, c0, wr, wc, achieved=True)
        rr, cc = np.meshgrid(np.arange(r0, r0 + wr), np.arange(c0, c0 + wc), indexing="ij")
        win = raw[r0:r0 + wr, c0:c0 + wc]
        ex, ey, ez = og.image_to_ground(rr + 0.5, cc + 0.5, win.astype(np.float64) * scale + offset)
        hole = (win == hole_value).reshape(-1)
        assert np.array_equal((st & _capi.ST_DEM_NODATA) != 0, hole)
        assert np.all(np.isnan(line[hole])) and np.all(np.isnan(samp[hole])) and np.all(np.isnan(ach[hole]))
        assert np.all(st[hole] == (_capi.ST_DEM_NODATA | _capi.ST_NONFINITE))
        ok = ~hole
        rl, rs, ra, rst = rm.ground_to_image(ex[ok], ey[ok], ez[ok])
        fin = np.isfinite(rl) & np.isfinite(rs)
        assert np.array_equal(np.isfinite(line[ok]) & np.isfinite(samp[ok]), fin)
        assert np.abs(line[ok][fin] - rl[fin]).max(initial=0) <= PX_TOL
        assert np.abs(samp[ok][fin] - rs[fin]).max(initial=0) <= PX_TOL
        fa = np.isfinite(ra)
        assert np.abs(ach[ok][fa] - ra[fa]).max(initial=0) <= PX_TOL
        assert np.array_equal(st[ok] & _capi.ST_REFERENCE_MASK, rst)
        # most of the window must actually be inside the image
        assert ((st[ok] & _capi.ST_OUT_OF_IMAGE) == 0).mean() > 0.05
    # device flavour: same bits as the host flavour
    tdem = ub.Dem(torch.as_tensor(raw, device="cuda:0"), dem.grid, nodata=hole_value, scale=scale, offset=offset)
    dl, ds, da, dst = gm.groundToImageDemGrid(tdem, 17, 9, 200, 256, achieved=True)
    assert np.array_equal(dl.cpu().numpy(), line, equal_nan=True) and np.array_equal(ds.cpu().numpy(), samp, equal_nan=True)
    assert np.array_equal(dst.cpu().numpy().view(np.uint32), st)
    # the fused path equals the two-step path (ground points, then groundToImage) bit for bit
    x, y, z, _ = dem.groundPoints(17, 9, 200, 256)
    l2, s2, _, st2 = gm.groundToImageBatch(x[ok], y[ok], z[ok])
    assert np.array_equal(l2, line[ok], equal_nan=True) and np.array_equal(s2, samp[ok], equal_nan=True)
    assert np.array_equal(st2, st[ok])


@pytest.mark.parametrize("kind", ["linescan", "pushframe", "frame", "sar"])
def test_image_to_ground_dem(ref, kind):
    """DEM-iterated intersection against the same loop around the reference's imageToGround."""
    import torch
    import usgscsm_b200 as ub
    from usgscsm_b200 import _capi
    state = _model_cases()[kind]
    gm, rm = models(ref, state)
    a, b = gm.getEllipsoid()
    rng = np.random.default_rng(31)
    nr, nc = 181, 203
    gt = _footprint_dem(rm, None, nr, nc, rng, margin=-0.1)   # smaller than the footprint: rays leave it
    r, c = np.meshgrid(np.arange(nr), np.arange(nc), indexing="ij")
    heights = (300.0 * np.sin(r / 17.0) * np.cos(c / 13.0)).astype(np.float32)
    heights[60:63, 80:84] = -9999.0
    dem = ub.Dem(heights, ub.GeoGrid.fromProjString("+proj=longlat +a=%.17g +b=%.17g" % (a, b), gt), nodata=-9999.0)
    og = geo.Geo(gt, projection=geo.PROJ_LONGLAT, proj_a=a, proj_b=b)
    nl, ns = rm.image_size()
    n = 4000
    line, samp = rng.uniform(0, nl, n), rng.uniform(0, ns, n)
    tol, iters, h0 = 1e-5, 12, 20.0
    x, y, z, h, st = gm.imageToGroundDem(dem, line, samp, h0=h0, tolerance=tol, maxIterations=iters)

    # the oracle: the same fixed point, one reference call per iteration
    hd = np.where(heights == np.float32(-9999.0), np.nan, heights.astype(np.float64))
    eh = np.full(n, h0)
    done = np.zeros(n, bool)
    est = np.zeros(n, np.uint32)
    ex = ey = ez = None
    for it in range(iters):
        rx, ry, rz, _, rst = rm.image_to_ground(line, samp, eh)
        if ex is None:
            ex, ey, ez = rx.copy(), ry.copy(), rz.copy()
        act = ~done
        ex[act], ey[act], ez[act] = rx[act], ry[act], rz[act]
        est[act] = rst[act]
        bad = act & ((rst & (_capi.ST_ERROR_ALGORITHM | _capi.ST_ERROR_UNKNOWN | _capi.ST_ERROR_OTHER | _capi.ST_NONFINITE)) != 0)
        done |= bad
        act = ~done
        l, s, _ = og.ground_to_image(rx, ry, rz)
        hn = geo.bilinear(hd, l, s)
        u, v = s - 0.5, l - 0.5
        outside = ~((u >= 0) & (v >= 0) & (u <= nc - 1) & (v <= nr - 1))
        est[act & outside] |= _capi.ST_DEM_OUTSIDE
        hole = act & ~outside & np.isnan(hn)
        est[hole] |= _capi.ST_DEM_NODATA
        done |= act & (outside | hole)
        act = ~done
        conv = act & (np.abs(hn - eh) <= tol)
        done |= conv
        act = ~done
        if it == iters - 1:
            est[act] |= _capi.ST_ITERATION_CAP
        else:
            eh[act] = hn[act]
    fin = np.isfinite(ex)
    assert np.array_equal(np.isfinite(x), fin)
    assert max(np.abs(x[fin] - ex[fin]).max(), np.abs(y[fin] - ey[fin]).max(), np.abs(z[fin] - ez[fin]).max()) <= M_TOL
    assert np.abs(h[fin] - eh[fin]).max() <= M_TOL
    geo_bits = _capi.ST_DEM_OUTSIDE | _capi.ST_DEM_NODATA | _capi.ST_ITERATION_CAP | _capi.ST_REFERENCE_MASK
    assert np.array_equal(st & geo_bits, est & geo_bits)
    # the test must exercise convergence, leaving the raster and the hole
    inside = (st & (_capi.ST_DEM_OUTSIDE | _capi.ST_DEM_NODATA)) == 0
    assert inside.mean() > 0.2 and (st & _capi.ST_DEM_OUTSIDE).any()
    # converged points sit on the DEM surface: height above the ellipsoid == DEM height there
    l, s, hh = og.ground_to_image(x[inside], y[inside], z[inside])
    on = geo.bilinear(hd, l, s)
    conv = inside.copy(); conv[inside] = ((st[inside] & _capi.ST_ITERATION_CAP) == 0) & np.isfinite(on)
    assert np.abs(hh[conv[inside]] - on[conv[inside]]).max() < 1e-3
    # device flavour agrees with the host flavour
    tdem = ub.Dem(torch.as_tensor(heights, device="cuda:0"), dem.grid, nodata=-9999.0)
    dx, dy, dz, dh, dst = gm.imageToGroundDem(tdem, torch.as_tensor(line, device="cuda:0"),
                                              torch.as_tensor(samp, device="cuda:0"), h0=h0, tolerance=tol,
                                              maxIterations=iters)
    assert np.array_equal(dx.cpu().numpy(), x, equal_nan=True) and np.array_equal(dh.cpu().numpy(), h, equal_nan=True)
    assert np.array_equal(dst.cpu().numpy().view(np.uint32), st)
