"""Full-size parity sweep (not a pytest: it takes a minute of host CPU): every pixel of the C2
line-scan image and large samples of the other models go image -> ground -> image through the
engine on cuda:0 AND through the reference (oracle/_ref, all host threads); results are compared
bit for bit and the counts are printed.  Run on the GPU box:
    python profiles/parity_sweep.py > gpurun_out/r2_parity_sweep.txt
"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import usgscsm_b200 as ub  # noqa: E402
from oracle import ref  # noqa: E402
from usgscsm_b200 import _capi, synth  # noqa: E402

THREADS = len(os.sched_getaffinity(0))
PREC = 1e-3


def sweep(name, state, max_points, adjust=None):
    gm = ub.UsgsAstroPlugin().constructModelFromState(state)
    rm = ref.RefModel(state)
    if adjust:
        for i, v in adjust.items():
            gm.setParameterValue(i, v)
            rm.set_parameter(i, v)
    if gm.getModelName() == "USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL":
        rm.ls_set_projective(True, gm.lsDerived()[5:19])
    sz = gm.getImageSize()
    nl, ns = int(sz.line), int(sz.samp)
    step = max(1, int(np.ceil(nl * ns / max_points)))
    rows = np.arange(0, nl, step, dtype=np.float64) + 0.5
    line = np.repeat(rows, ns)
    samp = np.tile(np.arange(ns, dtype=np.float64) + 0.5, rows.size)
    h = synth.synthetic_dem(line, samp)
    n = line.size
    t = time.perf_counter()
    gx, gy, gz, ga, gst = gm.imageToGroundBatch(line, samp, h, PREC)
    gl, gs, ga2, gst2 = gm.groundToImageBatch(gx, gy, gz, PREC)
    t_gpu = time.perf_counter() - t
    t = time.perf_counter()
    ex, ey, ez, ea, est = rm.image_to_ground(line, samp, h, PREC, threads=THREADS)
    el, es, ea2, est2 = rm.ground_to_image(gx, gy, gz, PREC, threads=THREADS)
    t_cpu = time.perf_counter() - t

    def same(a, b):
        return int(np.sum((a.view(np.uint64) != b.view(np.uint64)) & ~(np.isnan(a) & np.isnan(b))))
    bad_i2g = same(gx, ex) + same(gy, ey) + same(gz, ez)
    bad_g2i = same(gl, el) + same(gs, es)
    bad_st = int(np.sum((gst & _capi.ST_REFERENCE_MASK) != est)) + int(np.sum((gst2 & _capi.ST_REFERENCE_MASK) != est2))
    worst_m = float(np.nanmax(np.abs(np.stack([gx - ex, gy - ey, gz - ez])))) if n else 0.0
    worst_px = float(np.nanmax(np.abs(np.stack([gl - el, gs - es])))) if n else 0.0
    flagged = int(np.sum((est & ~np.uint32(_capi.ST_OUT_OF_IMAGE)) != 0)) + int(np.sum((est2 & ~np.uint32(_capi.ST_OUT_OF_IMAGE)) != 0))
    print("%-34s %11d points  i2g coords differing in any bit: %d (max %.3g m)  g2i: %d (max %.3g px)  "
          "status words differing: %d  points the reference flags: %d  [engine %.2f s host arrays, reference %.1f s on %d threads]"
          % (name, n, bad_i2g, worst_m, bad_g2i, worst_px, bad_st, flagged, t_gpu, t_cpu, THREADS), flush=True)
    return bad_i2g + bad_g2i + bad_st


def hostile(name, state, n, adjust=None):
    """The failure paths at scale: pixels up to half an image outside the image with heights of
    +-20 km, and ground points displaced by up to several image footprints, scaled off the body,
    mirrored to the far side, at the origin, NaN and infinite -- results AND status words against
    the reference's returns / exceptions / warnings."""
    rng = np.random.default_rng(99)
    gm = ub.UsgsAstroPlugin().constructModelFromState(state)
    rm = ref.RefModel(state)
    if adjust:
        for i, v in adjust.items():
            gm.setParameterValue(i, v)
            rm.set_parameter(i, v)
    if gm.getModelName() == "USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL":
        rm.ls_set_projective(True, gm.lsDerived()[5:19])
    sz = gm.getImageSize()
    nl, ns = float(sz.line), float(sz.samp)
    line = rng.uniform(-0.5 * nl, 1.5 * nl, n)
    samp = rng.uniform(-0.5 * ns, 1.5 * ns, n)
    h = rng.uniform(-20000.0, 20000.0, n)
    gx, gy, gz, ga, gst = gm.imageToGroundBatch(line, samp, h, PREC)
    ex, ey, ez, ea, est = rm.image_to_ground(line, samp, h, PREC, threads=THREADS)
    # ground points: the intersections, displaced
    x, y, z = ex.copy(), ey.copy(), ez.copy()
    bad = ~np.isfinite(x + y + z)
    x[bad], y[bad], z[bad] = 1.0e6, 1.0e6, 1.0e6
    q = n // 8
    ext = float(np.nanmax(np.abs(np.stack([x, y, z])))) * 0.05
    for a in (x, y, z):
        a[:q] += rng.normal(size=q) * ext                     # off by footprints
    x[q:2 * q] *= 1.2; y[q:2 * q] *= 1.2; z[q:2 * q] *= 1.2   # off the body
    x[2 * q:3 * q] *= -1; y[2 * q:3 * q] *= -1; z[2 * q:3 * q] *= -1   # far side
    x[3 * q:3 * q + 50] = 0; y[3 * q:3 * q + 50] = 0; z[3 * q:3 * q + 50] = 0
    x[3 * q + 50:3 * q + 100] = np.nan
    y[3 * q + 100:3 * q + 150] = np.inf
    gl, gs, ga2, gst2 = gm.groundToImageBatch(x, y, z, PREC)
    el, es, ea2, est2 = rm.ground_to_image(x, y, z, PREC, threads=THREADS)

    def same(a, b):
        return int(np.sum((a.view(np.uint64) != b.view(np.uint64)) & ~(np.isnan(a) & np.isnan(b))))
    bad_i2g = same(gx, ex) + same(gy, ey) + same(gz, ez)
    bad_g2i = same(gl, el) + same(gs, es)
    bad_st = int(np.sum((gst & _capi.ST_REFERENCE_MASK) != est)) + int(np.sum((gst2 & _capi.ST_REFERENCE_MASK) != est2))
    flagged = int(np.sum((est & ~np.uint32(_capi.ST_OUT_OF_IMAGE)) != 0)) + int(np.sum((est2 & ~np.uint32(_capi.ST_OUT_OF_IMAGE)) != 0))
    outside = int(np.sum((est & np.uint32(_capi.ST_OUT_OF_IMAGE)) != 0)) + int(np.sum((est2 & np.uint32(_capi.ST_OUT_OF_IMAGE)) != 0))
    print("%-34s %11d points  i2g coords differing in any bit: %d  g2i: %d  status words differing: %d  "
          "points the reference flags (throws / warns): %d, outside the image: %d"
          % (name + " HOSTILE", 2 * n, bad_i2g, bad_g2i, bad_st, flagged, outside), flush=True)
    return bad_i2g + bad_g2i + bad_st


if __name__ == "__main__":
    total = 0
    # every 3rd line of the C2 grid: 88 M points keep the host arrays of both sides under 20 GB
    total += sweep("linescan C2 5064x52224 (every 3rd line)", synth.linescan_state(), 100_000_000)
    total += sweep("linescan C2, 6 parameters adjusted", synth.linescan_state(), 40_000_000,
                   adjust={0: 2.5, 1: -1.5, 4: -0.02, 7: 1e-4, 11: 2e-6, 15: 0.03})
    total += sweep("linescan radial 4096x4096", synth.linescan_state(n_lines=4096, n_samples=4096, seed=3,
                                                                    distortion_type=synth.RADIAL, dist_coeffs=(1e-5, 2e-9, 0.0)), 1 << 40)
    total += sweep("pushframe C3 4096x4096", synth.pushframe_state(n_framelets=256, framelet_height=16, n_samples=4096), 1 << 40)
    total += sweep("frame radial C3 4096x4096", synth.frame_state(n_lines=4096, n_samples=4096), 1 << 40)
    total += sweep("frame transverse 4096x4096", synth.frame_state(
        n_lines=4096, n_samples=4096, distortion_type=synth.TRANSVERSE,
        dist_coeffs=(0.0, 1.0, 0.0, 1e-6, 2e-6, -1e-6, 1e-8, 0.0, 2e-8, 0.0,
                     0.0, 0.0, 1.0, -1e-6, 1e-6, 2e-6, 0.0, 1e-8, 0.0, -2e-8)), 1 << 40)
    total += sweep("sar C4 64576x2367 (every 16th line)", synth.sar_state(), 10_000_000)
    mars = dict(major_axis=3396190.0, minor_axis=3376200.0, gm=4.2828e13, altitude=300000.0)
    total += sweep("linescan oblate body 8192x5064", synth.linescan_state(n_lines=8192, **mars), 1 << 40)
    total += sweep("pushframe oblate, flipped+reversed", synth.pushframe_state(
        n_framelets=256, framelet_height=16, n_samples=4096, flipped=True, reversed_order=True, line_sign=1.0, **mars), 1 << 40)
    total += sweep("pushframe C3, 16 parameters adjusted", synth.pushframe_state(n_framelets=256, framelet_height=16, n_samples=4096),
                   1 << 40, adjust={0: 3.0, 2: 1.5, 4: -0.01, 6: 0.5, 9: 0.05, 12: 1e-4, 15: 0.01})
    total += hostile("linescan C2", synth.linescan_state(), 4_000_000)
    total += hostile("linescan C2 adjusted", synth.linescan_state(), 2_000_000,
                     adjust={0: 2.5, 1: -1.5, 4: -0.02, 7: 1e-4, 11: 2e-6, 15: 0.03})
    total += hostile("linescan oblate", synth.linescan_state(n_lines=8192, **mars), 2_000_000)
    total += hostile("pushframe C3", synth.pushframe_state(n_framelets=256, framelet_height=16, n_samples=4096), 4_000_000)
    total += hostile("frame radial C3", synth.frame_state(n_lines=4096, n_samples=4096), 4_000_000)
    total += hostile("sar C4", synth.sar_state(), 2_000_000)
    print("TOTAL differing values: %d" % total)
    sys.exit(1 if total else 0)
