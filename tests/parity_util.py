"""Shared helpers of the GPU parity tests: engine model + reference model on the same state,
and the comparison at the bar of BASELINE.json (<= 1e-6 px, <= 1e-4 m, observable status bits
bit-exact).  Everything on the engine side goes through the C-ABI."""
import numpy as np

PX_TOL = 1e-6   # pixels
M_TOL = 1e-4    # metres


def models(ref, state, device=0):
    import usgscsm_b200 as ub
    gm = ub.UsgsAstroPlugin(device=device).constructModelFromState(state)
    rm = ref.RefModel(state)
    return gm, rm


def check_i2g(gm, rm, line, samp, h, prec=0.001, tol=M_TOL):
    from usgscsm_b200 import _capi
    x, y, z, a, st = gm.imageToGroundBatch(line, samp, h, prec)
    rx, ry, rz, ra, rst = rm.image_to_ground(line, samp, h, prec)
    ok = np.isfinite(rx) & np.isfinite(ry) & np.isfinite(rz)
    assert np.array_equal(np.isfinite(x) & np.isfinite(y) & np.isfinite(z), ok)
    for u, v in ((x, rx), (y, ry), (z, rz)):
        assert np.abs(u[ok] - v[ok]).max(initial=0.0) <= tol
    fa = np.isfinite(ra)
    assert np.array_equal(np.isfinite(a), fa)
    assert np.abs(a[fa] - ra[fa]).max(initial=0.0) <= tol
    assert np.array_equal(st & _capi.ST_REFERENCE_MASK, rst)
    return x, y, z


def check_g2i(gm, rm, x, y, z, prec=0.001, tol=PX_TOL, achieved=True):
    from usgscsm_b200 import _capi
    l, s, a, st = gm.groundToImageBatch(x, y, z, prec)
    rl, rs, ra, rst = rm.ground_to_image(x, y, z, prec)
    ok = np.isfinite(rl) & np.isfinite(rs)
    assert np.array_equal(np.isfinite(l) & np.isfinite(s), ok)
    assert np.abs(l[ok] - rl[ok]).max(initial=0.0) <= tol
    assert np.abs(s[ok] - rs[ok]).max(initial=0.0) <= tol
    if achieved:
        fa = np.isfinite(ra)
        assert np.array_equal(np.isfinite(a), fa)
        assert np.abs(a[fa] - ra[fa]).max(initial=0.0) <= tol
    assert np.array_equal(st & _capi.ST_REFERENCE_MASK, rst)
    return l, s


def pixels(rng, n, nl, ns):
    return rng.uniform(0, nl, n), rng.uniform(0, ns, n)


def nasty_ground_points(rng, x, y, z, radius):
    """Ground points that exercise the failure paths: scaled, mirrored (far side), random on
    a sphere, the origin, NaN."""
    pts = np.stack([x[:600], y[:600], z[:600]]).copy()
    pts[:, 100:200] *= 1.01
    pts[:, 200:300] *= -1.0
    pts[:, 300:400] = rng.normal(size=(3, 100)) * radius
    pts[:, 400:410] = 0.0
    pts[:, 410:420] = np.nan
    return pts[0], pts[1], pts[2]
