"""The device building blocks (distortion models, Lagrange interpolation) through the C-ABI:
(1) the reference's own known-answer vectors (tests/golden/reference_kat.json), (2) random
inputs against oracle/_ref, bit for bit."""
import ctypes as C
import json
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
KAT = json.load(open(os.path.join(HERE, "golden", "reference_kat.json")))


def _close(got, want, tol):
    got, want = np.asarray(got, dtype=np.float64).reshape(-1), np.asarray(want, dtype=np.float64)
    if tol == 0:
        return bool(np.all(np.abs(got - want) <= 4 * np.spacing(np.abs(want))))
    return bool(np.all(np.abs(got - want) <= tol))


def test_kat_lagrange():
    import usgscsm_b200 as ub
    for k in KAT["lagrange"]:
        out = ub.lagrangeInterp(k["n"], k["values"], k["t0"], k["dt"], [k["time"]], k["vlen"], k["order"])
        assert _close(out, k["expect"], k["tol"]), k["src"]
    k = KAT["lagrange_error"]
    with pytest.raises(ub.Error) as e:
        ub.lagrangeInterp(k["n"], k["values"], k["t0"], k["dt"], [k["time"]], k["vlen"], k["order"])
    assert e.value.getError() == ub.Error.INDEX_OUT_OF_RANGE


def test_kat_distortion():
    import usgscsm_b200 as ub
    for k in KAT["apply_distortion"]:
        dx, dy, st = ub.applyDistortion([k["in"][0]], [k["in"][1]], k["coeffs"], k["focal"], k["type"], k["prec"])
        assert _close([dx[0], dy[0]], k["expect"], k["tol"]), k["src"]
    for k in KAT["remove_distortion"]:
        ux, uy, st = ub.removeDistortion([k["in"][0]], [k["in"][1]], k["coeffs"], k["focal"], k["type"], k["tol_arg"])
        assert _close([ux[0], uy[0]], k["expect"], k["tol"]), k["src"]
    for k in KAT["round_trip"]:
        ux, uy, _ = ub.removeDistortion([k["distorted"][0]], [k["distorted"][1]], k["coeffs"], k["focal"],
                                        k["type"], k["remove_tol"])
        dx, dy, _ = ub.applyDistortion(ux, uy, k["coeffs"], k["focal"], k["type"], k["apply_prec"])
        assert _close([dx[0], dy[0]], k["distorted"], k["tol"]), k["src"]


MODELS = {
    0: (0.03, 0.00001, 0.000004),
    1: (0.0, 1.0, 0.0, 1e-4, 2e-4, -1e-4, 1e-6, 0.0, 2e-6, 0.0,
        0.0, 0.0, 1.0, -1e-4, 1e-4, 2e-4, 0.0, 1e-6, 0.0, -2e-6),
    2: (-0.0725, -0.0009649900000000001, 0.00098441, 8.5773e-06, -3.7438e-06,
        0.0214, -0.0013796, 1.3502e-05, 2.7251e-06, -6.193800000000001e-06),
    3: (8.4e-06,),
    4: (1.81e-5,),
    5: (1e-5, -2e-8, 1e-11, 0.01, -0.02),
    6: (0.0,),
    7: (0.000031, -0.000056, 1.3e-5, -1.7e-6, 2.9e-8),
    8: (-1.741e-05,),
}


@pytest.mark.parametrize("dtype", sorted(MODELS))
def test_distortion_bit_exact_vs_reference(ref, dtype):
    """All nine distortion models, both directions, 4000 random focal-plane points each
    (including |y| > 40 mm for the LRO/KPLO bail-out and huge radii for the radial bail-out)."""
    import usgscsm_b200 as ub
    rng = np.random.default_rng(100 + dtype)
    x = np.concatenate([rng.uniform(-20, 20, 3600), rng.uniform(-60, 60, 300), rng.uniform(-1e6, 1e6, 100)])
    y = np.concatenate([rng.uniform(-20, 20, 3600), rng.uniform(-60, 60, 300), rng.uniform(-1e6, 1e6, 100)])
    c = np.ascontiguousarray(MODELS[dtype], dtype=np.float64)
    focal = 700.0
    L = ref.lib()
    for prec in (1e-6, 1e-3):
        dx, dy, st = ub.applyDistortion(x, y, c, focal, dtype, prec)
        ux, uy, st2 = ub.removeDistortion(x, y, c, focal, dtype, prec)
        a, b = C.c_double(), C.c_double()
        for i in range(x.size):
            assert L.ref_apply_distortion(x[i], y[i], C.addressof(a), C.addressof(b), c.ctypes.data, c.size,
                                          focal, dtype, prec, 1e-6) == 0
            assert (a.value == dx[i] or (np.isnan(a.value) and np.isnan(dx[i]))) and \
                   (b.value == dy[i] or (np.isnan(b.value) and np.isnan(dy[i]))), (i, x[i], y[i])
            rc = L.ref_remove_distortion(x[i], y[i], C.addressof(a), C.addressof(b), c.ctypes.data, c.size,
                                         focal, dtype, prec)
            if rc != 0:
                assert st2[i] & 32  # ST_ERROR_ALGORITHM: the reference threw
                continue
            assert (a.value == ux[i] or (np.isnan(a.value) and np.isnan(ux[i]))) and \
                   (b.value == uy[i] or (np.isnan(b.value) and np.isnan(uy[i]))), (i, x[i], y[i])


def test_lagrange_bit_exact_vs_reference(ref):
    import usgscsm_b200 as ub
    rng = np.random.default_rng(7)
    L = ref.lib()
    for n, vlen, order in ((30, 3, 8), (30, 4, 8), (30, 3, 4), (5, 4, 8), (2, 3, 8), (9, 1, 6), (12, 2, 2)):
        vals = np.ascontiguousarray(rng.normal(size=n * vlen) * 1e6)
        t0, dt = -3.7, 0.37
        times = np.concatenate([rng.uniform(t0 - 3 * dt, t0 + (n + 2) * dt, 3000),
                                t0 + dt * np.arange(-2, n + 2), [np.nan, 1e300, -1e300]])
        got = ub.lagrangeInterp(n, vals, t0, dt, times, vlen, order)
        out = np.zeros(vlen)
        for i, t in enumerate(times):
            assert L.ref_lagrange(n, vals.ctypes.data, t0, dt, float(t), vlen, order, out.ctypes.data) == 0
            assert np.array_equal(out, got[i], equal_nan=True), (n, vlen, order, t)


def test_division_with_hoisted_reciprocal_is_ieee_division():
    """The kernels' divideBy(a, b, refinedReciprocal(b)) against a / b on 2 * 10^9 operand pairs
    (all exponents, subnormals, inf/NaN, the constant divisors): zero differing bit patterns."""
    import ctypes as C
    from usgscsm_b200 import _capi
    bad = C.c_uint64(123)
    for seed in (1, 20261104):
        _capi.check(_capi.lib.usgscsm_b200_selftest_division(0, 1_000_000_000, seed, C.byref(bad)))
        assert bad.value == 0


# ---- tests/UtilitiesTests.cpp:55-284, 443-475 on the device functions ------------------------------
def test_kat_focal_plane_affine():
    """computeDistortedFocalPlaneCoordinates (UtilitiesTests.cpp:55-101) and computePixel (:238-282):
    EXPECT_DOUBLE_EQ = 4 ulp."""
    from usgscsm_b200 import plugin as P
    S, L = [0.0, 0.0, 10.0], [0.0, 10.0, 0.0]
    for args, expect, tol in (
            ((0.5, 4.0, 8.0, 0.5, 1.0, 1.0, 0.0, 0.0, S, L), (0.0, -0.4), 0),                 # :55
            ((2.0, 4.0, 8.0, 8.0, 2.0, 2.0, 0.0, 0.0, S, L), (-0.4, 0.0), 0),                 # :66 Summing
            ((11.0, -9.0, 0.0, 0.0, 1.0, 1.0, 0.0, 0.0, [-10.0, 0.0, 0.1], [10.0, -0.1, 0.0]), (-10.0, 10.0), 1e-12),  # :77
            ((2.0, 4.0, 8.0, 8.0, 1.0, 1.0, 2.0, 1.0, S, L), (-0.5, -0.2), 0)):               # :88 Start
        x, y = P.computeDistortedFocalPlaneCoordinates([args[0]], [args[1]], *args[2:])
        assert _close([x[0], y[0]], expect, tol), args
    for args, expect in (
            ((0.0, -0.4, 8.0, 0.5, 1.0, 1.0, 0.0, 0.0, S, L), (0.5, 4.0)),                    # :238
            ((-0.4, 0.0, 8.0, 8.0, 2.0, 2.0, 0.0, 0.0, S, L), (2.0, 4.0)),                    # :248
            ((-0.5, -0.2, 8.0, 8.0, 1.0, 1.0, 2.0, 1.0, S, L), (2.0, 4.0)),                   # :258
            ((-0.5, -0.2, 8.0, 8.0, 2.0, 4.0, 2.0, 1.0, S, L), (0.5, 2.0))):                  # :268
        line, samp = P.computePixel([args[0]], [args[1]], *args[2:])
        assert _close([line[0], samp[0]], expect, 0), args


def test_kat_camera_look_vector():
    """createCameraLookVector (UtilitiesTests.cpp:278-284)."""
    from usgscsm_b200 import plugin as P
    look = P.createCameraLookVector([0.0], [-0.4], 1, 50)[0]
    assert _close(look, [0.0, 0.007999744, -0.999968001], 1e-8)


def test_kat_jitter():
    """removeJitter / addJitter (UtilitiesTests.cpp:103-236)."""
    import usgscsm_b200 as ub
    from usgscsm_b200 import plugin as P
    lj, sj, lt = [2.0, 3.0, 4.0], [-10.0, -20.0, -30.0], [0.0, 1.0, 2.0, 3.0, 4.0]
    rm = lambda l, s: tuple(v[0] for v in P.removeJitter([l], [s], lj, sj, lt))
    assert _close(rm(1, 0), (1, 0), 0)                                            # t = 0: nothing
    assert _close(rm(2, 5.0), (2 - (2.0 + 3.0 + 4.0), 5.0 + (10.0 + 20.0 + 30.0)), 0)   # t = 1
    assert _close(rm(3, 5.0), (3 - (8 * 2.0 + 4 * 3.0 + 2 * 4.0), 5.0 + (8 * 10.0 + 4 * 20.0 + 2 * 30.0)), 0)
    a, b = rm(1, 0), rm(-100, 0)                                                   # bounded to the image
    assert _close((-100 - b[0], 0 - b[1]), (1 - a[0], 0 - a[1]), 0)
    a, b = rm(5.0, 0), rm(100, 0)
    assert _close((100 - b[0], 0 - b[1]), (5.0 - a[0], 0 - a[1]), 0)
    a = rm(1.0, 5.0)                                                               # rounding of the line
    for frac in (0.75, 1.35):
        b = rm(frac, 5.0)
        assert _close((frac - b[0], 5.0 - b[1]), (1.0 - a[0], 5.0 - a[1]), 0)
    for bad in ((lj[:2], sj, lt), (lj, sj + [-40.0], lt), (lj, sj, [])):           # removeJitterErrors
        with pytest.raises(ub.Error):
            P.removeJitter([4], [15], *bad)
    lj2, sj2 = [0.01, 0.02, 0.03], [-0.01, -0.02, -0.03]                           # addJitter :203-236
    dl, ds = P.removeJitter([4], [15.0], lj2, sj2, lt)
    jl, js = P.addJitter(dl, ds, 1e-7, 30, lj2, sj2, lt)
    assert abs(jl[0] - 4) <= 1e-7 and abs(js[0] - 15.0) <= 1e-7


def test_kat_polynomial_and_roots():
    """polynomialEval, polynomialRoot, brentRoot (UtilitiesTests.cpp:443-475)."""
    from usgscsm_b200 import plugin as P, _capi
    c = [-12.0, 4.0, -3.0, 1.0]
    x = [-1.0, 0.0, 2.0, 3.5]
    assert _close(P.evaluatePolynomial(c, x), [-20.0, -12.0, -8.0, 8.125], 0)
    assert _close(P.evaluatePolynomialDerivative(c, x), [13.0, 4.0, 4.0, 19.75], 0)
    with pytest.raises(ValueError):
        P.evaluatePolynomial([], [0.0])
    root, st = P.polynomialRoot(c, [0.0])
    assert abs(root[0] - 3.0) <= 1e-10 and st[0] == 0
    root, st = P.polynomialRoot([4.0, 0.0, 1.0], [0.0])        # no real root: the reference throws
    assert st[0] & _capi.ST_ERROR_UNKNOWN and np.isnan(root[0])
    # (x - 2)(x + 1)(x + 7) = x^3 + 6 x^2 - 9 x - 14
    poly = [-14.0, -9.0, 6.0, 1.0]
    root, st = P.brentRoot([1.0, 0.0, -3.0], [3.0, -3.0, 3.0], poly, 1e-10)
    assert abs(root[0] - 2.0) <= 1e-10 and abs(root[1] + 1.0) <= 1e-10
    assert not (st[:2] & _capi.ST_ERROR_UNKNOWN).any()
    assert st[2] & _capi.ST_ERROR_UNKNOWN and np.isnan(root[2])   # brentRoot(-3, 3): std::invalid_argument
