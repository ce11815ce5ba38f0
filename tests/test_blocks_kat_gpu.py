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
