"""Pin the oracle: replay the reference's own known-answer tests for the hot-path building
blocks (tests/golden/reference_kat.json, restated from tests/DistortionTests.cpp,
tests/UtilitiesTests.cpp, tests/EigenUtilitiesTests.cpp of the reference) against
oracle/_ref, the reference code compiled here with the ALE / Eigen / PROJ shims.  This is
what shows that the shims did not change the arithmetic.  CPU only."""
import ctypes as C
import json
import math
import os

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
KAT = json.load(open(os.path.join(HERE, "golden", "reference_kat.json")))


def _close(got, want, tol):
    got, want = np.asarray(got, dtype=np.float64), np.asarray(want, dtype=np.float64)
    if tol == 0:  # EXPECT_DOUBLE_EQ: within 4 ulp
        return bool(np.all(np.abs(got - want) <= 4 * np.spacing(np.abs(want))))
    return bool(np.all(np.abs(got - want) <= tol))


def _arr(v):
    return np.ascontiguousarray(v, dtype=np.float64)


@pytest.mark.parametrize("k", KAT["lagrange"], ids=lambda k: "%s@%g" % (k["src"], k["time"]))
def test_lagrange(ref, k):
    out = np.zeros(k["vlen"])
    v = _arr(k["values"])
    rc = ref.lib().ref_lagrange(k["n"], v.ctypes.data, k["t0"], k["dt"], k["time"], k["vlen"],
                                k["order"], out.ctypes.data)
    assert rc == 0 and _close(out, k["expect"], k["tol"])


def test_lagrange_too_few_points(ref):
    k = KAT["lagrange_error"]
    out, v = np.zeros(1), _arr(k["values"])
    assert ref.lib().ref_lagrange(k["n"], v.ctypes.data, k["t0"], k["dt"], k["time"], k["vlen"],
                                  k["order"], out.ctypes.data) != 0


def test_rotations(ref):
    for k in KAT["rotation_euler"]:
        e, m = _arr(k["euler"]), np.zeros(9)
        ref.lib().ref_rotation_from_euler(e.ctypes.data, m.ctypes.data)
        assert _close(m, k["expect"], k["tol"])
    for k in KAT["rotation_quaternion"]:
        q, m = _arr(k["q"]), np.zeros(9)
        ref.lib().ref_rotation_from_quaternion(q.ctypes.data, m.ctypes.data)
        assert _close(m, k["expect"], k["tol"])


def _ref_apply(ref, k, x, y, prec):
    dx, dy, c = C.c_double(), C.c_double(), _arr(k["coeffs"])
    rc = ref.lib().ref_apply_distortion(x, y, C.addressof(dx), C.addressof(dy), c.ctypes.data,
                                        c.size, k["focal"], k["type"], prec, 1e-6)
    assert rc == 0
    return dx.value, dy.value


def _ref_remove(ref, k, x, y, tol):
    ux, uy, c = C.c_double(), C.c_double(), _arr(k["coeffs"])
    rc = ref.lib().ref_remove_distortion(x, y, C.addressof(ux), C.addressof(uy), c.ctypes.data,
                                         c.size, k["focal"], k["type"], tol)
    assert rc == 0
    return ux.value, uy.value


@pytest.mark.parametrize("k", KAT["apply_distortion"], ids=lambda k: k["src"])
def test_apply_distortion(ref, k):
    assert _close(_ref_apply(ref, k, k["in"][0], k["in"][1], k["prec"]), k["expect"], k["tol"])


@pytest.mark.parametrize("k", KAT["remove_distortion"], ids=lambda k: k["src"])
def test_remove_distortion(ref, k):
    assert _close(_ref_remove(ref, k, k["in"][0], k["in"][1], k["tol_arg"]), k["expect"], k["tol"])


@pytest.mark.parametrize("k", KAT["round_trip"], ids=lambda k: k["src"])
def test_distortion_round_trip(ref, k):
    ux, uy = _ref_remove(ref, k, k["distorted"][0], k["distorted"][1], k["remove_tol"])
    assert _close(_ref_apply(ref, k, ux, uy, k["apply_prec"]), k["distorted"], k["tol"])


def test_projective_fit_recovers_transform(ref):
    """tests/EigenUtilitiesTests.cpp:8-45 against the Eigen-free restatement in oracle/ref_shim."""
    u = np.array([math.sin(i) for i in range(14)])
    rows = []
    for r in range(5):
        for c in range(5):
            x, y = float(c), float(r)
            z = 0.3 * x + 2 * y + 0.04 * x * y
            p = (u[0] + u[1] * x + u[2] * y + u[3] * z) / (1 + u[4] * x + u[5] * y + u[6] * z)
            q = (u[7] + u[8] * x + u[9] * y + u[10] * z) / (1 + u[11] * x + u[12] * y + u[13] * z)
            rows.append((p, q, x, y, z))
    a = np.ascontiguousarray(np.array(rows).T)
    out = np.zeros(14)
    rc = ref.lib().ref_projective_fit(25, a[0].ctypes.data, a[1].ctypes.data, a[2].ctypes.data,
                                      a[3].ctypes.data, a[4].ctypes.data, out.ctypes.data)
    assert rc == 0 and np.abs(out - u).max() <= 1e-12
