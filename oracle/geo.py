"""ctypes view of oracle/_ref/libgeo_oracle.so (oracle/geo_restate.c): the PROJ arithmetic of
UsgsAstroProjectedSensorModel restated on the CPU.  TEST INFRASTRUCTURE -- only tests/, smoke()
and bench.py's cpu_baseline leg may import this."""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libgeo_oracle.so")
PROJ_LONGLAT, PROJ_EQC, PROJ_SINU = 0, 1, 2
_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(LIB_PATH)
        _lib.geo_sizeof.restype = C.c_int
    return _lib


def _a(v):
    return np.ascontiguousarray(v, dtype=np.float64).reshape(-1)


class Geo(object):
    def __init__(self, geotransform, projection, lat_ts=0.0, lat_0=0.0, lon_0=0.0, x_0=0.0, y_0=0.0,
                 proj_a=0.0, proj_b=0.0, cart_a=0.0, cart_b=0.0):
        L = lib()
        self._g = C.create_string_buffer(L.geo_sizeof())
        gt = (C.c_double * 6)(*[float(v) for v in geotransform])
        L.geo_setup(gt, C.c_int(projection), *[C.c_double(float(v)) for v in (
            lat_ts, lat_0, lon_0, x_0, y_0, proj_a, proj_b, cart_a, cart_b)], self._g)

    def image_to_ground(self, line, samp, height=None):
        line, samp = _a(line), _a(samp)
        n = line.size
        h = None if height is None else np.ascontiguousarray(np.broadcast_to(_a(height), (n,)))
        x, y, z = np.empty(n), np.empty(n), np.empty(n)
        lib().geo_image_to_ground(self._g, C.c_int64(n), C.c_void_p(line.ctypes.data),
                                  C.c_void_p(samp.ctypes.data),
                                  C.c_void_p(h.ctypes.data) if h is not None else None,
                                  C.c_void_p(x.ctypes.data), C.c_void_p(y.ctypes.data),
                                  C.c_void_p(z.ctypes.data))
        return x, y, z

    def ground_to_image(self, x, y, z):
        x, y, z = _a(x), _a(y), _a(z)
        n = x.size
        line, samp, h = np.empty(n), np.empty(n), np.empty(n)
        lib().geo_ground_to_image(self._g, C.c_int64(n), C.c_void_p(x.ctypes.data),
                                  C.c_void_p(y.ctypes.data), C.c_void_p(z.ctypes.data),
                                  C.c_void_p(line.ctypes.data), C.c_void_p(samp.ctypes.data),
                                  C.c_void_p(h.ctypes.data))
        return line, samp, h

    def cart_forward_deg(self, lon, lat, h):
        out = (C.c_double * 3)()
        lib().geo_cart_forward_deg(self._g, C.c_double(lon), C.c_double(lat), C.c_double(h), out)
        return list(out)

    def cart_inverse_deg(self, x, y, z):
        xyz = (C.c_double * 3)(x, y, z)
        out = (C.c_double * 3)()
        lib().geo_cart_inverse_deg(self._g, xyz, out)
        return list(out)

    def proj_forward_deg(self, lon, lat):
        out = (C.c_double * 2)()
        lib().geo_proj_forward_deg(self._g, C.c_double(lon), C.c_double(lat), out)
        return list(out)

    def proj_inverse_deg(self, mx, my):
        out = (C.c_double * 2)()
        lib().geo_proj_inverse_deg(self._g, C.c_double(mx), C.c_double(my), out)
        return list(out)


def bilinear(dem, line, samp):
    """DEM height at raster coordinates (pixel centres at +0.5), the formula of the engine's
    DEM-iterated intersection: rows first along samples, then along lines.  NaN outside."""
    dem = np.asarray(dem, dtype=np.float64)
    nr, nc = dem.shape
    u, v = np.asarray(samp) - 0.5, np.asarray(line) - 0.5
    inside = (u >= 0) & (v >= 0) & (u <= nc - 1) & (v <= nr - 1)
    uu, vv = np.where(inside, u, 0.0), np.where(inside, v, 0.0)
    c0, r0 = np.floor(uu).astype(np.int64), np.floor(vv).astype(np.int64)
    c1, r1 = np.minimum(c0 + 1, nc - 1), np.minimum(r0 + 1, nr - 1)
    fu, fv = uu - c0, vv - r0
    top = dem[r0, c0] * (1.0 - fu) + dem[r0, c1] * fu
    bot = dem[r1, c0] * (1.0 - fu) + dem[r1, c1] * fu
    return np.where(inside, top * (1.0 - fv) + bot * fv, np.nan)
