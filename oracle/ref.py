"""ctypes bridge to oracle/_ref/libusgscsm_ref.so -- TEST INFRASTRUCTURE, not the product.

The library is the UNMODIFIED reference sensor-model code compiled by oracle/Makefile from
/root/reference (plus ALE/PROJ/Eigen shims and the C driver oracle/ref_driver.cpp).  Only
tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; nothing under usgscsm_b200/ does.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libusgscsm_ref.so")

_dp = C.c_void_p
_lib = None


def available():
    return os.path.exists(LIB_PATH)


def lib():
    global _lib
    if _lib is None:
        if not available():
            raise RuntimeError("oracle/_ref/libusgscsm_ref.so is missing: run `make -C oracle ref` "
                               "in the build container (needs /root/reference)")
        L = C.CDLL(LIB_PATH)
        L.ref_last_error.restype = C.c_char_p
        L.ref_max_threads.restype = C.c_int
        L.ref_model_from_state.restype = C.c_void_p
        L.ref_model_from_state.argtypes = [C.c_char_p, C.c_size_t]
        L.ref_model_from_isd.restype = C.c_void_p
        L.ref_model_from_isd.argtypes = [C.c_char_p, C.c_char_p]
        L.ref_state_from_isd.restype = C.c_size_t
        L.ref_state_from_isd.argtypes = [C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]
        L.ref_ls_refit_projective.argtypes = [C.c_void_p]
        L.ref_model_free.argtypes = [C.c_void_p]
        L.ref_model_state.restype = C.c_size_t
        L.ref_model_state.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
        L.ref_ls_derived.argtypes = [C.c_void_p, _dp]
        L.ref_ls_set_projective.argtypes = [C.c_void_p, C.c_int, _dp]
        L.ref_set_parameter.argtypes = [C.c_void_p, C.c_int, C.c_double]
        L.ref_image_size.argtypes = [C.c_void_p, C.POINTER(C.c_int), C.POINTER(C.c_int)]
        L.ref_ground_to_image.argtypes = [C.c_void_p, C.c_long, _dp, _dp, _dp, C.c_double, _dp,
                                          _dp, _dp, _dp, C.c_int]
        L.ref_image_to_ground.argtypes = [C.c_void_p, C.c_long, _dp, _dp, _dp, C.c_int,
                                          C.c_double, _dp, _dp, _dp, _dp, _dp, C.c_int]
        L.ref_remote_locus.argtypes = [C.c_void_p, C.c_long, _dp, _dp, C.c_double] + [_dp] * 8 + [C.c_int]
        L.ref_proximate_locus.argtypes = [C.c_void_p, C.c_long] + [_dp] * 5 + [C.c_double] + [_dp] * 8 + [C.c_int]
        L.ref_sensor_partials.argtypes = [C.c_void_p, C.c_int, C.c_long] + [_dp] * 5 + [C.c_double] + [_dp] * 4 + [C.c_int]
        L.ref_all_sensor_partials.argtypes = [C.c_void_p, C.c_long] + [_dp] * 5 + [C.c_double] + [_dp] * 3 + [C.c_int]
        L.ref_image_sensor_state.argtypes = [C.c_void_p, C.c_long, _dp, _dp, _dp, _dp, C.c_int]
        L.ref_illumination_direction.argtypes = [C.c_void_p, C.c_long, _dp, _dp, _dp, _dp, _dp, C.c_int]
        L.ref_set_parameter_covariance.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_double]
        L.ref_ground_to_image_covar.argtypes = [C.c_void_p, C.c_long] + [_dp] * 4 + [C.c_double] + [_dp] * 5 + [C.c_int]
        L.ref_image_to_ground_covar.argtypes = ([C.c_void_p, C.c_long] + [_dp] * 4 + [C.c_int, _dp, C.c_int, C.c_double]
                                                + [_dp] * 6 + [C.c_int])
        L.ref_ground_partials.argtypes = [C.c_void_p, C.c_long, _dp, _dp, _dp, _dp, _dp, C.c_int]
        L.ref_set_ellipsoid.argtypes = [C.c_void_p, C.c_double, C.c_double]
        L.ref_lagrange.argtypes = [C.c_int, _dp, C.c_double, C.c_double, C.c_double, C.c_int,
                                   C.c_int, _dp]
        L.ref_remove_distortion.argtypes = [C.c_double, C.c_double, _dp, _dp, _dp, C.c_int,
                                            C.c_double, C.c_int, C.c_double]
        L.ref_apply_distortion.argtypes = [C.c_double, C.c_double, _dp, _dp, _dp, C.c_int,
                                           C.c_double, C.c_int, C.c_double, C.c_double]
        L.ref_ellipsoid_elevation.restype = C.c_double
        L.ref_ellipsoid_elevation.argtypes = [C.c_double] * 6 + [_dp]
        L.ref_rotation_from_quaternion.argtypes = [_dp, _dp]
        L.ref_rotation_from_euler.argtypes = [_dp, _dp]
        L.ref_projective_fit.argtypes = [C.c_int, _dp, _dp, _dp, _dp, _dp, _dp]
        _lib = L
    return _lib


def _a(x, n=None):
    x = np.ascontiguousarray(x, dtype=np.float64).reshape(-1)
    if n is not None and x.size != n:
        raise ValueError("length mismatch")
    return x


class RefModel(object):
    """A reference csm::RasterGM built by the reference plugin from a model-state string."""

    def __init__(self, state, _handle=None):
        self._L = lib()
        if _handle is not None:
            self._h = _handle
            return
        raw = state if isinstance(state, bytes) else state.encode("utf-8")
        self._h = self._L.ref_model_from_state(raw, len(raw))
        if not self._h:
            raise RuntimeError("reference could not build the model: " +
                               self._L.ref_last_error().decode("utf-8", "replace"))

    @classmethod
    def from_isd(cls, image_path, model_name=None):
        """UsgsAstroPlugin::constructModelFromISD(csm::Isd(image_path), model_name): the plugin
        reads the sidecar <basename>.json next to the (possibly absent) image file."""
        L = lib()
        h = L.ref_model_from_isd(os.fsencode(image_path),
                                 None if model_name is None else model_name.encode())
        if not h:
            raise RuntimeError("reference could not build the model from the ISD: " +
                               L.ref_last_error().decode("utf-8", "replace"))
        return cls(None, _handle=h)

    def ls_refit_projective(self):
        if self._L.ref_ls_refit_projective(self._h) != 0:
            raise RuntimeError("not a line-scan model")

    def __del__(self):
        if getattr(self, "_h", None):
            self._L.ref_model_free(self._h)
            self._h = None

    def state(self):
        n = self._L.ref_model_state(self._h, None, 0)
        buf = C.create_string_buffer(n)
        self._L.ref_model_state(self._h, buf, n)
        return buf.raw.decode("utf-8")

    def image_size(self):
        l, s = C.c_int(), C.c_int()
        self._L.ref_image_size(self._h, C.byref(l), C.byref(s))
        return l.value, s.value

    def ls_derived(self):
        out = np.zeros(22)
        if self._L.ref_ls_derived(self._h, out.ctypes.data) != 0:
            raise RuntimeError("not a line-scan model")
        return out

    def ls_set_projective(self, use, coeffs14=None):
        c = None if coeffs14 is None else _a(coeffs14, 14)
        self._L.ref_ls_set_projective(self._h, int(bool(use)), None if c is None else c.ctypes.data)

    def set_parameter(self, index, value):
        if self._L.ref_set_parameter(self._h, index, float(value)) != 0:
            raise RuntimeError(self._L.ref_last_error().decode())

    def ground_to_image(self, x, y, z, prec=0.001, threads=1):
        x = _a(x)
        n = x.size
        y, z = _a(y, n), _a(z, n)
        line, samp, ach = np.empty(n), np.empty(n), np.empty(n)
        st = np.empty(n, dtype=np.uint32)
        self._L.ref_ground_to_image(self._h, n, x.ctypes.data, y.ctypes.data, z.ctypes.data, prec,
                                    line.ctypes.data, samp.ctypes.data, ach.ctypes.data,
                                    st.ctypes.data, threads)
        return line, samp, ach, st

    def image_to_ground(self, line, samp, height, prec=0.001, threads=1):
        line = _a(line)
        n = line.size
        samp = _a(samp, n)
        height = _a(height)
        stride = 0 if (height.size == 1 and n != 1) else 1
        x, y, z, ach = np.empty(n), np.empty(n), np.empty(n), np.empty(n)
        st = np.empty(n, dtype=np.uint32)
        self._L.ref_image_to_ground(self._h, n, line.ctypes.data, samp.ctypes.data,
                                    height.ctypes.data, stride, prec, x.ctypes.data,
                                    y.ctypes.data, z.ctypes.data, ach.ctypes.data,
                                    st.ctypes.data, threads)
        return x, y, z, ach, st


    def remote_locus(self, line, samp, prec=0.001, threads=1):
        """imageToRemoteImagingLocus per pixel -> (px, py, pz, dx, dy, dz, achieved, status)."""
        line = _a(line)
        n = line.size
        samp = _a(samp, n)
        out = [np.empty(n) for _ in range(7)]
        st = np.empty(n, dtype=np.uint32)
        self._L.ref_remote_locus(self._h, n, line.ctypes.data, samp.ctypes.data, prec,
                                 *[o.ctypes.data for o in out], st.ctypes.data, threads)
        return tuple(out) + (st,)

    def proximate_locus(self, line, samp, x, y, z, prec=0.001, threads=1):
        """imageToProximateImagingLocus per (pixel, ground point)."""
        line = _a(line)
        n = line.size
        samp, x, y, z = _a(samp, n), _a(x, n), _a(y, n), _a(z, n)
        out = [np.empty(n) for _ in range(7)]
        st = np.empty(n, dtype=np.uint32)
        self._L.ref_proximate_locus(self._h, n, line.ctypes.data, samp.ctypes.data, x.ctypes.data,
                                    y.ctypes.data, z.ctypes.data, prec,
                                    *[o.ctypes.data for o in out], st.ctypes.data, threads)
        return tuple(out) + (st,)

    def sensor_partials(self, index, x, y, z, prec=0.001, image_pts=None, threads=1):
        """computeSensorPartials per point -> (dline, dsamp, achieved, status)."""
        x = _a(x)
        n = x.size
        y, z = _a(y, n), _a(z, n)
        il = isamp = None
        if image_pts is not None:
            il, isamp = _a(image_pts[0], n), _a(image_pts[1], n)
        dl, ds, ach = np.empty(n), np.empty(n), np.empty(n)
        st = np.empty(n, dtype=np.uint32)
        self._L.ref_sensor_partials(self._h, int(index), n, x.ctypes.data, y.ctypes.data, z.ctypes.data,
                                    None if il is None else il.ctypes.data,
                                    None if isamp is None else isamp.ctypes.data, prec,
                                    dl.ctypes.data, ds.ctypes.data, ach.ctypes.data, st.ctypes.data, threads)
        return dl, ds, ach, st

    def all_sensor_partials(self, num_params, x, y, z, prec=0.001, image_pts=None, threads=1):
        """computeAllSensorPartials (csm::param::VALID) per point ->
        ((n, num_params, 2) partials, achieved, status)."""
        x = _a(x)
        n = x.size
        y, z = _a(y, n), _a(z, n)
        il = isamp = None
        if image_pts is not None:
            il, isamp = _a(image_pts[0], n), _a(image_pts[1], n)
        p, ach = np.empty((n, num_params, 2)), np.empty(n)
        st = np.empty(n, dtype=np.uint32)
        got = self._L.ref_all_sensor_partials(
            self._h, n, x.ctypes.data, y.ctypes.data, z.ctypes.data,
            None if il is None else il.ctypes.data, None if isamp is None else isamp.ctypes.data,
            prec, p.ctypes.data, ach.ctypes.data, st.ctypes.data, threads)
        if got != num_params:
            raise RuntimeError("the reference model has %d parameters, not %d" % (got, num_params))
        return p, ach, st

    def image_sensor_state(self, line, samp, threads=1):
        """getImageTime / getSensorPosition / getSensorVelocity per point -> ((n, 7), status)."""
        line = _a(line)
        n = line.size
        samp = _a(samp, n)
        out = np.empty((n, 7))
        st = np.empty(n, dtype=np.uint32)
        self._L.ref_image_sensor_state(self._h, n, line.ctypes.data, samp.ctypes.data,
                                       out.ctypes.data, st.ctypes.data, threads)
        return out, st

    def illumination_direction(self, x, y, z, threads=1):
        """getIlluminationDirection per point -> ((n, 3), status)."""
        x = _a(x)
        n = x.size
        y, z = _a(y, n), _a(z, n)
        out = np.empty((n, 3))
        st = np.empty(n, dtype=np.uint32)
        self._L.ref_illumination_direction(self._h, n, x.ctypes.data, y.ctypes.data, z.ctypes.data,
                                           out.ctypes.data, st.ctypes.data, threads)
        return out, st

    def set_parameter_covariance(self, i, j, v):
        self._L.ref_set_parameter_covariance(self._h, int(i), int(j), float(v))

    def ground_to_image_covar(self, x, y, z, ground_cov, prec=0.001, threads=1):
        """groundToImage(EcefCoordCovar) per point -> (line, samp, (n, 4) cov, achieved, status)."""
        x = _a(x)
        n = x.size
        y, z = _a(y, n), _a(z, n)
        gc = np.ascontiguousarray(np.asarray(ground_cov, dtype=np.float64).reshape(n, 9))
        line, samp, ach, ic = np.empty(n), np.empty(n), np.empty(n), np.empty((n, 4))
        st = np.empty(n, dtype=np.uint32)
        self._L.ref_ground_to_image_covar(self._h, n, x.ctypes.data, y.ctypes.data, z.ctypes.data,
                                          gc.ctypes.data, prec, line.ctypes.data, samp.ctypes.data,
                                          ic.ctypes.data, ach.ctypes.data, st.ctypes.data, threads)
        return line, samp, ic, ach, st

    def image_to_ground_covar(self, line, samp, image_cov, height, height_var, prec=0.001, threads=1):
        """imageToGround(ImageCoordCovar, h, hVar) per point -> (x, y, z, (n, 9) cov, achieved, status)."""
        line = _a(line)
        n = line.size
        samp = _a(samp, n)
        ic = np.ascontiguousarray(np.asarray(image_cov, dtype=np.float64).reshape(n, 4))
        h, hv = _a(height, n), _a(height_var, n)
        x, y, z, ach, gc = np.empty(n), np.empty(n), np.empty(n), np.empty(n), np.empty((n, 9))
        st = np.empty(n, dtype=np.uint32)
        self._L.ref_image_to_ground_covar(self._h, n, line.ctypes.data, samp.ctypes.data, ic.ctypes.data,
                                          h.ctypes.data, 1, hv.ctypes.data, 1, prec, x.ctypes.data,
                                          y.ctypes.data, z.ctypes.data, gc.ctypes.data, ach.ctypes.data,
                                          st.ctypes.data, threads)
        return x, y, z, gc, ach, st

    def ground_partials(self, x, y, z, threads=1):
        """computeGroundPartials per point -> ((n, 6) partials, status)."""
        x = _a(x)
        n = x.size
        y, z = _a(y, n), _a(z, n)
        p = np.empty((n, 6))
        st = np.empty(n, dtype=np.uint32)
        self._L.ref_ground_partials(self._h, n, x.ctypes.data, y.ctypes.data, z.ctypes.data,
                                    p.ctypes.data, st.ctypes.data, threads)
        return p, st

    def set_ellipsoid(self, a, b):
        if self._L.ref_set_ellipsoid(self._h, float(a), float(b)) != 0:
            raise RuntimeError("model is not a csm::SettableEllipsoid")


def state_to_msgpack(state):
    """nlohmann::json::to_msgpack of a model state string (the reference's binary state)."""
    L = lib()
    L.ref_state_to_msgpack.restype = C.c_size_t
    L.ref_state_to_msgpack.argtypes = [C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]
    raw = state.encode("utf-8") if isinstance(state, str) else state
    n = L.ref_state_to_msgpack(raw, len(raw), None, 0)
    if n == 0:
        raise RuntimeError(L.ref_last_error().decode())
    buf = C.create_string_buffer(n)
    L.ref_state_to_msgpack(raw, len(raw), buf, n)
    return buf.raw


def state_from_isd(isd_text):
    """<Model>::constructStateFromIsd of an ISD JSON text -> the state as a JSON string (before
    populateModel derives anything)."""
    L = lib()
    raw = isd_text.encode("utf-8") if isinstance(isd_text, str) else isd_text
    n = L.ref_state_from_isd(raw, len(raw), None, 0)
    if n == 0:
        raise RuntimeError(L.ref_last_error().decode("utf-8", "replace"))
    buf = C.create_string_buffer(n)
    L.ref_state_from_isd(raw, len(raw), buf, n)
    return buf.raw.decode("utf-8")


def max_threads():
    return lib().ref_max_threads()
