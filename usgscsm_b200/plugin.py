"""Host-side mirror of the reference's plugin / sensor-model interface for the hot path.

Names, argument meaning and error behaviour follow the reference:

* ``UsgsAstroPlugin``  -- src/UsgsAstroPlugin.cpp (constructModelFromState :535-555,
  getModelNameFromModelState :176-218, canModelBeConstructedFromState :229-250);
* ``RasterGM.groundToImage / imageToGround`` -- csm/RasterGM.h, one point per call, warnings
  appended to an optional list, hard failures raised as ``Error`` (csm/Error.h:40-48);
* new here: ``groundToImageBatch / imageToGroundBatch / imageToGroundGrid`` -- whole arrays per
  call through the C-ABI (include/usgscsm_b200.h) into the sm_100a kernels.

Every call, scalar or batched, runs on the GPU; there is no CPU path.
"""
import collections
import ctypes as C
import os
import json

import numpy as np

from . import _capi
from ._capi import lib, check

ImageCoord = collections.namedtuple("ImageCoord", "line samp")
EcefCoord = collections.namedtuple("EcefCoord", "x y z")


class Warning(object):  # csm/Warning.h:39-47
    UNKNOWN_WARNING, DATA_NOT_AVAILABLE, PRECISION_NOT_MET, NEGATIVE_PRECISION, \
        IMAGE_COORD_OUT_OF_BOUNDS, IMAGE_ID_TOO_LONG, NO_INTERSECTION, DEPRECATED_FUNCTION = range(1, 9)

    def __init__(self, warning, message, function):
        self.warning, self.message, self.function = warning, message, function

    def getWarning(self):
        return self.warning

    def getMessage(self):
        return self.message

    def getFunction(self):
        return self.function


class Error(Exception):  # csm/Error.h:40-48
    ALGORITHM, BOUNDS, FILE_READ, FILE_WRITE, ILLEGAL_MATH_OPERATION, INDEX_OUT_OF_RANGE, \
        INVALID_SENSOR_MODEL_STATE, INVALID_USE, ISD_NOT_SUPPORTED, MEMORY, \
        SENSOR_MODEL_NOT_CONSTRUCTIBLE, SENSOR_MODEL_NOT_SUPPORTED, STRING_TOO_LONG, \
        UNKNOWN_ERROR, UNSUPPORTED_FUNCTION, UNKNOWN_SUPPORT_DATA = range(1, 17)

    def __init__(self, error, message, function):
        super().__init__(message)
        self.error, self.message, self.function = error, message, function

    def getError(self):
        return self.error


_MODEL_NAMES = {
    "USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL": _capi.MODEL_LINE_SCANNER,
    "USGS_ASTRO_PUSH_FRAME_SENSOR_MODEL": _capi.MODEL_PUSH_FRAME,
    "USGS_ASTRO_FRAME_SENSOR_MODEL": _capi.MODEL_FRAME,
    "USGS_ASTRO_SAR_SENSOR_MODEL": _capi.MODEL_SAR,
}


def _is_torch(a):
    return hasattr(a, "data_ptr") and hasattr(a, "is_cuda")


def _host(a, n=None):
    a = np.ascontiguousarray(a, dtype=np.float64).reshape(-1)
    if n is not None and a.size != n:
        raise ValueError("array length mismatch")
    return a


def _ptr(a):
    if a is None:
        return None
    if _is_torch(a):
        # raw addresses cross the C-ABI: a strided or CPU tensor would be read as garbage
        if not (a.is_cuda and a.is_contiguous()):
            raise ValueError("torch arguments must be contiguous CUDA tensors")
        return a.data_ptr()
    return a.ctypes.data


def _dev_inputs(model, n, *tensors, **kw):
    """Checks of the device-pointer flavours' inputs: CUDA, float64, contiguous, `per` x n
    elements, on the model's device.  The C-ABI sees raw addresses and cannot check any of it."""
    import torch
    per = kw.get("per", 1)
    for t in tensors:
        if t is None:
            continue
        if not _is_torch(t) or not t.is_cuda:
            raise ValueError("expected a CUDA tensor")
        if t.dtype != torch.float64:
            raise ValueError("expected a float64 tensor, got %s" % t.dtype)
        if not t.is_contiguous():
            raise ValueError("expected a contiguous tensor")
        if t.numel() != per * n:
            raise ValueError("tensor length %d, expected %d" % (t.numel(), per * n))
        if t.device.index != model.device:
            raise ValueError("tensor on cuda:%s, the model lives on cuda:%d" % (t.device.index, model.device))


class RasterGM(object):
    """A sensor model living on one GPU (csm::RasterGM of the reference)."""

    def __init__(self, handle, state):
        self._h = handle
        self._state = state

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            lib.usgscsm_b200_model_free(h)

    # ---- csm::Model / GeometricModel subset --------------------------------------------
    def getModelName(self):
        kind = lib.usgscsm_b200_model_kind(self._h)
        return [k for k, v in _MODEL_NAMES.items() if v == kind][0]

    def getModelState(self):
        """csm::Model::getModelState: the state with the current parameter values, parameter
        covariance and ellipsoid written back."""
        need = C.c_size_t()
        check(lib.usgscsm_b200_model_get_state(self._h, None, 0, C.byref(need)))
        buf = C.create_string_buffer(need.value)
        check(lib.usgscsm_b200_model_get_state(self._h, buf, need.value, C.byref(need)))
        return buf.raw[:need.value].decode("utf-8")

    def getImageSize(self):
        l, s = C.c_int(), C.c_int()
        check(lib.usgscsm_b200_model_image_size(self._h, C.byref(l), C.byref(s)))
        return ImageCoord(l.value, s.value)

    def getValidImageRange(self):
        sz = self.getImageSize()
        return ImageCoord(0.0, 0.0), ImageCoord(float(sz.line), float(sz.samp))

    def getNumParameters(self):
        return lib.usgscsm_b200_model_num_parameters(self._h)

    def getParameterValue(self, index):
        v = C.c_double()
        check(lib.usgscsm_b200_model_get_parameter(self._h, index, C.byref(v)))
        return v.value

    def setParameterValue(self, index, value):
        check(lib.usgscsm_b200_model_set_parameter(self._h, index, float(value)))

    @property
    def device(self):
        return lib.usgscsm_b200_model_device(self._h)

    def lsDerived(self):
        out = (C.c_double * 22)()
        check(lib.usgscsm_b200_model_ls_derived(self._h, out))
        return np.array(out[:], dtype=np.float64)

    def lsSetProjective(self, use, coeffs14=None):
        c = None
        if coeffs14 is not None:
            c = (C.c_double * 14)(*[float(v) for v in coeffs14])
        check(lib.usgscsm_b200_model_ls_set_projective(self._h, int(bool(use)), c))

    # ---- batched per-point math ---------------------------------------------------------
    def groundToImageBatch(self, x, y, z, desiredPrecision=0.001, achieved=True, status=True,
                           stream=None, out=None):
        """groundToImage for arrays.  numpy in -> numpy out (host flavour, copies included);
        torch CUDA tensors in -> torch CUDA tensors out (device flavour, enqueued on `stream`
        or torch's current stream, not synchronised).  Returns (line, samp, achieved, status)."""
        if _is_torch(x):
            import torch
            n = x.numel()
            _dev_inputs(self, n, x, y, z)
            if out is not None:
                line, samp, ach, st = out
            else:
                line, samp = torch.empty_like(x), torch.empty_like(x)
                ach = torch.empty_like(x) if achieved else None
                st = torch.empty(n, dtype=torch.int32, device=x.device) if status else None
            if stream is None:
                stream = torch.cuda.current_stream(x.device).cuda_stream
            check(lib.usgscsm_b200_ground_to_image_dev(
                self._h, n, _ptr(x), _ptr(y), _ptr(z), desiredPrecision, _ptr(line), _ptr(samp),
                _ptr(ach), _ptr(st), stream))
            return line, samp, ach, st
        x = _host(x)
        n = x.size
        y, z = _host(y, n), _host(z, n)
        line, samp = np.empty(n), np.empty(n)
        ach = np.empty(n) if achieved else None
        st = np.empty(n, dtype=np.uint32) if status else None
        check(lib.usgscsm_b200_ground_to_image(self._h, n, _ptr(x), _ptr(y), _ptr(z),
                                               desiredPrecision, _ptr(line), _ptr(samp),
                                               _ptr(ach), _ptr(st)))
        return line, samp, ach, st

    def imageToGroundBatch(self, line, samp, height, desiredPrecision=0.001, achieved=True,
                           status=True, stream=None):
        """imageToGround for arrays; `height` is a scalar or an array of the same length.
        Returns (x, y, z, achieved, status)."""
        if _is_torch(line):
            import torch
            n = line.numel()
            _dev_inputs(self, n, line, samp)
            if not _is_torch(height):
                height = torch.full((1,), float(height), dtype=torch.float64, device=line.device)
            stride = 0 if height.numel() == 1 else 1
            _dev_inputs(self, n if stride else 1, height)
            x, y, z = torch.empty_like(line), torch.empty_like(line), torch.empty_like(line)
            ach = torch.empty_like(line) if achieved else None
            st = torch.empty(n, dtype=torch.int32, device=line.device) if status else None
            if stream is None:
                stream = torch.cuda.current_stream(line.device).cuda_stream
            check(lib.usgscsm_b200_image_to_ground_dev(
                self._h, n, _ptr(line), _ptr(samp), _ptr(height), stride, desiredPrecision,
                _ptr(x), _ptr(y), _ptr(z), _ptr(ach), _ptr(st), stream))
            return x, y, z, ach, st
        line = _host(line)
        n = line.size
        samp = _host(samp, n)
        height = _host(height)
        if height.size not in (1, n):
            raise ValueError("height must be a scalar or match line/samp")
        stride = 0 if height.size == 1 and n != 1 else 1
        x, y, z = np.empty(n), np.empty(n), np.empty(n)
        ach = np.empty(n) if achieved else None
        st = np.empty(n, dtype=np.uint32) if status else None
        check(lib.usgscsm_b200_image_to_ground(self._h, n, _ptr(line), _ptr(samp), _ptr(height),
                                               stride, desiredPrecision, _ptr(x), _ptr(y),
                                               _ptr(z), _ptr(ach), _ptr(st)))
        return x, y, z, ach, st

    def imageToGroundGrid(self, line0, samp0, dline, dsamp, n_lines, n_samps, height,
                          desiredPrecision=0.001, achieved=False, status=True, out=None,
                          stream=None):
        """Dense-grid imageToGround (bin/usgscsm_cam_test.cc:294-309) with pixel coordinates
        generated in the kernel.  `height`: torch CUDA tensor with 1 or n_lines*n_samps
        entries.  Returns torch CUDA tensors (x, y, z, achieved, status)."""
        n = int(n_lines) * int(n_samps)
        if not _is_torch(height):   # host arrays in, host arrays out (copies inside the call)
            hh = _host(height)
            if hh.size not in (1, n):
                raise ValueError("height must hold 1 or n_lines * n_samps entries")
            x, y, z = np.empty(n), np.empty(n), np.empty(n)
            ach = np.empty(n) if achieved else None
            st = np.empty(n, dtype=np.uint32) if status else None
            check(lib.usgscsm_b200_image_to_ground_grid(
                self._h, float(line0), float(samp0), float(dline), float(dsamp), int(n_lines),
                int(n_samps), _ptr(hh), 0 if hh.size == 1 and n != 1 else 1, desiredPrecision,
                _ptr(x), _ptr(y), _ptr(z), _ptr(ach), _ptr(st)))
            return x, y, z, ach, st
        import torch
        dev = height.device
        stride = 0 if height.numel() == 1 else 1
        _dev_inputs(self, n if stride else 1, height)
        if out is None:
            x, y, z = (torch.empty(n, dtype=torch.float64, device=dev) for _ in range(3))
            ach = torch.empty(n, dtype=torch.float64, device=dev) if achieved else None
            st = torch.empty(n, dtype=torch.int32, device=dev) if status else None
        else:
            x, y, z, ach, st = out
        if stream is None:
            stream = torch.cuda.current_stream(dev).cuda_stream
        check(lib.usgscsm_b200_image_to_ground_grid_dev(
            self._h, float(line0), float(samp0), float(dline), float(dsamp), int(n_lines),
            int(n_samps), _ptr(height), stride, desiredPrecision, _ptr(x), _ptr(y), _ptr(z),
            _ptr(ach), _ptr(st), stream))
        return x, y, z, ach, st

    def imageToRemoteImagingLocusBatch(self, line, samp, desiredPrecision=0.001):
        """csm::RasterGM::imageToRemoteImagingLocus for arrays (host numpy or torch CUDA).
        Returns (point (n,3)-like tuple px,py,pz, direction dx,dy,dz, achieved, status)."""
        if _is_torch(line):
            import torch
            n = line.numel()
            _dev_inputs(self, n, line, samp)
            out = [torch.empty_like(line) for _ in range(7)]
            st = torch.empty(n, dtype=torch.int32, device=line.device)
            stream = torch.cuda.current_stream(line.device).cuda_stream
            check(lib.usgscsm_b200_image_to_remote_imaging_locus_dev(
                self._h, n, _ptr(line), _ptr(samp), desiredPrecision, *[_ptr(o) for o in out],
                _ptr(st), stream))
            return tuple(out) + (st,)
        line = _host(line)
        n = line.size
        samp = _host(samp, n)
        out = [np.empty(n) for _ in range(7)]
        st = np.empty(n, dtype=np.uint32)
        check(lib.usgscsm_b200_image_to_remote_imaging_locus(
            self._h, n, _ptr(line), _ptr(samp), desiredPrecision, *[_ptr(o) for o in out], _ptr(st)))
        return tuple(out) + (st,)

    def imageToProximateImagingLocusBatch(self, line, samp, x, y, z, desiredPrecision=0.001):
        """csm::RasterGM::imageToProximateImagingLocus for arrays (host numpy or torch CUDA):
        (px, py, pz, dx, dy, dz, achieved, status)."""
        if _is_torch(line):
            import torch
            n = line.numel()
            _dev_inputs(self, n, line, samp, x, y, z)
            out = [torch.empty_like(line) for _ in range(7)]
            st = torch.empty(n, dtype=torch.int32, device=line.device)
            stream = torch.cuda.current_stream(line.device).cuda_stream
            check(lib.usgscsm_b200_image_to_proximate_imaging_locus_dev(
                self._h, n, _ptr(line), _ptr(samp), _ptr(x), _ptr(y), _ptr(z), desiredPrecision,
                *[_ptr(o) for o in out], _ptr(st), stream))
            return tuple(out) + (st,)
        line = _host(line)
        n = line.size
        samp, x, y, z = _host(samp, n), _host(x, n), _host(y, n), _host(z, n)
        out = [np.empty(n) for _ in range(7)]
        st = np.empty(n, dtype=np.uint32)
        check(lib.usgscsm_b200_image_to_proximate_imaging_locus(
            self._h, n, _ptr(line), _ptr(samp), _ptr(x), _ptr(y), _ptr(z), desiredPrecision,
            *[_ptr(o) for o in out], _ptr(st)))
        return tuple(out) + (st,)

    def computeGroundPartialsBatch(self, x, y, z):
        """csm::RasterGM::computeGroundPartials for arrays -> ((n, 6) partials, status)."""
        if _is_torch(x):
            import torch
            n = x.numel()
            _dev_inputs(self, n, x, y, z)
            p = torch.empty((n, 6), dtype=torch.float64, device=x.device)
            st = torch.empty(n, dtype=torch.int32, device=x.device)
            stream = torch.cuda.current_stream(x.device).cuda_stream
            check(lib.usgscsm_b200_ground_partials_dev(self._h, n, _ptr(x), _ptr(y), _ptr(z),
                                                       _ptr(p), _ptr(st), stream))
            return p, st
        x = _host(x)
        n = x.size
        y, z = _host(y, n), _host(z, n)
        p = np.empty((n, 6))
        st = np.empty(n, dtype=np.uint32)
        check(lib.usgscsm_b200_ground_partials(self._h, n, _ptr(x), _ptr(y), _ptr(z), _ptr(p), _ptr(st)))
        return p, st

    def computeSensorPartialsBatch(self, index, x, y, z, desiredPrecision=0.001, imagePts=None):
        """csm::RasterGM::computeSensorPartials for arrays: returns
        (dLine/dParam, dSamp/dParam, achieved, status).  `imagePts` = (line, samp) arrays selects
        the overload that takes the image point."""
        il, isamp = (None, None) if imagePts is None else imagePts
        if _is_torch(x):
            import torch
            n = x.numel()
            _dev_inputs(self, n, x, y, z)
            dl, ds, ach = (torch.empty_like(x) for _ in range(3))
            st = torch.empty(n, dtype=torch.int32, device=x.device)
            stream = torch.cuda.current_stream(x.device).cuda_stream
            check(lib.usgscsm_b200_sensor_partials_dev(
                self._h, int(index), n, _ptr(x), _ptr(y), _ptr(z), _ptr(il), _ptr(isamp),
                desiredPrecision, _ptr(dl), _ptr(ds), _ptr(ach), _ptr(st), stream))
            return dl, ds, ach, st
        x = _host(x)
        n = x.size
        y, z = _host(y, n), _host(z, n)
        if il is not None:
            il, isamp = _host(il, n), _host(isamp, n)
        dl, ds, ach = np.empty(n), np.empty(n), np.empty(n)
        st = np.empty(n, dtype=np.uint32)
        check(lib.usgscsm_b200_sensor_partials(
            self._h, int(index), n, _ptr(x), _ptr(y), _ptr(z), _ptr(il), _ptr(isamp),
            desiredPrecision, _ptr(dl), _ptr(ds), _ptr(ach), _ptr(st)))
        return dl, ds, ach, st

    def computeAllSensorPartialsBatch(self, x, y, z, desiredPrecision=0.001, imagePts=None,
                                      indices=None):
        """csm::RasterGM::computeAllSensorPartials for arrays: returns ((n, P, 2) partials --
        (dLine, dSamp) per requested parameter --, achieved, status).  `indices` = the parameter
        indices of a csm::param::Set (default: all parameters); `imagePts` = (line, samp)
        arrays selects the overload that takes the image point."""
        il, isamp = (None, None) if imagePts is None else imagePts
        if indices is None:
            npar, idx = self.getNumParameters(), None
        else:
            npar = len(indices)
            idx = (C.c_int * max(npar, 1))(*[int(i) for i in indices])
        if _is_torch(x):
            import torch
            n = x.numel()
            _dev_inputs(self, n, x, y, z)
            p = torch.empty((n, npar, 2), dtype=torch.float64, device=x.device)
            ach = torch.empty_like(x)
            st = torch.empty(n, dtype=torch.int32, device=x.device)
            stream = torch.cuda.current_stream(x.device).cuda_stream
            check(lib.usgscsm_b200_all_sensor_partials_dev(
                self._h, npar, idx, n, _ptr(x), _ptr(y), _ptr(z), _ptr(il), _ptr(isamp),
                desiredPrecision, _ptr(p), _ptr(ach), _ptr(st), stream))
            return p, ach, st
        x = _host(x)
        n = x.size
        y, z = _host(y, n), _host(z, n)
        if il is not None:
            il, isamp = _host(il, n), _host(isamp, n)
        p, ach = np.empty((n, npar, 2)), np.empty(n)
        st = np.empty(n, dtype=np.uint32)
        check(lib.usgscsm_b200_all_sensor_partials(
            self._h, npar, idx, n, _ptr(x), _ptr(y), _ptr(z), _ptr(il), _ptr(isamp),
            desiredPrecision, _ptr(p), _ptr(ach), _ptr(st)))
        return p, ach, st

    def imageSensorStateBatch(self, line, samp):
        """getImageTime / getSensorPosition / getSensorVelocity of image points, for arrays:
        returns (time, (px, py, pz), (vx, vy, vz), status)."""
        if _is_torch(line):
            import torch
            n = line.numel()
            _dev_inputs(self, n, line, samp)
            o = [torch.empty_like(line) for _ in range(7)]
            st = torch.empty(n, dtype=torch.int32, device=line.device)
            check(lib.usgscsm_b200_image_sensor_state_dev(
                self._h, n, _ptr(line), _ptr(samp), *[_ptr(t) for t in o], _ptr(st),
                torch.cuda.current_stream(line.device).cuda_stream))
        else:
            line = _host(line)
            n = line.size
            samp = _host(samp, n)
            o = [np.empty(n) for _ in range(7)]
            st = np.empty(n, dtype=np.uint32)
            check(lib.usgscsm_b200_image_sensor_state(
                self._h, n, _ptr(line), _ptr(samp), *[_ptr(t) for t in o], _ptr(st)))
        return o[0], tuple(o[1:4]), tuple(o[4:7]), st

    def getImageTime(self, imagePt):
        return float(self.imageSensorStateBatch([imagePt[0]], [imagePt[1]])[0][0])

    def getSensorPosition(self, imagePt):
        _, p, _, _ = self.imageSensorStateBatch([imagePt[0]], [imagePt[1]])
        return tuple(float(c[0]) for c in p)

    def getSensorVelocity(self, imagePt):
        _, _, v, st = self.imageSensorStateBatch([imagePt[0]], [imagePt[1]])
        self._raise_or_warn(int(st[0]) & ~_capi.ST_NONFINITE, None, "getSensorVelocity")
        return tuple(float(c[0]) for c in v)

    def getIlluminationDirectionBatch(self, x, y, z):
        """csm::RasterGM::getIlluminationDirection for arrays: returns (dx, dy, dz, status)."""
        if _is_torch(x):
            import torch
            n = x.numel()
            _dev_inputs(self, n, x, y, z)
            d = [torch.empty_like(x) for _ in range(3)]
            st = torch.empty(n, dtype=torch.int32, device=x.device)
            check(lib.usgscsm_b200_illumination_direction_dev(
                self._h, n, _ptr(x), _ptr(y), _ptr(z), *[_ptr(t) for t in d], _ptr(st),
                torch.cuda.current_stream(x.device).cuda_stream))
        else:
            x = _host(x)
            n = x.size
            y, z = _host(y, n), _host(z, n)
            d = [np.empty(n) for _ in range(3)]
            st = np.empty(n, dtype=np.uint32)
            check(lib.usgscsm_b200_illumination_direction(
                self._h, n, _ptr(x), _ptr(y), _ptr(z), *[_ptr(t) for t in d], _ptr(st)))
        return d[0], d[1], d[2], st

    def getIlluminationDirection(self, groundPt):
        dx, dy, dz, st = self.getIlluminationDirectionBatch([groundPt[0]], [groundPt[1]], [groundPt[2]])
        self._raise_or_warn(int(st[0]) & ~_capi.ST_NONFINITE, None, "getIlluminationDirection")
        return float(dx[0]), float(dy[0]), float(dz[0])

    # ---- error propagation (csm::EcefCoordCovar / csm::ImageCoordCovar overloads) -------------
    def getParameterCovariance(self, index1, index2):
        v = C.c_double()
        check(lib.usgscsm_b200_model_get_parameter_covariance(self._h, int(index1), int(index2),
                                                              C.byref(v)))
        return v.value

    def setParameterCovariance(self, index1, index2, covariance):
        check(lib.usgscsm_b200_model_set_parameter_covariance(self._h, int(index1), int(index2),
                                                              float(covariance)))

    def groundToImageCovarBatch(self, x, y, z, groundCov, desiredPrecision=0.001):
        """groundToImage(csm::EcefCoordCovar) for arrays: groundCov (n, 9) row-major 3x3 ->
        (line, samp, imageCov (n, 4), achieved, status)."""
        if _is_torch(x):
            import torch
            n = x.numel()
            _dev_inputs(self, n, x, y, z)
            line, samp, ach = (torch.empty_like(x) for _ in range(3))
            icov = torch.empty((n, 4), dtype=torch.float64, device=x.device)
            st = torch.empty(n, dtype=torch.int32, device=x.device)
            check(lib.usgscsm_b200_ground_to_image_covar_dev(
                self._h, n, _ptr(x), _ptr(y), _ptr(z), _ptr(groundCov.contiguous()),
                desiredPrecision, _ptr(line), _ptr(samp), _ptr(icov), _ptr(ach), _ptr(st),
                torch.cuda.current_stream(x.device).cuda_stream))
            return line, samp, icov, ach, st
        x = _host(x)
        n = x.size
        y, z = _host(y, n), _host(z, n)
        gcov = np.ascontiguousarray(np.broadcast_to(np.asarray(groundCov, dtype=np.float64).reshape(-1, 9), (n, 9)))
        line, samp, ach, icov = np.empty(n), np.empty(n), np.empty(n), np.empty((n, 4))
        st = np.empty(n, dtype=np.uint32)
        check(lib.usgscsm_b200_ground_to_image_covar(
            self._h, n, _ptr(x), _ptr(y), _ptr(z), _ptr(gcov), desiredPrecision, _ptr(line),
            _ptr(samp), _ptr(icov), _ptr(ach), _ptr(st)))
        return line, samp, icov, ach, st

    def imageToGroundCovarBatch(self, line, samp, imageCov, height, heightVariance,
                                desiredPrecision=0.001):
        """imageToGround(csm::ImageCoordCovar, height, heightVariance) for arrays: imageCov
        (n, 4); height / heightVariance scalars or per-point arrays ->
        (x, y, z, groundCov (n, 9), achieved, status)."""
        if _is_torch(line):
            import torch
            n = line.numel()
            _dev_inputs(self, n, line, samp)
            dev = line.device

            def per_point(v):
                if _is_torch(v):
                    return v.contiguous(), (1 if v.numel() > 1 or n == 1 else 0)
                return torch.full((1,), float(v), dtype=torch.float64, device=dev), 0
            h, hs = per_point(height)
            hv, vs = per_point(heightVariance)
            x, y, z, ach = (torch.empty_like(line) for _ in range(4))
            gcov = torch.empty((n, 9), dtype=torch.float64, device=dev)
            st = torch.empty(n, dtype=torch.int32, device=dev)
            check(lib.usgscsm_b200_image_to_ground_covar_dev(
                self._h, n, _ptr(line), _ptr(samp), _ptr(imageCov.contiguous()), _ptr(h), hs,
                _ptr(hv), vs, desiredPrecision, _ptr(x), _ptr(y), _ptr(z), _ptr(gcov), _ptr(ach),
                _ptr(st), torch.cuda.current_stream(dev).cuda_stream))
            return x, y, z, gcov, ach, st
        line = _host(line)
        n = line.size
        samp = _host(samp, n)
        icov = np.ascontiguousarray(np.broadcast_to(np.asarray(imageCov, dtype=np.float64).reshape(-1, 4), (n, 4)))
        h = np.ascontiguousarray(np.atleast_1d(np.asarray(height, dtype=np.float64)))
        hv = np.ascontiguousarray(np.atleast_1d(np.asarray(heightVariance, dtype=np.float64)))
        for a in (h, hv):
            if a.size not in (1, n):
                raise ValueError("height / heightVariance must hold 1 or n entries")
        x, y, z, ach, gcov = np.empty(n), np.empty(n), np.empty(n), np.empty(n), np.empty((n, 9))
        st = np.empty(n, dtype=np.uint32)
        check(lib.usgscsm_b200_image_to_ground_covar(
            self._h, n, _ptr(line), _ptr(samp), _ptr(icov), _ptr(h), 1 if h.size == n else 0,
            _ptr(hv), 1 if hv.size == n else 0, desiredPrecision, _ptr(x), _ptr(y), _ptr(z),
            _ptr(gcov), _ptr(ach), _ptr(st)))
        return x, y, z, gcov, ach, st

    def computeGroundPartials(self, groundPt):
        """csm::RasterGM::computeGroundPartials (one point): a list of six partials."""
        p, st = self.computeGroundPartialsBatch([groundPt[0]], [groundPt[1]], [groundPt[2]])
        self._raise_or_warn(int(st[0]), None, "computeGroundPartials")
        return [float(v) for v in p[0]]

    def setEllipsoid(self, semiMajor, semiMinor):
        """csm::SettableEllipsoid::setEllipsoid."""
        check(lib.usgscsm_b200_model_set_ellipsoid(self._h, float(semiMajor), float(semiMinor)))

    def getEllipsoid(self):
        a, b = C.c_double(), C.c_double()
        check(lib.usgscsm_b200_model_get_ellipsoid(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    # ---- csm::RasterGM per-point API (a batch of one) ---------------------------------------
    @staticmethod
    def _raise_or_warn(st, warnings, function):
        if st & _capi.ST_ERROR_ALGORITHM:
            raise Error(Error.ALGORITHM, "Unable to apply transform.", function)
        if st & _capi.ST_ERROR_UNKNOWN:
            raise Error(Error.UNKNOWN_ERROR, "Could not calculate the image coordinate.", function)
        if st & _capi.ST_ERROR_OTHER:
            raise Error(Error.UNKNOWN_ERROR, "per-point failure", function)
        if warnings is not None:
            if st & _capi.ST_NO_INTERSECTION:
                warnings.append(Warning(Warning.NO_INTERSECTION, "No Intersection", function))
            if st & _capi.ST_PRECISION_NOT_MET:
                warnings.append(Warning(Warning.PRECISION_NOT_MET,
                                        "Desired precision not achieved.", function))

    def groundToImage(self, groundPt, desiredPrecision=0.001, achievedPrecision=None,
                      warnings=None):
        """csm::RasterGM::groundToImage.  `achievedPrecision`: optional list, receives one
        value; `warnings`: optional list, receives Warning objects."""
        l, s, a, st = self.groundToImageBatch([groundPt[0]], [groundPt[1]], [groundPt[2]],
                                              desiredPrecision)
        self._raise_or_warn(int(st[0]), warnings, "groundToImage")
        if achievedPrecision is not None:
            achievedPrecision.append(float(a[0]))
        return ImageCoord(float(l[0]), float(s[0]))

    def imageToGround(self, imagePt, height, desiredPrecision=0.001, achievedPrecision=None,
                      warnings=None):
        """csm::RasterGM::imageToGround."""
        x, y, z, a, st = self.imageToGroundBatch([imagePt[0]], [imagePt[1]], [height],
                                                 desiredPrecision)
        self._raise_or_warn(int(st[0]), warnings, "imageToGround")
        if achievedPrecision is not None:
            achievedPrecision.append(float(a[0]))
        return EcefCoord(float(x[0]), float(y[0]), float(z[0]))


class UsgsAstroPlugin(object):
    """src/UsgsAstroPlugin.cpp: the factory; models are built from model-state strings."""

    _PLUGIN_NAME = "UsgsAstroPluginCSM"
    _MANUFACTURER_NAME = "UsgsAstrogeology"
    _RELEASE_DATE = "20190222"

    def __init__(self, device=0):
        self.device = device

    def getPluginName(self):
        return self._PLUGIN_NAME

    def getManufacturer(self):
        return self._MANUFACTURER_NAME

    def getReleaseDate(self):
        return self._RELEASE_DATE

    def getNumModels(self):
        return len(_MODEL_NAMES)

    def getModelName(self, index):
        return list(_MODEL_NAMES)[index]

    @staticmethod
    def _state_bytes(state):
        return state if isinstance(state, bytes) else state.encode("utf-8")

    def getModelNameFromModelState(self, modelState):
        """UsgsAstroPlugin.cpp:176-218: JSON key m_modelName (after stateAsJson)."""
        raw = self._state_bytes(modelState)
        buf = C.create_string_buffer(128)
        code = lib.usgscsm_b200_model_name_from_state(raw, len(raw), buf, 128)
        if code == _capi.EUNSUPPORTED:
            raise Error(Error.SENSOR_MODEL_NOT_SUPPORTED, "Sensor model not supported.",
                        "UsgsAstroPlugin::getModelNameFromModelState")
        if code != _capi.OK:
            raise Error(Error.INVALID_SENSOR_MODEL_STATE, "No 'm_modelName' key in the model state object.",
                        "UsgsAstroPlugin::getModelNameFromModelState")
        return buf.value.decode("ascii")

    def canModelBeConstructedFromState(self, modelName, modelState, warnings=None):
        try:
            return self.getModelNameFromModelState(modelState) == modelName
        except Error:
            return False

    def constructModelFromState(self, modelState, warnings=None, device=None):
        raw = self._state_bytes(modelState)
        h = C.c_void_p()
        code = lib.usgscsm_b200_model_from_state(raw, len(raw), self.device if device is None else device,
                                                 C.byref(h))
        if code in (_capi.ESTATE, _capi.EUNSUPPORTED):
            raise Error(Error.SENSOR_MODEL_NOT_CONSTRUCTIBLE if code == _capi.ESTATE
                        else Error.SENSOR_MODEL_NOT_SUPPORTED,
                        lib.usgscsm_b200_last_error().decode("utf-8", "replace"),
                        "UsgsAstroPlugin::constructModelFromState")
        check(code)
        return RasterGM(h, modelState)


    # ---- ISD side (include/usgscsm/UsgsAstroPlugin.h:42-72) ---------------------------------------
    @staticmethod
    def loadImageSupportData(imageSupportData):
        """src/UsgsAstroPlugin.cpp:324-353: `imageSupportData` is a csm::Isd, i.e. the IMAGE file
        name (a str / os.PathLike, or an object with .filename()); the ISD is the sidecar
        "<base name>.json" next to it.  Returns (ISD text, image identifier)."""
        name = imageSupportData.filename() if hasattr(imageSupportData, "filename") else os.fspath(imageSupportData)
        base = name[:name.rfind(".")] if "." in name else name
        ident = base[base.rfind(os.sep) + 1:]
        try:
            with open(base + ".json", "rb") as f:
                return f.read(), ident
        except OSError as e:
            raise Error(Error.FILE_READ, "Could not read metadata file associated with image [%s.json] "
                        "with error [%s]" % (base, e), "UsgsAstroPlugin::loadImageSupportData")

    def _state_from_isd(self, imageSupportData, modelName, function):
        text, ident = self.loadImageSupportData(imageSupportData)
        need = C.c_size_t()
        name = None if not modelName else modelName.encode("ascii")
        code = lib.usgscsm_b200_state_from_isd(text, len(text), name, ident.encode("utf-8"), None, 0,
                                               C.byref(need))
        if code == _capi.ESTATE:
            raise Error(Error.SENSOR_MODEL_NOT_CONSTRUCTIBLE,
                        lib.usgscsm_b200_last_error().decode("utf-8", "replace"), function)
        check(code)
        buf = C.create_string_buffer(need.value)
        check(lib.usgscsm_b200_state_from_isd(text, len(text), name, ident.encode("utf-8"), buf,
                                              need.value, C.byref(need)))
        return buf.raw[:need.value].decode("utf-8")

    def constructModelFromISD(self, imageSupportData, modelName, warnings=None, device=None):
        """src/UsgsAstroPlugin.cpp:488-519 for the four un-projected models: ISD -> state
        (constructStateFromIsd, on the host) -> model on the device."""
        state = self._state_from_isd(imageSupportData, modelName, "UsgsAstroPlugin::constructModelFromISD")
        return self.constructModelFromState(state, warnings, device)

    def canModelBeConstructedFromISD(self, imageSupportData, modelName, warnings=None):
        """src/UsgsAstroPlugin.cpp:272-307: true when the model can be built; otherwise a warning
        with the reason is appended."""
        try:
            self.constructModelFromISD(imageSupportData, modelName, warnings)
            return True
        except (Error, RuntimeError) as e:
            if warnings is not None:
                warnings.append(Warning(Warning.UNKNOWN_WARNING,
                                        "Could not create model [%s] with error [%s]" % (modelName, e),
                                        "UsgsAstroPlugin::canModelBeConstructedFromISD()"))
            return False

    def convertISDToModelState(self, imageSupportData, modelName, warnings=None):
        """src/UsgsAstroPlugin.cpp:465-472: the state of the model built from the ISD (with what
        the model derived at load written back)."""
        return self.constructModelFromISD(imageSupportData, modelName, warnings).getModelState()

    def canISDBeConvertedToModelState(self, imageSupportData, modelName, warnings=None):
        try:
            self.convertISDToModelState(imageSupportData, modelName, warnings)
            return True
        except (Error, RuntimeError) as e:
            if warnings is not None:
                warnings.append(Warning(Warning.UNKNOWN_WARNING,
                                        "Could not create model [%s] state with error [%s]" % (modelName, e),
                                        "UsgsAstroPlugin::canISDBeConvertedToModelState()"))
            return False

    def getStateFromISD(self, imageSupportData):
        """src/UsgsAstroPlugin.cpp:459-463: convertISDToModelState with the ISD's own name_model."""
        return self.convertISDToModelState(imageSupportData, None)

    def constructStateFromIsd(self, imageSupportData, modelName=None):
        """<Model>::constructStateFromIsd: the state BEFORE the model derives anything.  Host only
        (no CUDA device needed)."""
        return self._state_from_isd(imageSupportData, modelName, "constructStateFromIsd")



# ---- georeferenced grids and DEMs (include/usgscsm_b200.h, "the callers either side of the path") ----
class GeoGrid(object):
    """Georeferencing of a raster: GDAL geotransform + map projection (longlat, eqc, sinu) +
    ellipsoid, i.e. m_geoTransform / m_projString of UsgsAstroProjectedSensorModel
    (src/UsgsAstroProjectedSensorModel.cpp:123-147)."""

    def __init__(self, c_grid):
        self.c = c_grid

    @staticmethod
    def fromProjString(projString, geoTransform, semiMajor=0.0, semiMinor=0.0):
        g = _capi.GeoGrid()
        gt = (C.c_double * 6)(*[float(v) for v in geoTransform])
        check(lib.usgscsm_b200_geogrid_from_proj_string(projString.encode("utf-8"), gt,
                                                        float(semiMajor), float(semiMinor), C.byref(g)))
        return GeoGrid(g)

    def imageToGround(self, line, samp, height=None, device=0):
        """UsgsAstroProjectedSensorModel::imageToGround for arrays: raster (line, samp) [+ height]
        -> body-fixed (x, y, z, status)."""
        if _is_torch(line):
            import torch
            n = line.numel()
            x, y, z = (torch.empty_like(line) for _ in range(3))
            st = torch.empty(n, dtype=torch.int32, device=line.device)
            hs = 0 if height is None or height.numel() == 1 else 1
            check(lib.usgscsm_b200_geogrid_image_to_ground_dev(
                C.byref(self.c), line.device.index, n, _ptr(line), _ptr(samp), _ptr(height), hs,
                _ptr(x), _ptr(y), _ptr(z), _ptr(st), torch.cuda.current_stream(line.device).cuda_stream))
            return x, y, z, st
        line = _host(line)
        n = line.size
        samp = _host(samp, n)
        h = None if height is None else _host(height)
        x, y, z = np.empty(n), np.empty(n), np.empty(n)
        st = np.empty(n, dtype=np.uint32)
        check(lib.usgscsm_b200_geogrid_image_to_ground(
            C.byref(self.c), device, n, _ptr(line), _ptr(samp), _ptr(h),
            0 if h is None or h.size == 1 else 1, _ptr(x), _ptr(y), _ptr(z), _ptr(st)))
        return x, y, z, st

    def groundToImage(self, x, y, z, device=0):
        """UsgsAstroProjectedSensorModel::groundToImage for arrays: body-fixed -> raster
        (line, samp, height above the ellipsoid, status)."""
        if _is_torch(x):
            import torch
            n = x.numel()
            line, samp, h = (torch.empty_like(x) for _ in range(3))
            st = torch.empty(n, dtype=torch.int32, device=x.device)
            check(lib.usgscsm_b200_geogrid_ground_to_image_dev(
                C.byref(self.c), x.device.index, n, _ptr(x), _ptr(y), _ptr(z), _ptr(line), _ptr(samp),
                _ptr(h), _ptr(st), torch.cuda.current_stream(x.device).cuda_stream))
            return line, samp, h, st
        x = _host(x)
        n = x.size
        y, z = _host(y, n), _host(z, n)
        line, samp, h = np.empty(n), np.empty(n), np.empty(n)
        st = np.empty(n, dtype=np.uint32)
        check(lib.usgscsm_b200_geogrid_ground_to_image(
            C.byref(self.c), device, n, _ptr(x), _ptr(y), _ptr(z), _ptr(line), _ptr(samp), _ptr(h),
            _ptr(st)))
        return line, samp, h, st


class Dem(object):
    """A DEM raster: a 2-D array of float64 / float32 / int16 samples (numpy = host entry points,
    torch CUDA tensor = device entry points) on a GeoGrid; height = raw * scale + offset."""

    def __init__(self, samples, grid, nodata=None, scale=1.0, offset=0.0):
        self.torch = _is_torch(samples)
        if self.torch:
            import torch
            types = {torch.float64: _capi.DEM_F64, torch.float32: _capi.DEM_F32, torch.int16: _capi.DEM_I16}
            if not (samples.is_cuda and samples.is_contiguous() and samples.dim() == 2 and samples.dtype in types):
                raise ValueError("DEM samples: contiguous 2-D CUDA tensor of float64 / float32 / int16")
            kind, rows, cols = types[samples.dtype], samples.shape[0], samples.shape[1]
        else:
            samples = np.ascontiguousarray(samples)
            types = {np.dtype("float64"): _capi.DEM_F64, np.dtype("float32"): _capi.DEM_F32,
                     np.dtype("int16"): _capi.DEM_I16}
            if samples.ndim != 2 or samples.dtype not in types:
                raise ValueError("DEM samples: 2-D array of float64 / float32 / int16")
            kind, rows, cols = types[samples.dtype], samples.shape[0], samples.shape[1]
        self.samples, self.grid = samples, grid
        self.c = _capi.Dem(_ptr(samples), kind, 0 if nodata is None else 1,
                           0.0 if nodata is None else float(nodata), float(scale), float(offset),
                           rows, cols, grid.c)

    @property
    def shape(self):
        return int(self.c.n_rows), int(self.c.n_cols)

    def _window(self, row0, col0, n_rows, n_cols):
        n_rows = self.c.n_rows - row0 if n_rows is None else n_rows
        n_cols = self.c.n_cols - col0 if n_cols is None else n_cols
        return int(row0), int(col0), int(n_rows), int(n_cols)

    def groundPoints(self, row0=0, col0=0, n_rows=None, n_cols=None, device=0):
        """Body-fixed ground points of the window's pixel centres, height = the sample."""
        row0, col0, n_rows, n_cols = self._window(row0, col0, n_rows, n_cols)
        n = n_rows * n_cols
        if self.torch:
            import torch
            dev = self.samples.device
            x, y, z = (torch.empty(n, dtype=torch.float64, device=dev) for _ in range(3))
            st = torch.empty(n, dtype=torch.int32, device=dev)
            check(lib.usgscsm_b200_dem_ground_points_dev(
                C.byref(self.c), dev.index, row0, col0, n_rows, n_cols, _ptr(x), _ptr(y), _ptr(z),
                _ptr(st), torch.cuda.current_stream(dev).cuda_stream))
            return x, y, z, st
        x, y, z = np.empty(n), np.empty(n), np.empty(n)
        st = np.empty(n, dtype=np.uint32)
        check(lib.usgscsm_b200_dem_ground_points(C.byref(self.c), device, row0, col0, n_rows, n_cols,
                                                 _ptr(x), _ptr(y), _ptr(z), _ptr(st)))
        return x, y, z, st


def _ground_to_image_dem_grid(self, dem, row0=0, col0=0, n_rows=None, n_cols=None,
                              desiredPrecision=0.001, achieved=False, status=True, out=None, stream=None):
    """groundToImage of every ground point of a DEM window, the points generated in the kernel
    (orthorectification's inner loop).  Returns (line, samp, achieved, status), row-major over the
    window."""
    row0, col0, n_rows, n_cols = dem._window(row0, col0, n_rows, n_cols)
    n = n_rows * n_cols
    if dem.torch:
        import torch
        dev = dem.samples.device
        if dev.index != self.device:
            raise ValueError("DEM on cuda:%s, the model lives on cuda:%d" % (dev.index, self.device))
        if out is None:
            line, samp = (torch.empty(n, dtype=torch.float64, device=dev) for _ in range(2))
            ach = torch.empty(n, dtype=torch.float64, device=dev) if achieved else None
            st = torch.empty(n, dtype=torch.int32, device=dev) if status else None
        else:
            line, samp, ach, st = out
        if stream is None:
            stream = torch.cuda.current_stream(dev).cuda_stream
        check(lib.usgscsm_b200_ground_to_image_dem_grid_dev(
            self._h, C.byref(dem.c), row0, col0, n_rows, n_cols, desiredPrecision, _ptr(line), _ptr(samp),
            _ptr(ach), _ptr(st), stream))
        return line, samp, ach, st
    if out is None:
        line, samp = np.empty(n), np.empty(n)
        ach = np.empty(n) if achieved else None
        st = np.empty(n, dtype=np.uint32) if status else None
    else:
        line, samp, ach, st = out
    check(lib.usgscsm_b200_ground_to_image_dem_grid(
        self._h, C.byref(dem.c), row0, col0, n_rows, n_cols, desiredPrecision, _ptr(line), _ptr(samp),
        _ptr(ach), _ptr(st)))
    return line, samp, ach, st


def _image_to_ground_dem(self, dem, line, samp, h0=0.0, tolerance=0.01, maxIterations=8,
                         desiredPrecision=0.001):
    """DEM-iterated ground intersection: imageToGround at the height of the DEM where the ray
    meets it.  Returns (x, y, z, height, status)."""
    if dem.torch:
        import torch
        n = line.numel()
        _dev_inputs(self, n, line, samp)
        x, y, z, h = (torch.empty_like(line) for _ in range(4))
        st = torch.empty(n, dtype=torch.int32, device=line.device)
        check(lib.usgscsm_b200_image_to_ground_dem_dev(
            self._h, C.byref(dem.c), n, _ptr(line), _ptr(samp), float(h0), float(tolerance),
            int(maxIterations), desiredPrecision, _ptr(x), _ptr(y), _ptr(z), _ptr(h), _ptr(st),
            torch.cuda.current_stream(line.device).cuda_stream))
        return x, y, z, h, st
    line = _host(line)
    n = line.size
    samp = _host(samp, n)
    x, y, z, h = np.empty(n), np.empty(n), np.empty(n), np.empty(n)
    st = np.empty(n, dtype=np.uint32)
    check(lib.usgscsm_b200_image_to_ground_dem(
        self._h, C.byref(dem.c), n, _ptr(line), _ptr(samp), float(h0), float(tolerance),
        int(maxIterations), desiredPrecision, _ptr(x), _ptr(y), _ptr(z), _ptr(h), _ptr(st)))
    return x, y, z, h, st


RasterGM.groundToImageDemGrid = _ground_to_image_dem_grid
RasterGM.imageToGroundDem = _image_to_ground_dem


# ---- many small images per call ----------------------------------------------------------------------
def _handles(models):
    return (C.c_void_p * len(models))(*[m._h for m in models])


def groundToImageImages(models, counts, x, y, z, desiredPrecision=0.001, achieved=False, status=True,
                        out=None, stream=None):
    """groundToImage of several images in one call: models[k] takes the next counts[k] points of
    the concatenated torch CUDA tensors; the images' kernels overlap on internal streams.
    Returns (line, samp, achieved, status)."""
    import torch
    n = int(sum(counts))
    _dev_inputs(models[0], n, x, y, z)
    if out is None:
        line, samp = torch.empty_like(x), torch.empty_like(x)
        ach = torch.empty_like(x) if achieved else None
        st = torch.empty(n, dtype=torch.int32, device=x.device) if status else None
    else:
        line, samp, ach, st = out
    if stream is None:
        stream = torch.cuda.current_stream(x.device).cuda_stream
    cnt = (C.c_int64 * len(models))(*[int(c) for c in counts])
    check(lib.usgscsm_b200_ground_to_image_images_dev(
        _handles(models), len(models), cnt, _ptr(x), _ptr(y), _ptr(z), desiredPrecision, _ptr(line),
        _ptr(samp), _ptr(ach), _ptr(st), stream))
    return line, samp, ach, st


def imageToGroundImages(models, counts, line, samp, height, desiredPrecision=0.001, achieved=False,
                        status=True, stream=None):
    """imageToGround of several images in one call (explicit pixel arrays, concatenated).
    Returns (x, y, z, achieved, status)."""
    import torch
    n = int(sum(counts))
    _dev_inputs(models[0], n, line, samp)
    if not _is_torch(height):
        height = torch.full((1,), float(height), dtype=torch.float64, device=line.device)
    stride = 0 if height.numel() == 1 else 1
    _dev_inputs(models[0], n if stride else 1, height)
    x, y, z = (torch.empty_like(line) for _ in range(3))
    ach = torch.empty_like(line) if achieved else None
    st = torch.empty(n, dtype=torch.int32, device=line.device) if status else None
    if stream is None:
        stream = torch.cuda.current_stream(line.device).cuda_stream
    cnt = (C.c_int64 * len(models))(*[int(c) for c in counts])
    check(lib.usgscsm_b200_image_to_ground_images_dev(
        _handles(models), len(models), cnt, _ptr(line), _ptr(samp), _ptr(height), stride,
        desiredPrecision, _ptr(x), _ptr(y), _ptr(z), _ptr(ach), _ptr(st), stream))
    return x, y, z, ach, st


def imageToGroundGridImages(models, line0, samp0, dline, dsamp, n_lines, n_samps, height,
                            desiredPrecision=0.001, achieved=False, status=True, out=None, stream=None):
    """Dense-grid imageToGround of several images in one call: image k is the n_lines[k] x
    n_samps[k] grid; `height`: torch CUDA tensor with 1 entry or one per point of all grids.
    Returns (x, y, z, achieved, status), the images' points one after the other."""
    import torch
    n = int(sum(int(a) * int(b) for a, b in zip(n_lines, n_samps)))
    stride = 0 if height.numel() == 1 else 1
    _dev_inputs(models[0], n if stride else 1, height)
    dev = height.device
    if out is None:
        x, y, z = (torch.empty(n, dtype=torch.float64, device=dev) for _ in range(3))
        ach = torch.empty(n, dtype=torch.float64, device=dev) if achieved else None
        st = torch.empty(n, dtype=torch.int32, device=dev) if status else None
    else:
        x, y, z, ach, st = out
    if stream is None:
        stream = torch.cuda.current_stream(dev).cuda_stream
    nl = (C.c_int64 * len(models))(*[int(v) for v in n_lines])
    ns = (C.c_int64 * len(models))(*[int(v) for v in n_samps])
    check(lib.usgscsm_b200_image_to_ground_grid_images_dev(
        _handles(models), len(models), float(line0), float(samp0), float(dline), float(dsamp), nl, ns,
        _ptr(height), stride, desiredPrecision, _ptr(x), _ptr(y), _ptr(z), _ptr(ach), _ptr(st), stream))
    return x, y, z, ach, st


# ---- the reference's remaining free functions (known-answer replay on the device functions) ---------
def _affine12(iTransS, iTransL, sampleOrigin, lineOrigin, sampleSumming, lineSumming, startingSample,
              startingLine):
    return np.array(list(iTransS) + list(iTransL) + [sampleOrigin, lineOrigin, sampleSumming, lineSumming,
                                                     startingSample, startingLine], dtype=np.float64)


def computeDistortedFocalPlaneCoordinates(line, sample, sampleOrigin, lineOrigin, sampleSumming, lineSumming,
                                          startingSample, startingLine, iTransS, iTransL, device=0):
    """src/Utilities.cpp:129-151, same argument order; arrays in, (x, y) arrays out."""
    line = _host(line)
    sample = _host(sample, line.size)
    a = _affine12(iTransS, iTransL, sampleOrigin, lineOrigin, sampleSumming, lineSumming, startingSample, startingLine)
    x, y = np.empty(line.size), np.empty(line.size)
    check(lib.usgscsm_b200_focal_plane_coordinates(device, _ptr(a), line.size, _ptr(line), _ptr(sample), _ptr(x), _ptr(y)))
    return x, y


def computePixel(distortedX, distortedY, sampleOrigin, lineOrigin, sampleSumming, lineSumming, startingSample,
                 startingLine, iTransS, iTransL, device=0):
    """src/Utilities.cpp:317-331, same argument order; returns (line, sample) arrays."""
    fx = _host(distortedX)
    fy = _host(distortedY, fx.size)
    a = _affine12(iTransS, iTransL, sampleOrigin, lineOrigin, sampleSumming, lineSumming, startingSample, startingLine)
    line, sample = np.empty(fx.size), np.empty(fx.size)
    check(lib.usgscsm_b200_compute_pixel(device, _ptr(a), fx.size, _ptr(fx), _ptr(fy), _ptr(line), _ptr(sample)))
    return line, sample


def createCameraLookVector(undistortedX, undistortedY, zDirection, focalLength, device=0):
    """src/Utilities.cpp:360-373; returns an (n, 3) array of unit look vectors."""
    ux = _host(undistortedX)
    uy = _host(undistortedY, ux.size)
    look = np.empty((ux.size, 3))
    check(lib.usgscsm_b200_camera_look_vector(device, float(zDirection), float(focalLength), ux.size, _ptr(ux),
                                              _ptr(uy), _ptr(look)))
    return look


def _jitter(remove, line, sample, lineJitterCoeffs, sampleJitterCoeffs, lineTimes, tolerance, maxIts, device):
    lj, sj, lt = _host(lineJitterCoeffs), _host(sampleJitterCoeffs), _host(lineTimes)
    if lj.size != sj.size or lt.size == 0:   # U:186-197: csm::Error(INDEX_OUT_OF_RANGE)
        raise Error(Error.INDEX_OUT_OF_RANGE, "Jitter coefficient arrays must be the same length / line times "
                                              "must not be empty.", "removeJitter")
    line = _host(line)
    sample = _host(sample, line.size)
    ol, os_ = np.empty(line.size), np.empty(line.size)
    check(lib.usgscsm_b200_jitter(device, remove, lj.size, _ptr(lj), _ptr(sj), lt.size, _ptr(lt), float(tolerance),
                                  int(maxIts), line.size, _ptr(line), _ptr(sample), _ptr(ol), _ptr(os_)))
    return ol, os_


def removeJitter(line, sample, lineJitterCoeffs, sampleJitterCoeffs, lineTimes, device=0):
    """src/Utilities.cpp:182-213; returns (dejitteredLine, dejitteredSample)."""
    return _jitter(1, line, sample, lineJitterCoeffs, sampleJitterCoeffs, lineTimes, 0.0, 0, device)


def addJitter(line, sample, tolerance, maxIts, lineJitterCoeffs, sampleJitterCoeffs, lineTimes, device=0):
    """src/Utilities.cpp:244-282; returns (jitteredLine, jitteredSample)."""
    return _jitter(0, line, sample, lineJitterCoeffs, sampleJitterCoeffs, lineTimes, tolerance, maxIts, device)


def _cubic(coeffs):
    c = [float(v) for v in coeffs]
    if not c:
        raise ValueError("Polynomial coeffs must be non-empty.")   # U:698-700 throws std::invalid_argument
    if len(c) > 4:
        raise ValueError("the engine evaluates polynomials of degree <= 3 (the SAR range polynomial)")
    return np.array(c + [0.0] * (4 - len(c)), dtype=np.float64)


def evaluatePolynomial(coeffs, x, device=0):
    """src/Utilities.cpp:696-710 (degree <= 3)."""
    c, x = _cubic(coeffs), _host(x)
    v = np.empty(x.size)
    check(lib.usgscsm_b200_polynomial(device, _ptr(c), x.size, _ptr(x), _ptr(v), None))
    return v


def evaluatePolynomialDerivative(coeffs, x, device=0):
    """src/Utilities.cpp:712-727 (degree <= 3)."""
    c, x = _cubic(coeffs), _host(x)
    v = np.empty(x.size)
    check(lib.usgscsm_b200_polynomial(device, _ptr(c), x.size, _ptr(x), None, _ptr(v)))
    return v


def polynomialRoot(coeffs, guess, threshold=1.0e-10, device=0):
    """src/Utilities.cpp:771-791 (maxIterations 30); returns (root, status): status has
    ST_ERROR_UNKNOWN where the reference throws std::invalid_argument."""
    c, g = _cubic(coeffs), _host(guess)
    r, st = np.empty(g.size), np.empty(g.size, dtype=np.uint32)
    check(lib.usgscsm_b200_polynomial_root(device, _ptr(c), float(threshold), g.size, _ptr(g), _ptr(r), _ptr(st)))
    return r, st


def brentRoot(lowerBound, upperBound, coeffs, epsilon=1e-10, device=0):
    """src/Utilities.cpp:545-622 on the polynomial `coeffs` (degree <= 3); returns (root, status)."""
    c, lo = _cubic(coeffs), _host(lowerBound)
    hi = _host(upperBound, lo.size)
    r, st = np.empty(lo.size), np.empty(lo.size, dtype=np.uint32)
    check(lib.usgscsm_b200_brent_root(device, _ptr(c), float(epsilon), lo.size, _ptr(lo), _ptr(hi), _ptr(r), _ptr(st)))
    return r, st

def stateToMsgpack(modelState):
    """The msgpack binary form of a model state, as the reference writes it
    (nlohmann::json::to_msgpack; bin/usgscsm_cam_test.cc --output-binary-model-state)."""
    raw = modelState if isinstance(modelState, bytes) else modelState.encode("utf-8")
    need = C.c_size_t()
    check(lib.usgscsm_b200_state_to_msgpack(raw, len(raw), None, 0, C.byref(need)))
    buf = C.create_string_buffer(need.value)
    check(lib.usgscsm_b200_state_to_msgpack(raw, len(raw), buf, need.value, C.byref(need)))
    return buf.raw


def stateToJson(modelState):
    """A model state in any accepted encoding (msgpack bytes, JSON text, .sup) as
    "<model name>\\n<JSON>" text, doubles written so that they parse back to the same bits."""
    raw = modelState if isinstance(modelState, bytes) else modelState.encode("utf-8")
    need = C.c_size_t()
    check(lib.usgscsm_b200_state_to_json(raw, len(raw), None, 0, C.byref(need)))
    buf = C.create_string_buffer(need.value)
    check(lib.usgscsm_b200_state_to_json(raw, len(raw), buf, need.value, C.byref(need)))
    return buf.raw[:need.value].decode("utf-8")


# ---- building blocks (include/usgscsm/Distortion.h, Utilities.h of the reference) ------------
def applyDistortion(ux, uy, opticalDistCoeffs, focalLength, distortionType,
                    desiredPrecision=1.0e-6, device=0):
    """applyDistortion (src/Distortion.cpp:441-775) for arrays -> (dx, dy, status)."""
    ux = _host(ux)
    uy = _host(uy, ux.size)
    c = _host(opticalDistCoeffs)
    dx, dy = np.empty(ux.size), np.empty(ux.size)
    st = np.empty(ux.size, dtype=np.uint32)
    check(lib.usgscsm_b200_apply_distortion(device, int(distortionType), _ptr(c), c.size,
                                            float(focalLength), float(desiredPrecision), ux.size,
                                            _ptr(ux), _ptr(uy), _ptr(dx), _ptr(dy), _ptr(st)))
    return dx, dy, st


def removeDistortion(dx, dy, opticalDistCoeffs, focalLength, distortionType, tolerance=1.0e-6,
                     device=0):
    """removeDistortion (src/Distortion.cpp:208-424) for arrays -> (ux, uy, status)."""
    dx = _host(dx)
    dy = _host(dy, dx.size)
    c = _host(opticalDistCoeffs)
    ux, uy = np.empty(dx.size), np.empty(dx.size)
    st = np.empty(dx.size, dtype=np.uint32)
    check(lib.usgscsm_b200_remove_distortion(device, int(distortionType), _ptr(c), c.size,
                                             float(focalLength), float(tolerance), dx.size,
                                             _ptr(dx), _ptr(dy), _ptr(ux), _ptr(uy), _ptr(st)))
    return ux, uy, st


def lagrangeInterp(numTime, valueArray, startTime, delTime, times, vectorLength, order, device=0):
    """lagrangeInterp (src/Utilities.cpp:409-513) at an array of times -> (n, vectorLength)."""
    v = _host(valueArray)
    t = _host(times)
    out = np.empty(t.size * int(vectorLength))
    code = lib.usgscsm_b200_lagrange_interp(device, int(numTime), _ptr(v), float(startTime),
                                            float(delTime), int(vectorLength), int(order), t.size,
                                            _ptr(t), _ptr(out))
    if code == _capi.EINVAL and numTime < 2:
        raise Error(Error.INDEX_OUT_OF_RANGE, "Not enough data to interpolate", "lagrangeInterp")
    check(code)
    return out.reshape(t.size, int(vectorLength))
