"""ctypes binding of the C-ABI in include/usgscsm_b200.h (libusgscsm_b200.so).

The library is the product: there is no Python or CPU implementation of the per-point
math behind it.  Importing this module fails loudly when the shared library has not been
built (``python -c "import __graft_entry__ as g; g.build()"`` or ``make -C usgscsm_b200``).
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# USGSCSM_B200_LIB: another build of the same library (A/B experiments)
LIB_PATH = os.environ.get("USGSCSM_B200_LIB") or os.path.join(_HERE, "libusgscsm_b200.so")

OK = 0
EINVAL, ESTATE, ENODEVICE, ECUDA, EUNSUPPORTED = -1, -2, -3, -4, -5

MODEL_LINE_SCANNER, MODEL_PUSH_FRAME, MODEL_FRAME, MODEL_SAR = 1, 2, 3, 4

ST_NO_INTERSECTION = 1 << 0
ST_PRECISION_NOT_MET = 1 << 1
ST_ITERATION_CAP = 1 << 2
ST_SEARCH_STALL = 1 << 3
ST_DISTORTION_NOT_CONVERGED = 1 << 4
ST_ERROR_ALGORITHM = 1 << 5
ST_ERROR_UNKNOWN = 1 << 6
ST_NONFINITE = 1 << 7
ST_OUT_OF_IMAGE = 1 << 8
ST_ERROR_OTHER = 1 << 9
ST_DEM_NODATA = 1 << 10
ST_DEM_OUTSIDE = 1 << 11
ST_EVALS_SHIFT = 16
ST_EVALS_MASK = 0xff << 16
ST_REFERENCE_MASK = (ST_NO_INTERSECTION | ST_PRECISION_NOT_MET | ST_ERROR_ALGORITHM |
                     ST_ERROR_UNKNOWN | ST_NONFINITE | ST_OUT_OF_IMAGE | ST_ERROR_OTHER)

_dp = C.c_void_p  # double* / uint32_t* / stream, passed as raw addresses
_h = C.c_void_p   # usgscsm_b200_model*

PROJ_LONGLAT, PROJ_EQC, PROJ_SINU = 0, 1, 2
DEM_F64, DEM_F32, DEM_I16 = 0, 1, 2


class GeoGrid(C.Structure):
    """usgscsm_b200_geogrid"""
    _fields_ = [("geotransform", C.c_double * 6), ("projection", C.c_int), ("reserved", C.c_int),
                ("lat_ts_deg", C.c_double), ("lat_0_deg", C.c_double), ("lon_0_deg", C.c_double),
                ("x_0", C.c_double), ("y_0", C.c_double),
                ("proj_semi_major", C.c_double), ("proj_semi_minor", C.c_double),
                ("semi_major", C.c_double), ("semi_minor", C.c_double)]


class Dem(C.Structure):
    """usgscsm_b200_dem"""
    _fields_ = [("samples", C.c_void_p), ("sample_type", C.c_int), ("has_nodata", C.c_int),
                ("nodata", C.c_double), ("scale", C.c_double), ("offset", C.c_double),
                ("n_rows", C.c_int64), ("n_cols", C.c_int64), ("grid", GeoGrid)]


_gp = C.POINTER(GeoGrid)
_demp = C.POINTER(Dem)

# name -> (restype, argtypes); the order follows include/usgscsm_b200.h
SIGNATURES = {
    "usgscsm_b200_abi_version": (C.c_int, []),
    "usgscsm_b200_last_error": (C.c_char_p, []),
    "usgscsm_b200_device_count": (C.c_int, []),
    "usgscsm_b200_model_from_state": (C.c_int, [C.c_char_p, C.c_size_t, C.c_int, C.POINTER(_h)]),
    "usgscsm_b200_state_from_isd": (
        C.c_int, [C.c_char_p, C.c_size_t, C.c_char_p, C.c_char_p, C.c_char_p, C.c_size_t,
                  C.POINTER(C.c_size_t)]),
    "usgscsm_b200_model_from_isd": (
        C.c_int, [C.c_char_p, C.c_size_t, C.c_char_p, C.c_char_p, C.c_int, C.POINTER(_h)]),
    "usgscsm_b200_model_name_from_state": (C.c_int, [C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t]),
    "usgscsm_b200_state_to_msgpack": (
        C.c_int, [C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]),
    "usgscsm_b200_state_to_json": (
        C.c_int, [C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]),
    "usgscsm_b200_model_free": (None, [_h]),
    "usgscsm_b200_model_kind": (C.c_int, [_h]),
    "usgscsm_b200_model_device": (C.c_int, [_h]),
    "usgscsm_b200_model_image_size": (C.c_int, [_h, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "usgscsm_b200_model_num_parameters": (C.c_int, [_h]),
    "usgscsm_b200_model_get_parameter": (C.c_int, [_h, C.c_int, C.POINTER(C.c_double)]),
    "usgscsm_b200_model_set_parameter": (C.c_int, [_h, C.c_int, C.c_double]),
    "usgscsm_b200_model_set_ellipsoid": (C.c_int, [_h, C.c_double, C.c_double]),
    "usgscsm_b200_model_get_ellipsoid": (C.c_int, [_h, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "usgscsm_b200_model_ls_derived": (C.c_int, [_h, C.POINTER(C.c_double)]),
    "usgscsm_b200_model_ls_set_projective": (C.c_int, [_h, C.c_int, C.POINTER(C.c_double)]),
    "usgscsm_b200_sensor_state_at_time": (C.c_int, [_h, C.c_int64] + [_dp] * 8),
    "usgscsm_b200_sensor_state_at_time_dev": (C.c_int, [_h, C.c_int64] + [_dp] * 9),
    "usgscsm_b200_measure_fp64": (
        C.c_int, [C.c_int, C.c_int, C.c_double, C.POINTER(C.c_double), C.POINTER(C.c_double)]),
    "usgscsm_b200_ground_to_image": (
        C.c_int, [_h, C.c_int64, _dp, _dp, _dp, C.c_double, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_ground_to_image_dev": (
        C.c_int, [_h, C.c_int64, _dp, _dp, _dp, C.c_double, _dp, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_image_to_ground": (
        C.c_int, [_h, C.c_int64, _dp, _dp, _dp, C.c_int, C.c_double, _dp, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_image_to_ground_dev": (
        C.c_int,
        [_h, C.c_int64, _dp, _dp, _dp, C.c_int, C.c_double, _dp, _dp, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_ground_to_image_aos": (
        C.c_int, [_h, C.c_int64, _dp, C.c_double, _dp, _dp, _dp]),
    "usgscsm_b200_image_to_ground_aos": (
        C.c_int, [_h, C.c_int64, _dp, _dp, C.c_int, C.c_double, _dp, _dp, _dp]),
    "usgscsm_b200_image_to_remote_imaging_locus": (
        C.c_int, [_h, C.c_int64, _dp, _dp, C.c_double, _dp, _dp, _dp, _dp, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_image_to_remote_imaging_locus_dev": (
        C.c_int, [_h, C.c_int64, _dp, _dp, C.c_double, _dp, _dp, _dp, _dp, _dp, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_image_to_proximate_imaging_locus": (
        C.c_int, [_h, C.c_int64, _dp, _dp, _dp, _dp, _dp, C.c_double] + [_dp] * 8),
    "usgscsm_b200_image_to_proximate_imaging_locus_dev": (
        C.c_int, [_h, C.c_int64, _dp, _dp, _dp, _dp, _dp, C.c_double] + [_dp] * 9),
    "usgscsm_b200_ground_partials": (C.c_int, [_h, C.c_int64, _dp, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_ground_partials_dev": (C.c_int, [_h, C.c_int64, _dp, _dp, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_sensor_partials": (
        C.c_int, [_h, C.c_int, C.c_int64, _dp, _dp, _dp, _dp, _dp, C.c_double, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_sensor_partials_dev": (
        C.c_int, [_h, C.c_int, C.c_int64, _dp, _dp, _dp, _dp, _dp, C.c_double, _dp, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_all_sensor_partials": (
        C.c_int, [_h, C.c_int, C.POINTER(C.c_int), C.c_int64, _dp, _dp, _dp, _dp, _dp, C.c_double,
                  _dp, _dp, _dp]),
    "usgscsm_b200_all_sensor_partials_dev": (
        C.c_int, [_h, C.c_int, C.POINTER(C.c_int), C.c_int64, _dp, _dp, _dp, _dp, _dp, C.c_double,
                  _dp, _dp, _dp, _dp]),
    "usgscsm_b200_model_get_state": (C.c_int, [_h, C.c_char_p, C.c_size_t, C.POINTER(C.c_size_t)]),
    "usgscsm_b200_image_sensor_state": (C.c_int, [_h, C.c_int64] + [_dp] * 10),
    "usgscsm_b200_image_sensor_state_dev": (C.c_int, [_h, C.c_int64] + [_dp] * 11),
    "usgscsm_b200_illumination_direction": (C.c_int, [_h, C.c_int64] + [_dp] * 7),
    "usgscsm_b200_illumination_direction_dev": (C.c_int, [_h, C.c_int64] + [_dp] * 8),
    "usgscsm_b200_ground_to_image_covar": (
        C.c_int, [_h, C.c_int64, _dp, _dp, _dp, _dp, C.c_double, _dp, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_ground_to_image_covar_dev": (
        C.c_int, [_h, C.c_int64, _dp, _dp, _dp, _dp, C.c_double, _dp, _dp, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_image_to_ground_covar": (
        C.c_int, [_h, C.c_int64, _dp, _dp, _dp, _dp, C.c_int, _dp, C.c_int, C.c_double, _dp, _dp,
                  _dp, _dp, _dp, _dp]),
    "usgscsm_b200_image_to_ground_covar_dev": (
        C.c_int, [_h, C.c_int64, _dp, _dp, _dp, _dp, C.c_int, _dp, C.c_int, C.c_double, _dp, _dp,
                  _dp, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_model_get_parameter_covariance": (
        C.c_int, [_h, C.c_int, C.c_int, C.POINTER(C.c_double)]),
    "usgscsm_b200_model_set_parameter_covariance": (C.c_int, [_h, C.c_int, C.c_int, C.c_double]),
    "usgscsm_b200_ground_to_image_multi": (
        C.c_int, [C.POINTER(_h), C.c_int, C.c_int64, _dp, _dp, _dp, C.c_double, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_image_to_ground_multi": (
        C.c_int, [C.POINTER(_h), C.c_int, C.c_int64, _dp, _dp, _dp, C.c_int, C.c_double, _dp, _dp,
                  _dp, _dp, _dp]),
    "usgscsm_b200_ground_to_image_images_dev": (
        C.c_int, [C.POINTER(_h), C.c_int, C.POINTER(C.c_int64), _dp, _dp, _dp, C.c_double, _dp, _dp, _dp,
                  _dp, _dp]),
    "usgscsm_b200_image_to_ground_images_dev": (
        C.c_int, [C.POINTER(_h), C.c_int, C.POINTER(C.c_int64), _dp, _dp, _dp, C.c_int, C.c_double, _dp,
                  _dp, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_image_to_ground_grid_images_dev": (
        C.c_int, [C.POINTER(_h), C.c_int, C.c_double, C.c_double, C.c_double, C.c_double,
                  C.POINTER(C.c_int64), C.POINTER(C.c_int64), _dp, C.c_int, C.c_double, _dp, _dp, _dp, _dp,
                  _dp, _dp]),
    "usgscsm_b200_image_to_ground_grid_dev": (
        C.c_int, [_h, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int64, C.c_int64, _dp,
                  C.c_int, C.c_double, _dp, _dp, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_image_to_ground_grid": (
        C.c_int, [_h, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int64, C.c_int64, _dp,
                  C.c_int, C.c_double, _dp, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_image_to_ground_grid_static_dev": (
        C.c_int, [_h, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int64, C.c_int64, _dp,
                  C.c_int, C.c_double, _dp, _dp, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_geogrid_from_proj_string": (
        C.c_int, [C.c_char_p, C.POINTER(C.c_double), C.c_double, C.c_double, _gp]),
    "usgscsm_b200_geogrid_image_to_ground": (
        C.c_int, [_gp, C.c_int, C.c_int64, _dp, _dp, _dp, C.c_int, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_geogrid_image_to_ground_dev": (
        C.c_int, [_gp, C.c_int, C.c_int64, _dp, _dp, _dp, C.c_int, _dp, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_geogrid_ground_to_image": (
        C.c_int, [_gp, C.c_int, C.c_int64, _dp, _dp, _dp, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_geogrid_ground_to_image_dev": (
        C.c_int, [_gp, C.c_int, C.c_int64, _dp, _dp, _dp, _dp, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_dem_ground_points": (
        C.c_int, [_demp, C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_int64, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_dem_ground_points_dev": (
        C.c_int, [_demp, C.c_int, C.c_int64, C.c_int64, C.c_int64, C.c_int64, _dp, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_ground_to_image_dem_grid": (
        C.c_int, [_h, _demp, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_double, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_ground_to_image_dem_grid_dev": (
        C.c_int, [_h, _demp, C.c_int64, C.c_int64, C.c_int64, C.c_int64, C.c_double, _dp, _dp, _dp, _dp,
                  _dp]),
    "usgscsm_b200_image_to_ground_dem": (
        C.c_int, [_h, _demp, C.c_int64, _dp, _dp, C.c_double, C.c_double, C.c_int, C.c_double,
                  _dp, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_image_to_ground_dem_dev": (
        C.c_int, [_h, _demp, C.c_int64, _dp, _dp, C.c_double, C.c_double, C.c_int, C.c_double,
                  _dp, _dp, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_apply_distortion": (
        C.c_int, [C.c_int, C.c_int, _dp, C.c_int, C.c_double, C.c_double, C.c_int64, _dp, _dp, _dp,
                  _dp, _dp]),
    "usgscsm_b200_remove_distortion": (
        C.c_int, [C.c_int, C.c_int, _dp, C.c_int, C.c_double, C.c_double, C.c_int64, _dp, _dp, _dp,
                  _dp, _dp]),
    "usgscsm_b200_lagrange_interp": (
        C.c_int, [C.c_int, C.c_int, _dp, C.c_double, C.c_double, C.c_int, C.c_int, C.c_int64, _dp,
                  _dp]),
    "usgscsm_b200_focal_plane_coordinates": (C.c_int, [C.c_int, _dp, C.c_int64, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_compute_pixel": (C.c_int, [C.c_int, _dp, C.c_int64, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_camera_look_vector": (
        C.c_int, [C.c_int, C.c_double, C.c_double, C.c_int64, _dp, _dp, _dp]),
    "usgscsm_b200_jitter": (
        C.c_int, [C.c_int, C.c_int, C.c_int, _dp, _dp, C.c_int, _dp, C.c_double, C.c_int, C.c_int64,
                  _dp, _dp, _dp, _dp]),
    "usgscsm_b200_polynomial": (C.c_int, [C.c_int, _dp, C.c_int64, _dp, _dp, _dp]),
    "usgscsm_b200_polynomial_root": (C.c_int, [C.c_int, _dp, C.c_double, C.c_int64, _dp, _dp, _dp]),
    "usgscsm_b200_brent_root": (C.c_int, [C.c_int, _dp, C.c_double, C.c_int64, _dp, _dp, _dp, _dp]),
    "usgscsm_b200_selftest_division": (C.c_int, [C.c_int, C.c_int64, C.c_uint64, C.POINTER(C.c_uint64)]),
    "usgscsm_b200_kernel_launches": (C.c_int64, []),
}


class EngineError(RuntimeError):
    """A C-ABI call returned a negative USGSCSM_B200_E* code."""

    def __init__(self, code, message):
        super().__init__("usgscsm_b200 error %d: %s" % (code, message))
        self.code = code


def _load():
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            "usgscsm_b200: %s is missing -- build the CUDA library first (there is no CPU "
            "fallback): python -c 'import __graft_entry__ as g; g.build()'" % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the library does not export it
        fn.restype = res
        fn.argtypes = args
    if lib.usgscsm_b200_abi_version() != 1:
        raise ImportError("usgscsm_b200: ABI version mismatch")
    return lib


lib = _load()


def check(code):
    if code != OK:
        raise EngineError(code, lib.usgscsm_b200_last_error().decode("utf-8", "replace"))
