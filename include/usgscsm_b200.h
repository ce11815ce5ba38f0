/* usgscsm_b200.h -- C-ABI of the B200-native batched sensor-model engine.
 *
 * This is the drop-in boundary for the ONE data-parallel hot path of DOI-USGS/usgscsm:
 * per-point csm::RasterGM::groundToImage / imageToGround of the USGS Astro line-scan,
 * push-frame, frame and SAR sensor models, evaluated for whole batches of points by
 * hand-written sm_100a CUDA kernels (one thread per point, FP64).
 *
 * Every entry point names the reference interface it replaces (paths relative to the
 * reference repository root).  Plain pointers and sizes only: no C++ or torch types.
 *
 * Conventions
 *   - all functions return 0 on success or a negative USGSCSM_B200_E* code; the message of
 *     the last failure on the calling thread is returned by usgscsm_b200_last_error();
 *   - per-point failures never fail the call: they are reported in the `status` words
 *     (bit-field below), which restate the reference's csm::Warning list and thrown
 *     csm::Error of the scalar call for that point;
 *   - the library has NO CPU implementation of the per-point math.  If no CUDA device is
 *     usable every compute entry point fails with USGSCSM_B200_ENODEVICE.
 */
#ifndef USGSCSM_B200_H_
#define USGSCSM_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define USGSCSM_B200_ABI_VERSION 1

/* every entry point below is exported from libusgscsm_b200.so */
#if defined(__GNUC__)
#define USGSCSM_B200_API __attribute__((visibility("default")))
#else
#define USGSCSM_B200_API
#endif

/* ---- error codes (function return values) ------------------------------------------ */
#define USGSCSM_B200_OK 0
#define USGSCSM_B200_EINVAL -1    /* bad argument                                        */
#define USGSCSM_B200_ESTATE -2    /* model state could not be parsed / is inconsistent   */
#define USGSCSM_B200_ENODEVICE -3 /* no usable CUDA device                               */
#define USGSCSM_B200_ECUDA -4     /* a CUDA runtime call failed                          */
#define USGSCSM_B200_EUNSUPPORTED -5

/* ---- model kinds: csm model names of src/UsgsAstroPlugin.cpp:62-69 ------------------ */
#define USGSCSM_B200_MODEL_LINE_SCANNER 1 /* USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL */
#define USGSCSM_B200_MODEL_PUSH_FRAME 2   /* USGS_ASTRO_PUSH_FRAME_SENSOR_MODEL   */
#define USGSCSM_B200_MODEL_FRAME 3        /* USGS_ASTRO_FRAME_SENSOR_MODEL        */
#define USGSCSM_B200_MODEL_SAR 4          /* USGS_ASTRO_SAR_SENSOR_MODEL          */

/* ---- per-point status bits ----------------------------------------------------------
 * Bits 0,1 restate csm::Warning::WarningType entries pushed by the scalar call
 * (csm/Warning.h:39-47); bits 5,6,9 restate the csm::Error the scalar call would have
 * thrown (csm/Error.h:40-48) -- the outputs of such a point are NaN; bits 7,8 are derived
 * the way a caller of the reference derives them; bits 2-4 are diagnostics only this
 * engine can see (the reference does not report them). */
#define USGSCSM_B200_ST_NO_INTERSECTION (1u << 0)   /* Warning::NO_INTERSECTION (LS:2433-2441)   */
#define USGSCSM_B200_ST_PRECISION_NOT_MET (1u << 1) /* Warning::PRECISION_NOT_MET (LS:882-887)   */
#define USGSCSM_B200_ST_ITERATION_CAP (1u << 2)     /* root search hit its iteration cap         */
#define USGSCSM_B200_ST_SEARCH_STALL (1u << 3)      /* secant stopped on equal residuals         */
#define USGSCSM_B200_ST_DISTORTION_NOT_CONVERGED (1u << 4)
#define USGSCSM_B200_ST_ERROR_ALGORITHM (1u << 5)   /* Error::ALGORITHM  (Distortion.cpp:357-365) */
#define USGSCSM_B200_ST_ERROR_UNKNOWN (1u << 6)     /* Error::UNKNOWN_ERROR (SAR:569-575)        */
#define USGSCSM_B200_ST_NONFINITE (1u << 7)         /* an output coordinate is NaN/Inf           */
#define USGSCSM_B200_ST_OUT_OF_IMAGE (1u << 8)      /* g2i result outside getValidImageRange()   */
#define USGSCSM_B200_ST_ERROR_OTHER (1u << 9)
/* DEM entry points only (no counterpart in the reference: its callers own the DEM) */
#define USGSCSM_B200_ST_DEM_NODATA (1u << 10)  /* the DEM sample (or a bilinear neighbour) is nodata */
#define USGSCSM_B200_ST_DEM_OUTSIDE (1u << 11) /* the ray left the DEM raster                        */
/* Bits 16-23 (diagnostic): how many times the root search of groundToImage evaluated its
 * function for this point (line-scan: detector views of the secant search; push-frame: frame
 * distances; SAR: Doppler evaluations of Brent's method).  bench.py derives the algorithmic
 * flop count of a pass from it. */
#define USGSCSM_B200_ST_EVALS_SHIFT 16
#define USGSCSM_B200_ST_EVALS_MASK (0xffu << USGSCSM_B200_ST_EVALS_SHIFT)
/* the bits the reference itself makes observable (compared bit-exactly in the tests) */
#define USGSCSM_B200_ST_REFERENCE_MASK \
  (USGSCSM_B200_ST_NO_INTERSECTION | USGSCSM_B200_ST_PRECISION_NOT_MET | \
   USGSCSM_B200_ST_ERROR_ALGORITHM | USGSCSM_B200_ST_ERROR_UNKNOWN |     \
   USGSCSM_B200_ST_NONFINITE | USGSCSM_B200_ST_OUT_OF_IMAGE | USGSCSM_B200_ST_ERROR_OTHER)

typedef struct usgscsm_b200_model usgscsm_b200_model;

/* ---- library ------------------------------------------------------------------------ */
USGSCSM_B200_API int usgscsm_b200_abi_version(void);
USGSCSM_B200_API const char *usgscsm_b200_last_error(void);
/* Number of CUDA devices the engine can use (0 when there is none). */
USGSCSM_B200_API int usgscsm_b200_device_count(void);

/* ---- model construction ---------------------------------------------------------------
 * Replaces UsgsAstroPlugin::constructModelFromState (src/UsgsAstroPlugin.cpp:535-555) +
 * <Model>::populateModel (LS: src/UsgsAstroLsSensorModel.cpp:141-293): `state` is a model
 * state string as written by getModelState() -- "<MODEL_NAME>\n{json}", bare JSON, the
 * body of a GXP .sup file (non-printable bytes are treated as newlines and the text is cut
 * to the outermost braces, src/Utilities.cpp:2489-2500) -- or the msgpack binary form of the
 * state (nlohmann::json::to_msgpack; recognised by its first byte like isMsgpack,
 * src/Utilities.cpp:2511-2530; bin/usgscsm_cam_test.cc:157-170).  The packed device state is
 * uploaded to CUDA device `device`.  Derived quantities the reference computes at load time
 * with its own per-point calls (updateState LS:649-735, createProjectiveApproximation
 * LS:2745-2803) are computed here by running the batch kernels on those points. */
USGSCSM_B200_API int usgscsm_b200_model_from_state(const char *state, size_t len, int device,
                                  usgscsm_b200_model **out);
/* ---- ISD -> model state ---------------------------------------------
 * Replaces <Model>::constructStateFromIsd (LS: src/UsgsAstroLsSensorModel.cpp:2818-3152, PF:
 * src/UsgsAstroPushFrameSensorModel.cpp:2114-2463, FR: src/UsgsAstroFrameSensorModel.cpp:1660-1903,
 * SAR: src/UsgsAstroSarSensorModel.cpp:52-230) and the ALE calls inside them: `isd` is the JSON
 * text of an ALE-format ISD (what UsgsAstroPlugin::loadImageSupportData reads from the image's
 * sidecar file, src/UsgsAstroPlugin.cpp:324-353); image_id (nullable) is stored as
 * m_imageIdentifier, as loadImageSupportData does with the image's base name; model_name
 * (nullable / "") defaults to the ISD's own "name_model" (UsgsAstroPlugin::getStateFromISD,
 * src/UsgsAstroPlugin.cpp:459-463).  Writes "<model name>\n<JSON>", the state BEFORE the model
 * derives its reference point / gsd (those are derived when the state is loaded, as in the
 * reference).  Host only: no device needed.  Call with out = NULL to learn the size.
 * USGSCSM_B200_ESTATE: the ISD cannot be converted (csm::Error SENSOR_MODEL_NOT_CONSTRUCTIBLE /
 * ISD_NOT_SUPPORTED in the reference); the message names the missing key. */
USGSCSM_B200_API int usgscsm_b200_state_from_isd(const char *isd, size_t len, const char *model_name,
                                const char *image_id, char *out, size_t capacity,
                                size_t *needed);
/* UsgsAstroPlugin::constructModelFromISD (src/UsgsAstroPlugin.cpp:488-519) for the four
 * un-projected models: usgscsm_b200_state_from_isd, then usgscsm_b200_model_from_state. */
USGSCSM_B200_API int usgscsm_b200_model_from_isd(const char *isd, size_t len, const char *model_name,
                                const char *image_id, int device, usgscsm_b200_model **out);

/* UsgsAstroPlugin::getModelNameFromModelState (src/UsgsAstroPlugin.cpp:176-218): the value of
 * the "m_modelName" key, written NUL-terminated into name[cap].  Host only (no device needed).
 * USGSCSM_B200_ESTATE: no such key / unparsable; USGSCSM_B200_EUNSUPPORTED: not one of the
 * four model names above. */
USGSCSM_B200_API int usgscsm_b200_model_name_from_state(const char *state, size_t len, char *name,
                                       size_t cap);
/* getModelState (LS:345-483, PF:346-455, FR:1241-1300, SAR:395-455): "<model name>\n<JSON>" of
 * the state the model was built from, with what the caller may have changed written back:
 * m_currentParameterValue, m_covariance, m_majorAxis / m_minorAxis.  Doubles are written with
 * 17 significant digits, so a model built from this text computes the same bits.  Call with
 * out = NULL to learn the size (*needed, without a terminator). */
USGSCSM_B200_API int usgscsm_b200_model_get_state(const usgscsm_b200_model *m, char *out, size_t capacity,
                                 size_t *needed);

/* A model state (any accepted encoding) -> its msgpack binary form, byte for byte what the
 * reference writes with nlohmann::json::to_msgpack (bin/usgscsm_cam_test.cc
 * --output-binary-model-state).  *needed receives the size; bytes are written when out holds
 * at least that many (call with out = NULL to size the buffer).  Host only. */
USGSCSM_B200_API int usgscsm_b200_state_to_msgpack(const char *state, size_t len, unsigned char *out,
                                  size_t cap, size_t *needed);
/* The other direction: a state in any accepted encoding (msgpack, JSON text, .sup) ->
 * "<model name>\n<JSON>", the text getModelState writes (json::from_msgpack + dump in the
 * reference, bin/usgscsm_cam_test.cc:157-170).  Doubles with 17 significant digits: the text
 * parses back to the same bits.  No device needed. */
USGSCSM_B200_API int usgscsm_b200_state_to_json(const char *state, size_t len, char *out, size_t capacity,
                               size_t *needed);
/* csm::Model destructor. */
USGSCSM_B200_API void usgscsm_b200_model_free(usgscsm_b200_model *m);
/* USGSCSM_B200_MODEL_* */
USGSCSM_B200_API int usgscsm_b200_model_kind(const usgscsm_b200_model *m);
USGSCSM_B200_API int usgscsm_b200_model_device(const usgscsm_b200_model *m);
/* csm::RasterGM::getImageSize (LS:1845-1852). */
USGSCSM_B200_API int usgscsm_b200_model_image_size(const usgscsm_b200_model *m, int *lines, int *samps);
/* csm::GeometricModel::getNumParameters / getParameterValue / setParameterValue
 * (LS:1585-1620).  Setting a parameter re-packs the device state. */
USGSCSM_B200_API int usgscsm_b200_model_num_parameters(const usgscsm_b200_model *m);
USGSCSM_B200_API int usgscsm_b200_model_get_parameter(const usgscsm_b200_model *m, int index, double *value);
USGSCSM_B200_API int usgscsm_b200_model_set_parameter(usgscsm_b200_model *m, int index, double value);
/* csm::SettableEllipsoid::setEllipsoid / getEllipsoid (LS:2116-2119, PF:1663-1666, FR:2235-2240,
 * SAR:1884-1887): replaces the two radii; as in the reference nothing else is re-derived. */
USGSCSM_B200_API int usgscsm_b200_model_set_ellipsoid(usgscsm_b200_model *m, double semi_major,
                                     double semi_minor);
USGSCSM_B200_API int usgscsm_b200_model_get_ellipsoid(const usgscsm_b200_model *m, double *semi_major,
                                     double *semi_minor);
/* Host-derived inputs of the line-scan per-point math:
 * out[0..3] = gsd, flyingHeight, halfSwath, halfTime; out[4] = useApproxInitTrans;
 * out[5..18] = the 14 projective-seed coefficients (LS:2688-2736); out[19..21] = reference
 * point.  Only meaningful for the line-scan model. */
USGSCSM_B200_API int usgscsm_b200_model_ls_derived(const usgscsm_b200_model *m, double *out22);
/* Override the projective seed (m_useApproxInitTrans, m_projTransCoeffs). */
USGSCSM_B200_API int usgscsm_b200_model_ls_set_projective(usgscsm_b200_model *m, int use,
                                         const double *coeffs14);

/* ---- per-point math, batched ------------------------------------------------------------
 * groundToImage: replaces csm::RasterGM::groundToImage(const EcefCoord&, double
 * desiredPrecision, double* achievedPrecision, WarningList*) (csm/RasterGM.h;
 * LS:757-890, PF:715-778, FR:137-232, SAR:541-593) called once per point.
 *   x,y,z     body-fixed ground coordinates, metres, n each (SoA)
 *   prec      desiredPrecision (reference default 0.001)
 *   line,samp outputs, n each
 *   achieved  nullable; the value the scalar call writes to *achievedPrecision
 *   status    nullable; USGSCSM_B200_ST_* bits
 * The plain flavour takes HOST pointers and performs the host<->device copies itself
 * (chunked and overlapped with the kernels); the _dev flavour takes DEVICE pointers on
 * the model's device and enqueues on `stream` (a cudaStream_t; NULL = default stream)
 * without synchronising. */
USGSCSM_B200_API int usgscsm_b200_ground_to_image(usgscsm_b200_model *m, int64_t n, const double *x,
                                 const double *y, const double *z, double prec, double *line,
                                 double *samp, double *achieved, uint32_t *status);
USGSCSM_B200_API int usgscsm_b200_ground_to_image_dev(usgscsm_b200_model *m, int64_t n, const double *x,
                                     const double *y, const double *z, double prec,
                                     double *line, double *samp, double *achieved,
                                     uint32_t *status, void *stream);

/* imageToGround: replaces csm::RasterGM::imageToGround(const ImageCoord&, double height,
 * double desiredPrecision, double* achievedPrecision, WarningList*) (LS:980-1009,
 * PF:842-870, FR:280-372, SAR:775-861).
 *   height         heights above the ellipsoid, metres; n entries when height_stride = 1,
 *                  one entry when height_stride = 0 */
USGSCSM_B200_API int usgscsm_b200_image_to_ground(usgscsm_b200_model *m, int64_t n, const double *line,
                                 const double *samp, const double *height, int height_stride,
                                 double prec, double *x, double *y, double *z, double *achieved,
                                 uint32_t *status);
USGSCSM_B200_API int usgscsm_b200_image_to_ground_dev(usgscsm_b200_model *m, int64_t n, const double *line,
                                     const double *samp, const double *height,
                                     int height_stride, double prec, double *x, double *y,
                                     double *z, double *achieved, uint32_t *status,
                                     void *stream);

/* AoS flavours of the two host-pointer calls, for callers that hold arrays of the CSM value
 * types (csm/csm.h:212-308): xyz = n csm::EcefCoord {x,y,z} (3n doubles), line_samp =
 * n csm::ImageCoord {line,samp} (2n doubles).  Split / joined on the device. */
USGSCSM_B200_API int usgscsm_b200_ground_to_image_aos(usgscsm_b200_model *m, int64_t n, const double *xyz,
                                     double prec, double *line_samp, double *achieved,
                                     uint32_t *status);
USGSCSM_B200_API int usgscsm_b200_image_to_ground_aos(usgscsm_b200_model *m, int64_t n,
                                     const double *line_samp, const double *height,
                                     int height_stride, double prec, double *xyz,
                                     double *achieved, uint32_t *status);

/* imageToRemoteImagingLocus: replaces csm::RasterGM::imageToRemoteImagingLocus(const ImageCoord&,
 * double desiredPrecision, double* achievedPrecision, WarningList*) (LS:1254-1282, PF:1060-1082,
 * FR:488-543, SAR:1009-1030) called once per pixel -- the call ray-tracing callers (DEM
 * intersection) make.  Outputs SoA: locus point (px,py,pz) and direction (dx,dy,dz) of
 * csm::EcefLocus (csm/csm.h:420-445).  Line-scan / push-frame: point = sensor position,
 * direction = the look vector as the reference returns it; frame: point = parameters 0..2,
 * unit direction; SAR: point = imageToGround at height 0, unit direction from the sensor.
 * achieved (nullable): 0 for the optical models, the imageToGround value for SAR. */
USGSCSM_B200_API int usgscsm_b200_image_to_remote_imaging_locus(usgscsm_b200_model *m, int64_t n,
                                               const double *line, const double *samp,
                                               double prec, double *px, double *py, double *pz,
                                               double *dx, double *dy, double *dz,
                                               double *achieved, uint32_t *status);
USGSCSM_B200_API int usgscsm_b200_image_to_remote_imaging_locus_dev(usgscsm_b200_model *m, int64_t n,
                                                   const double *line, const double *samp,
                                                   double prec, double *px, double *py,
                                                   double *pz, double *dx, double *dy, double *dz,
                                                   double *achieved, uint32_t *status,
                                                   void *stream);

/* imageToProximateImagingLocus: replaces csm::RasterGM::imageToProximateImagingLocus(const
 * ImageCoord&, const EcefCoord& groundPt, double desiredPrecision, double* achievedPrecision,
 * WarningList*) (LS:1175-1235, PF:995-1055, FR:424-462, SAR:959-993): the point of the pixel's
 * imaging locus nearest the given ground point and the locus direction there.  Inputs SoA
 * (line, samp, x, y, z); outputs as for the remote locus. */
USGSCSM_B200_API int usgscsm_b200_image_to_proximate_imaging_locus(
    usgscsm_b200_model *m, int64_t n, const double *line, const double *samp, const double *x,
    const double *y, const double *z, double prec, double *px, double *py, double *pz, double *dx,
    double *dy, double *dz, double *achieved, uint32_t *status);
USGSCSM_B200_API int usgscsm_b200_image_to_proximate_imaging_locus_dev(
    usgscsm_b200_model *m, int64_t n, const double *line, const double *samp, const double *x,
    const double *y, const double *z, double prec, double *px, double *py, double *pz, double *dx,
    double *dy, double *dz, double *achieved, uint32_t *status, void *stream);

/* computeGroundPartials: replaces csm::RasterGM::computeGroundPartials(const EcefCoord&)
 * (LS:1299-1325, PF:1091-1116, FR:955-1003, SAR:1384-1406), four groundToImage searches per
 * point fused in one kernel.  partials: 6 doubles per point, [dLine/dX, dLine/dY, dLine/dZ,
 * dSamp/dX, dSamp/dY, dSamp/dZ] (the std::vector the scalar call returns).  status: only the
 * ERROR bits and NONFINITE (the scalar call has no warning list); a point whose scalar call
 * would have thrown gets NaN partials. */
USGSCSM_B200_API int usgscsm_b200_ground_partials(usgscsm_b200_model *m, int64_t n, const double *x,
                                 const double *y, const double *z, double *partials,
                                 uint32_t *status);
USGSCSM_B200_API int usgscsm_b200_ground_partials_dev(usgscsm_b200_model *m, int64_t n, const double *x,
                                     const double *y, const double *z, double *partials,
                                     uint32_t *status, void *stream);

/* computeSensorPartials: replaces csm::RasterGM::computeSensorPartials(int index, const
 * EcefCoord&, double desiredPrecision, double* achievedPrecision, WarningList*) (LS:1336-1348)
 * and, when img_line / img_samp are given, the overload that takes the image point
 * (LS:1373-1396; PF:1121-1162, FR:814-873, SAR:1308-1371): the partial derivatives of line and sample with respect to adjustable
 * parameter `index`, by a forward difference of two groundToImage searches, the second with
 * the parameter moved by the model's own step: m_gsd (LS:1386, PF:1153), 1.0 or 0.01 for the
 * quaternion parameters (FR:853-859), m_scaledPixelWidth (SAR:1360).
 * computeAllSensorPartials (csm/RasterGM.cpp:52-100) is this call once per parameter index. */
USGSCSM_B200_API int usgscsm_b200_sensor_partials(usgscsm_b200_model *m, int index, int64_t n,
                                 const double *x, const double *y, const double *z,
                                 const double *img_line, const double *img_samp, double prec,
                                 double *dline, double *dsamp, double *achieved,
                                 uint32_t *status);
USGSCSM_B200_API int usgscsm_b200_sensor_partials_dev(usgscsm_b200_model *m, int index, int64_t n,
                                     const double *x, const double *y, const double *z,
                                     const double *img_line, const double *img_samp,
                                     double prec, double *dline, double *dsamp,
                                     double *achieved, uint32_t *status, void *stream);

/* computeAllSensorPartials(ground_pt | image_pt + ground_pt, pSet, prec, *ach, *warn)
 * (csm/RasterGM.cpp:52-140 through LS:1413-1455, PF:1170-1201, FR:894-940; SAR takes the
 * RasterGM default): the image point once (skipped when img_line / img_samp are given), then
 * one adjusted search per requested parameter.  indices = NULL: all parameters, in order
 * (n_indices must then be usgscsm_b200_model_num_parameters; csm::param::VALID for these
 * models); otherwise what getParameterSetIndices(pSet) would return.  partials: per point,
 * per requested parameter, (dline, dsamp) -- n * n_indices * 2 doubles, the batch form of
 * std::vector<csm::RasterGM::SensorPartials>.  achieved (nullable): the largest achieved
 * precision over the parameters, starting from 0.  status (nullable): warning and error bits
 * of all the searches; with an error bit (the scalar call would have thrown) the point's
 * whole list is NaN. */
USGSCSM_B200_API int usgscsm_b200_all_sensor_partials(usgscsm_b200_model *m, int n_indices, const int *indices,
                                     int64_t n, const double *x, const double *y, const double *z,
                                     const double *img_line, const double *img_samp,
                                     double desired_precision, double *partials,
                                     double *achieved, uint32_t *status);
USGSCSM_B200_API int usgscsm_b200_all_sensor_partials_dev(usgscsm_b200_model *m, int n_indices, const int *indices,
                                         int64_t n, const double *x, const double *y, const double *z,
                                         const double *img_line, const double *img_samp,
                                         double desired_precision, double *partials,
                                         double *achieved, uint32_t *status, void *stream);

/* getImageTime(ImageCoord), getSensorPosition(ImageCoord) and getSensorVelocity(ImageCoord)
 * for arrays (LS:1543-1653, PF:1258-1320, FR:665-765, SAR:1113-1238): the model-relative
 * image time and the sensor's ECEF position / velocity at it, with the model's current
 * parameter values.  Every output array may be NULL.  status: USGSCSM_B200_ST_ERROR_OTHER where
 * the scalar call throws csm::Error::BOUNDS (frame model, velocity of a point outside the
 * image, FR:754-760). */
USGSCSM_B200_API int usgscsm_b200_image_sensor_state(usgscsm_b200_model *m, int64_t n, const double *line,
                                    const double *samp, double *time, double *px, double *py,
                                    double *pz, double *vx, double *vy, double *vz,
                                    uint32_t *status);
USGSCSM_B200_API int usgscsm_b200_image_sensor_state_dev(usgscsm_b200_model *m, int64_t n, const double *line,
                                        const double *samp, double *time, double *px, double *py,
                                        double *pz, double *vx, double *vy, double *vz,
                                        uint32_t *status, void *stream);

/* getSensorPosition(double time) / getSensorVelocity(double time) for arrays of image times
 * (LS:1599-1653, PF:1281-1320, SAR:1160-1260; frame FR:707-795: the only valid time is 0.0,
 * any other gets USGSCSM_B200_ST_ERROR_OTHER = csm::Error::BOUNDS and NaN outputs). */
USGSCSM_B200_API int usgscsm_b200_sensor_state_at_time(usgscsm_b200_model *m, int64_t n, const double *time,
                                      double *px, double *py, double *pz, double *vx, double *vy,
                                      double *vz, uint32_t *status);
USGSCSM_B200_API int usgscsm_b200_sensor_state_at_time_dev(usgscsm_b200_model *m, int64_t n, const double *time,
                                          double *px, double *py, double *pz, double *vx,
                                          double *vy, double *vz, uint32_t *status, void *stream);

/* getIlluminationDirection(const csm::EcefCoord&) for arrays (LS:1878-1899, PF:1477-1498,
 * SAR:1095-1102, FR:643-652): the unit vector from the sun to the ground point at the time the
 * point was imaged (frame model: ground - sun, not normalised, as the reference returns it).
 * status (nullable): an error bit where the groundToImage inside would have thrown. */
USGSCSM_B200_API int usgscsm_b200_illumination_direction(usgscsm_b200_model *m, int64_t n, const double *x,
                                        const double *y, const double *z, double *dx, double *dy,
                                        double *dz, uint32_t *status);
USGSCSM_B200_API int usgscsm_b200_illumination_direction_dev(usgscsm_b200_model *m, int64_t n, const double *x,
                                            const double *y, const double *z, double *dx,
                                            double *dy, double *dz, uint32_t *status, void *stream);

/* ---- error propagation -------------------------------------------------------------------
 * groundToImage(const csm::EcefCoordCovar&, prec, *ach, *warn) -> csm::ImageCoordCovar and
 * imageToGround(const csm::ImageCoordCovar&, height, heightVariance, prec, *ach, *warn) ->
 * csm::EcefCoordCovar of the line-scan, push-frame and SAR models (LS:910-964 / 1075-1157,
 * PF:784-840 / 909-990, SAR:709-758 / 864-940).  ground_cov: 9 doubles per point (row-major
 * 3x3, csm::EcefCoordCovar::covariance), image_cov: 4 doubles per point (2x2,
 * csm::ImageCoordCovar::covariance).  The sensor term uses the model's parameter covariance
 * (m_covariance of the state, usgscsm_b200_model_set_parameter_covariance) and all sensor
 * partials at the default precision 0.001, as determineSensorCovarianceInImageSpace does
 * (LS:1025-1054).  status carries the warnings of the plain projection; an error bit means one
 * of the scalar calls inside would have thrown, and every output of the point is NaN.  The
 * frame model's overloads are stubs in the reference (FR:234-278, 374-395):
 * USGSCSM_B200_EUNSUPPORTED. */
USGSCSM_B200_API int usgscsm_b200_ground_to_image_covar(usgscsm_b200_model *m, int64_t n, const double *x,
                                       const double *y, const double *z, const double *ground_cov,
                                       double desired_precision, double *line, double *samp,
                                       double *image_cov, double *achieved, uint32_t *status);
USGSCSM_B200_API int usgscsm_b200_ground_to_image_covar_dev(usgscsm_b200_model *m, int64_t n, const double *x,
                                           const double *y, const double *z,
                                           const double *ground_cov, double desired_precision,
                                           double *line, double *samp, double *image_cov,
                                           double *achieved, uint32_t *status, void *stream);
/* height_stride / variance_stride: 0 = one value for all points, otherwise one per point. */
USGSCSM_B200_API int usgscsm_b200_image_to_ground_covar(usgscsm_b200_model *m, int64_t n, const double *line,
                                       const double *samp, const double *image_cov,
                                       const double *height, int height_stride,
                                       const double *height_variance, int variance_stride,
                                       double desired_precision, double *x, double *y, double *z,
                                       double *ground_cov, double *achieved, uint32_t *status);
USGSCSM_B200_API int usgscsm_b200_image_to_ground_covar_dev(usgscsm_b200_model *m, int64_t n, const double *line,
                                           const double *samp, const double *image_cov,
                                           const double *height, int height_stride,
                                           const double *height_variance, int variance_stride,
                                           double desired_precision, double *x, double *y,
                                           double *z, double *ground_cov, double *achieved,
                                           uint32_t *status, void *stream);
/* getParameterCovariance / setParameterCovariance (LS:1468-1497, PF:1207-1230, SAR:1604-1623) */
USGSCSM_B200_API int usgscsm_b200_model_get_parameter_covariance(const usgscsm_b200_model *m, int index1,
                                                int index2, double *value);
USGSCSM_B200_API int usgscsm_b200_model_set_parameter_covariance(usgscsm_b200_model *m, int index1, int index2,
                                                double value);

/* Several GPUs from one process (row-band / chunk sharding, no exchange step): models[g] are
 * the SAME model state loaded on n_models different devices; contiguous chunk g of the host
 * arrays (sizes differ by at most one point) is processed on models[g], all concurrently.
 * For a dense grid stored row-major, contiguous chunks are contiguous row bands. */
USGSCSM_B200_API int usgscsm_b200_ground_to_image_multi(usgscsm_b200_model *const *models, int n_models,
                                       int64_t n, const double *x, const double *y,
                                       const double *z, double prec, double *line, double *samp,
                                       double *achieved, uint32_t *status);
USGSCSM_B200_API int usgscsm_b200_image_to_ground_multi(usgscsm_b200_model *const *models, int n_models,
                                       int64_t n, const double *line, const double *samp,
                                       const double *height, int height_stride, double prec,
                                       double *x, double *y, double *z, double *achieved,
                                       uint32_t *status);

/* Many small images in one call ("small images can instead be distributed
 * whole"; config C5 packs several per launch): models[k] (all of one device) processes the next
 * counts[k] points of the concatenated device arrays.  A 4096 x 4096 frame or push-frame image is
 * a 0.2 - 0.7 ms kernel that spends a good part of its life ramping up and draining; here the
 * images' kernels are spread over a few internal streams forked from / joined to `stream`, so one
 * image's tail overlaps the next image's body.  Nothing is synchronised with the host. */
USGSCSM_B200_API int usgscsm_b200_ground_to_image_images_dev(usgscsm_b200_model *const *models, int n_models,
                                                            const int64_t *counts, const double *x,
                                                            const double *y, const double *z, double prec,
                                                            double *line, double *samp, double *achieved,
                                                            uint32_t *status, void *stream);
USGSCSM_B200_API int usgscsm_b200_image_to_ground_images_dev(usgscsm_b200_model *const *models, int n_models,
                                                            const int64_t *counts, const double *line,
                                                            const double *samp, const double *height,
                                                            int height_stride, double prec, double *x,
                                                            double *y, double *z, double *achieved,
                                                            uint32_t *status, void *stream);
/* The same over dense pixel grids: image k is the grid (line0 + r * dline, samp0 + c * dsamp),
 * r < n_lines[k], c < n_samps[k]; its n_lines[k] * n_samps[k] points follow image k - 1's. */
USGSCSM_B200_API int usgscsm_b200_image_to_ground_grid_images_dev(
    usgscsm_b200_model *const *models, int n_models, double line0, double samp0, double dline,
    double dsamp, const int64_t *n_lines, const int64_t *n_samps, const double *height,
    int height_stride, double prec, double *x, double *y, double *z, double *achieved,
    uint32_t *status, void *stream);

/* Dense-grid imageToGround: the double loop of bin/usgscsm_cam_test.cc:294-309 (and of an
 * orthorectifier) with the pixel coordinates generated in the kernel.  Point k =
 * r * n_samps + c is pixel (line0 + r * dline, samp0 + c * dsamp), r < n_lines,
 * c < n_samps; `height` as above, indexed by k.  Device pointers. */
USGSCSM_B200_API int usgscsm_b200_image_to_ground_grid_dev(usgscsm_b200_model *m, double line0, double samp0,
                                          double dline, double dsamp, int64_t n_lines,
                                          int64_t n_samps, const double *height,
                                          int height_stride, double prec, double *x, double *y,
                                          double *z, double *achieved, uint32_t *status,
                                          void *stream);

/* usgscsm_b200_image_to_ground_grid_dev for callers that may not allocate device memory (code
 * under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  This entry
 * point has no counterpart in the reference.
 * Same arguments, same results (the same two kernels); line-scan models only
 * (USGSCSM_B200_EUNSUPPORTED otherwise: the other models' grid kernels use no scratch).  The
 * per-row / per-column tables go to a file-scope device array of 2 Mi doubles = 16 MiB, which
 * every GPU that loads the library holds until the process exits; a grid fits when
 * n_lines * 21 + n_samps * 4 <= 2097152 (config C2's 52224 x 5064 does), anything larger is
 * rejected with USGSCSM_B200_EINVAL.  The call makes no cudaMalloc / cudaFree / stream-ordered
 * allocation.  All calls share the one array per GPU, and it keeps the last call's tables
 * (every entry a call reads is rewritten by that call first): calls whose kernels overlap on
 * one GPU -- different streams or host threads -- race; order them on one stream. */
#define USGSCSM_B200_STATIC_GRID_SCRATCH_DOUBLES 2097152
USGSCSM_B200_API int usgscsm_b200_image_to_ground_grid_static_dev(
    usgscsm_b200_model *m, double line0, double samp0, double dline, double dsamp, int64_t n_lines,
    int64_t n_samps, const double *height, int height_stride, double prec, double *x, double *y,
    double *z, double *achieved, uint32_t *status, void *stream);

/* The same with HOST arrays: heights up (one double when height_stride = 0), ground points
 * down, chunked by whole rows through the copy pipeline. */
USGSCSM_B200_API int usgscsm_b200_image_to_ground_grid(usgscsm_b200_model *m, double line0, double samp0,
                                      double dline, double dsamp, int64_t n_lines,
                                      int64_t n_samps, const double *height, int height_stride,
                                      double prec, double *x, double *y, double *z,
                                      double *achieved, uint32_t *status);

/* ---- georeferenced grids: the callers either side of the path ------------
 * What an orthorectifier does around groundToImage / imageToGround, and what
 * UsgsAstroProjectedSensorModel wraps around a camera (src/UsgsAstroProjectedSensorModel.cpp:
 * 289-365): raster pixel -> map coordinates by the GDAL geotransform (pixelToMeter / meterToPixel,
 * src/Utilities.cpp:2720-2745; gt[2] and gt[4] are ignored as there) -> geographic coordinates by
 * the inverse map projection -> body-fixed cartesian coordinates.  The reference delegates the
 * last two steps to PROJ (proj_trans of "+proj=<...>" to "+proj=geocent +a +b",
 * UsgsAstroProjectedSensorModel.cpp:128-147); the engine evaluates three projections itself with
 * PROJ's arithmetic (PROJ/src/conversions/cart.cpp, projections/eqc.cpp, projections/gn_sinu.cpp,
 * mlfn.cpp, fwd.cpp, inv.cpp as vendored with the reference). */
#define USGSCSM_B200_PROJ_LONGLAT 0 /* +proj=longlat: map x = longitude, y = latitude, degrees */
#define USGSCSM_B200_PROJ_EQC 1     /* +proj=eqc  (+lat_ts +lat_0 +lon_0 +x_0 +y_0)            */
#define USGSCSM_B200_PROJ_SINU 2    /* +proj=sinu (+lon_0 +x_0 +y_0), sphere or ellipsoid      */

typedef struct usgscsm_b200_geogrid {
  double geotransform[6]; /* GDAL order (m_geoTransform): x = gt[0] + samp * gt[1], y = gt[3] + line * gt[5] */
  int projection;         /* USGSCSM_B200_PROJ_*                                                 */
  int reserved;
  double lat_ts_deg, lat_0_deg, lon_0_deg; /* eqc: +lat_ts +lat_0; eqc / sinu: +lon_0             */
  double x_0, y_0;                         /* false easting / northing, metres                  */
  double proj_semi_major, proj_semi_minor; /* ellipsoid of the projected CRS (+R, or +a +b)     */
  double semi_major, semi_minor; /* ellipsoid of the cartesian step; 0 = the projection's.  The
                                    projected sensor model passes the camera's radii
                                    (UsgsAstroProjectedSensorModel.cpp:131-139)                  */
} usgscsm_b200_geogrid;

/* Fills *out from a PROJ string (m_projString) and a geotransform.  Supported: +proj=longlat
 * (latlong, lonlat, latlon) | eqc | sinu with +lat_ts +lat_0 +lon_0 +x_0 +y_0 +R | +a +b | +a +rf |
 * +a +f, +units=m, +no_defs, +type=crs.  Anything else: USGSCSM_B200_EUNSUPPORTED.  cart_a / cart_b:
 * radii of the cartesian step (0 = the projection's ellipsoid).  Host only. */
USGSCSM_B200_API int usgscsm_b200_geogrid_from_proj_string(const char *proj_string,
                                                          const double geotransform[6], double cart_a,
                                                          double cart_b, usgscsm_b200_geogrid *out);

/* UsgsAstroProjectedSensorModel::imageToGround(ImageCoord, height, ...)
 * (src/UsgsAstroProjectedSensorModel.cpp:329-365) for arrays: raster (line, samp) -> body-fixed
 * (x, y, z).  height: nullable (0 for every point, what the reference effectively uses), n entries
 * when height_stride = 1, one when 0.  status (nullable): USGSCSM_B200_ST_NONFINITE outside the
 * projection's domain.  `device`: the CUDA device to run on (host flavour) / the arrays live on. */
USGSCSM_B200_API int usgscsm_b200_geogrid_image_to_ground(const usgscsm_b200_geogrid *grid, int device,
                                                         int64_t n, const double *line, const double *samp,
                                                         const double *height, int height_stride,
                                                         double *x, double *y, double *z, uint32_t *status);
USGSCSM_B200_API int usgscsm_b200_geogrid_image_to_ground_dev(const usgscsm_b200_geogrid *grid, int device,
                                                             int64_t n, const double *line, const double *samp,
                                                             const double *height, int height_stride,
                                                             double *x, double *y, double *z, uint32_t *status,
                                                             void *stream);
/* UsgsAstroProjectedSensorModel::groundToImage(EcefCoord, ...)
 * (src/UsgsAstroProjectedSensorModel.cpp:289-321): body-fixed (x, y, z) -> raster (line, samp);
 * height (nullable) receives the height above the ellipsoid PROJ's geodetic conversion yields. */
USGSCSM_B200_API int usgscsm_b200_geogrid_ground_to_image(const usgscsm_b200_geogrid *grid, int device,
                                                         int64_t n, const double *x, const double *y,
                                                         const double *z, double *line, double *samp,
                                                         double *height, uint32_t *status);
USGSCSM_B200_API int usgscsm_b200_geogrid_ground_to_image_dev(const usgscsm_b200_geogrid *grid, int device,
                                                             int64_t n, const double *x, const double *y,
                                                             const double *z, double *line, double *samp,
                                                             double *height, uint32_t *status, void *stream);

/* A DEM raster: heights above the ellipsoid on a georeferenced grid. */
#define USGSCSM_B200_DEM_F64 0
#define USGSCSM_B200_DEM_F32 1
#define USGSCSM_B200_DEM_I16 2
typedef struct usgscsm_b200_dem {
  const void *samples;  /* n_rows x n_cols, row-major; HOST memory for the plain entry points,
                           DEVICE memory for the _dev ones                                         */
  int sample_type;      /* USGSCSM_B200_DEM_*                                                      */
  int has_nodata;
  double nodata;        /* raw sample value that marks a hole                                      */
  double scale, offset; /* height = raw * scale + offset (GDAL band scale / offset; 1, 0 = metres) */
  int64_t n_rows, n_cols;
  usgscsm_b200_geogrid grid;
} usgscsm_b200_dem;

/* The ground points of a DEM window -- rows row0 .. row0 + n_rows - 1, columns col0 .. col0 +
 * n_cols - 1, pixel centres (row + 0.5, col + 0.5), point k = r * n_cols + c -- as the chain
 * above gives them with height = the sample.  For longlat and eqc grids the trigonometry is
 * evaluated once per row and per column ON THE HOST with the C library (the functions PROJ
 * calls), so the result equals PROJ's cart conversion of the pixel bit for bit; sinu is
 * evaluated per pixel with CUDA's sin / cos (<= 2 ulp).  Nodata samples: NaN and
 * USGSCSM_B200_ST_DEM_NODATA. */
USGSCSM_B200_API int usgscsm_b200_dem_ground_points(const usgscsm_b200_dem *dem, int device, int64_t row0,
                                                   int64_t col0, int64_t n_rows, int64_t n_cols, double *x,
                                                   double *y, double *z, uint32_t *status);
USGSCSM_B200_API int usgscsm_b200_dem_ground_points_dev(const usgscsm_b200_dem *dem, int device, int64_t row0,
                                                       int64_t col0, int64_t n_rows, int64_t n_cols,
                                                       double *x, double *y, double *z, uint32_t *status,
                                                       void *stream);

/* Orthorectification's inner loop in one call: groundToImage of every ground point of a DEM
 * window (as above), the ground points generated in the kernel -- 2 / 4 / 8 bytes per point go
 * up the host link instead of 24.  Line-scan models run one fused kernel; the other models a
 * ground-point kernel and their groundToImage kernel.  Outputs as usgscsm_b200_ground_to_image;
 * a nodata sample gives NaN and USGSCSM_B200_ST_DEM_NODATA | USGSCSM_B200_ST_NONFINITE. */
USGSCSM_B200_API int usgscsm_b200_ground_to_image_dem_grid(usgscsm_b200_model *m, const usgscsm_b200_dem *dem,
                                                          int64_t row0, int64_t col0, int64_t n_rows,
                                                          int64_t n_cols, double prec, double *line,
                                                          double *samp, double *achieved, uint32_t *status);
USGSCSM_B200_API int usgscsm_b200_ground_to_image_dem_grid_dev(usgscsm_b200_model *m,
                                                              const usgscsm_b200_dem *dem, int64_t row0,
                                                              int64_t col0, int64_t n_rows, int64_t n_cols,
                                                              double prec, double *line, double *samp,
                                                              double *achieved, uint32_t *status,
                                                              void *stream);

/* DEM-iterated ground intersection: imageToGround of each pixel at the height of the DEM where
 * the ray meets it, by the fixed point callers of the reference run around imageToGround(height):
 *   h = h0;  repeat { P = imageToGround(pixel, h);  (lon, lat) = geodetic(P);
 *                     hn = DEM bilinear at (lon, lat);  if |hn - h| <= tolerance stop;  h = hn }
 * at most max_iterations times (1..64).  The DEM is sampled bilinearly between pixel centres.
 * Outputs: the ground point, the height it was computed at (nullable), status: the bits of the
 * last imageToGround, USGSCSM_B200_ST_DEM_OUTSIDE / _DEM_NODATA where the ray leaves the raster
 * or meets a hole (the point of the last height is returned), USGSCSM_B200_ST_ITERATION_CAP
 * when the height did not settle.  The plain flavour uploads the whole DEM once per call. */
USGSCSM_B200_API int usgscsm_b200_image_to_ground_dem(usgscsm_b200_model *m, const usgscsm_b200_dem *dem,
                                                     int64_t n, const double *line, const double *samp,
                                                     double h0, double tolerance, int max_iterations,
                                                     double prec, double *x, double *y, double *z,
                                                     double *height, uint32_t *status);
USGSCSM_B200_API int usgscsm_b200_image_to_ground_dem_dev(usgscsm_b200_model *m, const usgscsm_b200_dem *dem,
                                                         int64_t n, const double *line, const double *samp,
                                                         double h0, double tolerance, int max_iterations,
                                                         double prec, double *x, double *y, double *z,
                                                         double *height, uint32_t *status, void *stream);

/* ---- building blocks of the per-point math, batched -------------------------------------
 * The reference exposes these as free functions and tests them directly
 * (tests/DistortionTests.cpp, tests/UtilitiesTests.cpp); exporting the device versions lets
 * the same known-answer vectors be replayed on the GPU.  HOST pointers, n entries each.
 *
 * applyDistortion / removeDistortion: include/usgscsm/Distortion.h:20-31
 * (src/Distortion.cpp:441-775, 208-424).  `type` is the DistortionType value
 * (include/usgscsm/Distortion.h:10-11), `coeffs` the opticalDistCoeffs (ncoeffs <= 20).
 * status: ST_ERROR_ALGORITHM where the reference throws, ST_DISTORTION_NOT_CONVERGED. */
USGSCSM_B200_API int usgscsm_b200_apply_distortion(int device, int type, const double *coeffs, int ncoeffs,
                                  double focal_length, double desired_precision, int64_t n,
                                  const double *ux, const double *uy, double *dx, double *dy,
                                  uint32_t *status);
USGSCSM_B200_API int usgscsm_b200_remove_distortion(int device, int type, const double *coeffs, int ncoeffs,
                                   double focal_length, double tolerance, int64_t n,
                                   const double *dx, const double *dy, double *ux, double *uy,
                                   uint32_t *status);
/* lagrangeInterp (include/usgscsm/Utilities.h; src/Utilities.cpp:409-513) of a uniformly
 * sampled table of `vector_length`-vectors (1..4) at n times; out is n * vector_length.
 * Fails with USGSCSM_B200_EINVAL where the reference throws INDEX_OUT_OF_RANGE
 * (num_time < 2). */
USGSCSM_B200_API int usgscsm_b200_lagrange_interp(int device, int num_time, const double *values,
                                 double start_time, double del_time, int vector_length, int order,
                                 int64_t n, const double *times, double *out);

/* The remaining free functions the reference tests directly (tests/UtilitiesTests.cpp:55-284,
 * 443-475), batched over n inputs with HOST pointers, so that its known-answer vectors can be
 * replayed on the device functions the kernels are built from.
 *
 * computeDistortedFocalPlaneCoordinates (src/Utilities.cpp:129-151): pixel -> focal plane (mm).
 * affine12 = iTransS[0..2], iTransL[0..2], sampleOrigin, lineOrigin, sampleSumming, lineSumming,
 * startingSample, startingLine. */
USGSCSM_B200_API int usgscsm_b200_focal_plane_coordinates(int device, const double affine12[12], int64_t n,
                                                         const double *line, const double *sample,
                                                         double *focal_x, double *focal_y);
/* computePixel (src/Utilities.cpp:317-331): focal plane (mm) -> pixel. */
USGSCSM_B200_API int usgscsm_b200_compute_pixel(int device, const double affine12[12], int64_t n,
                                               const double *focal_x, const double *focal_y, double *line,
                                               double *sample);
/* createCameraLookVector (src/Utilities.cpp:360-373): look = 3 doubles per input (x, y, z). */
USGSCSM_B200_API int usgscsm_b200_camera_look_vector(int device, double z_direction, double focal_length,
                                                    int64_t n, const double *focal_x, const double *focal_y,
                                                    double *look);
/* removeJitter / addJitter (src/Utilities.cpp:182-282): rolling-shutter jitter polynomial of the
 * frame model; remove != 0 selects removeJitter.  addJitter's tolerance / max_iterations as in the
 * reference's signature (defaults 1e-6, 20). */
USGSCSM_B200_API int usgscsm_b200_jitter(int device, int remove, int num_coeffs, const double *line_jitter,
                                        const double *sample_jitter, int num_line_times,
                                        const double *line_times, double tolerance, int max_iterations,
                                        int64_t n, const double *line, const double *sample,
                                        double *out_line, double *out_sample);
/* evaluatePolynomial / evaluatePolynomialDerivative (src/Utilities.cpp:696-727) of
 * c[0] + c[1] x + c[2] x^2 + c[3] x^3; either output may be NULL. */
USGSCSM_B200_API int usgscsm_b200_polynomial(int device, const double coeffs4[4], int64_t n, const double *x,
                                            double *value, double *derivative);
/* polynomialRoot (src/Utilities.cpp:771-791, maxIterations 30) of the same cubic from guess[i];
 * status: USGSCSM_B200_ST_ERROR_UNKNOWN where the reference throws. */
USGSCSM_B200_API int usgscsm_b200_polynomial_root(int device, const double coeffs4[4], double threshold,
                                                 int64_t n, const double *guess, double *root,
                                                 uint32_t *status);
/* brentRoot (src/Utilities.cpp:545-622) of the same cubic between lower[i] and upper[i]; status:
 * USGSCSM_B200_ST_ERROR_UNKNOWN where the reference throws (no sign change),
 * USGSCSM_B200_ST_ITERATION_CAP after 30 iterations; bits 16-23: function evaluations. */
USGSCSM_B200_API int usgscsm_b200_brent_root(int device, const double coeffs4[4], double epsilon, int64_t n,
                                            const double *lower, const double *upper, double *root,
                                            uint32_t *status);

/* Self-test of the engine's division: the kernels replace a / b by divideBy(a, b, r) with a
 * hoisted reciprocal (usgscsm_b200/csrc/device_math.cuh), which must equal the IEEE quotient
 * bit for bit.  Compares the two on `pairs` pseudo-random operand pairs (all exponents,
 * subnormals, infinities, NaNs, and the constant divisors the kernels use) on the GPU and
 * returns the number of differing results (0 expected). */
USGSCSM_B200_API int usgscsm_b200_selftest_division(int device, int64_t pairs, uint64_t seed,
                                   uint64_t *mismatches);

/* Measures the FP64 pipe the kernels are bound by (the roofline denominator of bench.py): a
 * dependent-free stream of one instruction kind on every SM for at least min_ms of kernel time.
 * op 0 = DFMA, 1 = DADD, 2 = DMUL, 3 = DADD / DMUL alternating (the mix the engine's bit-exact
 * arithmetic -- no FMA contraction, like the reference's x86-64 build -- can issue).  Writes FP64
 * thread-instructions per second (DFMA: x 2 for flops) and, nullable, the device's nominal SM
 * clock in MHz. */
USGSCSM_B200_API int usgscsm_b200_measure_fp64(int device, int op, double min_ms,
                              double *instructions_per_second, double *sm_clock_mhz);

/* Number of kernels this library has launched on the calling process so far (all models,
 * all streams).  bench.py reports the difference over its timed region as gpu_launches. */
USGSCSM_B200_API int64_t usgscsm_b200_kernel_launches(void);

#ifdef __cplusplus
}
#endif
#endif /* USGSCSM_B200_H_ */
