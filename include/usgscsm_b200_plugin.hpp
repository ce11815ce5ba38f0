// usgscsm_b200_plugin.hpp -- C++ host side of the engine, above the C-ABI (usgscsm_b200.h).
//
// Header-only mirror of the reference's plugin / sensor-model interface for the hot path:
// same names, argument meaning, defaults and error behaviour as
//   * UsgsAstroPlugin                 (include/usgscsm/UsgsAstroPlugin.h:42-72,
//                                      src/UsgsAstroPlugin.cpp:62-69, 158-250, 535-555),
//   * csm::RasterGM::groundToImage / imageToGround (csm/RasterGM.h; LS:757-890, 980-1009),
//   * csm value types ImageCoord / ImageVector / EcefCoord (csm/csm.h:212-445),
//     csm::Warning (csm/Warning.h:39-47) and csm::Error (csm/Error.h:40-48),
// plus the batched array entry points the reference does not have (groundToImageBatch,
// imageToGroundBatch, imageToGroundGrid).  Every call, scalar or batched, ends in a kernel
// launch through libusgscsm_b200.so: there is no CPU implementation behind this header, and
// with no CUDA device constructModelFromState throws csm::Error(SENSOR_MODEL_NOT_CONSTRUCTIBLE).
//
// The value types are declared here (namespace usgscsm_b200::csm) with the memory layout of
// the CSM API's, so arrays of ::csm::EcefCoord / ::csm::ImageCoord can be handed to the AoS
// batch calls with a reinterpret_cast.
#ifndef USGSCSM_B200_PLUGIN_HPP_
#define USGSCSM_B200_PLUGIN_HPP_

#include <cstddef>
#include <cstdint>
#include <cstdio>
#include <exception>
#include <list>
#include <algorithm>
#include <string>
#include <vector>

#include "usgscsm_b200.h"

namespace usgscsm_b200 {
namespace csm {

struct ImageCoord {  // csm/csm.h:212-226
  double line, samp;
  ImageCoord() : line(0.0), samp(0.0) {}
  ImageCoord(double aLine, double aSamp) : line(aLine), samp(aSamp) {}
};
struct ImageVector {  // csm/csm.h:258-272
  double line, samp;
  ImageVector() : line(0.0), samp(0.0) {}
  ImageVector(double aLine, double aSamp) : line(aLine), samp(aSamp) {}
};
struct EcefVector {  // csm/csm.h:360-376
  double x, y, z;
  EcefVector() : x(0.0), y(0.0), z(0.0) {}
  EcefVector(double aX, double aY, double aZ) : x(aX), y(aY), z(aZ) {}
};
struct EcefCoord {  // csm/csm.h:292-308
  double x, y, z;
  EcefCoord() : x(0.0), y(0.0), z(0.0) {}
  EcefCoord(double aX, double aY, double aZ) : x(aX), y(aY), z(aZ) {}
};

struct ImageCoordCovar : public ImageCoord {  // csm/csm.h:228-256
  double covariance[4];                        // 2x2, row-major
  ImageCoordCovar() : ImageCoord() { std::fill(covariance, covariance + 4, 0.0); }
  ImageCoordCovar(double aLine, double aSamp) : ImageCoord(aLine, aSamp) {
    std::fill(covariance, covariance + 4, 0.0);
  }
};
struct EcefCoordCovar : public EcefCoord {  // csm/csm.h:310-358
  double covariance[9];                      // 3x3, row-major
  EcefCoordCovar() : EcefCoord() { std::fill(covariance, covariance + 9, 0.0); }
  EcefCoordCovar(double aX, double aY, double aZ) : EcefCoord(aX, aY, aZ) {
    std::fill(covariance, covariance + 9, 0.0);
  }
};

struct EcefLocus {  // csm/csm.h:420-445
  EcefCoord point;
  EcefVector direction;
  EcefLocus() {}
  EcefLocus(double ptX, double ptY, double ptZ, double dirX, double dirY, double dirZ)
      : point(ptX, ptY, ptZ), direction(dirX, dirY, dirZ) {}
};

class Warning {  // csm/Warning.h:36-89
 public:
  enum WarningType {
    UNKNOWN_WARNING = 1,
    DATA_NOT_AVAILABLE,
    PRECISION_NOT_MET,
    NEGATIVE_PRECISION,
    IMAGE_COORD_OUT_OF_BOUNDS,
    IMAGE_ID_TOO_LONG,
    NO_INTERSECTION,
    DEPRECATED_FUNCTION
  };
  Warning() : theWarning(UNKNOWN_WARNING) {}
  Warning(const WarningType &aWarningType, const std::string &aMessage,
          const std::string &aFunction)
      : theWarning(aWarningType), theMessage(aMessage), theFunction(aFunction) {}
  WarningType getWarning() const { return theWarning; }
  const std::string &getMessage() const { return theMessage; }
  const std::string &getFunction() const { return theFunction; }

 private:
  WarningType theWarning;
  std::string theMessage, theFunction;
};
typedef std::list<Warning> WarningList;

class Error : public std::exception {  // csm/Error.h:37-100
 public:
  enum ErrorType {
    ALGORITHM = 1,
    BOUNDS,
    FILE_READ,
    FILE_WRITE,
    ILLEGAL_MATH_OPERATION,
    INDEX_OUT_OF_RANGE,
    INVALID_SENSOR_MODEL_STATE,
    INVALID_USE,
    ISD_NOT_SUPPORTED,
    MEMORY,
    SENSOR_MODEL_NOT_CONSTRUCTIBLE,
    SENSOR_MODEL_NOT_SUPPORTED,
    STRING_TOO_LONG,
    UNKNOWN_ERROR,
    UNSUPPORTED_FUNCTION,
    UNKNOWN_SUPPORT_DATA
  };
  Error() : theError(UNKNOWN_ERROR) {}
  Error(const ErrorType &aErrorType, const std::string &aMessage, const std::string &aFunction)
      : theError(aErrorType), theMessage(aMessage), theFunction(aFunction) {}
  virtual ~Error() throw() {}
  ErrorType getError() const { return theError; }
  const std::string &getMessage() const { return theMessage; }
  const std::string &getFunction() const { return theFunction; }
  virtual const char *what() const throw() { return theMessage.c_str(); }

 private:
  ErrorType theError;
  std::string theMessage, theFunction;
};

}  // namespace csm

// A sensor model living on one GPU: the hot-path subset of csm::RasterGM.
class RasterGM {
 public:
  RasterGM(usgscsm_b200_model *handle, const std::string &state) : m_h(handle), m_state(state) {}
  ~RasterGM() { usgscsm_b200_model_free(m_h); }  // csm::Model::~Model; the caller owns the model

  // ---- csm::Model / GeometricModel subset ------------------------------------------------
  std::string getModelName() const {
    switch (usgscsm_b200_model_kind(m_h)) {
      case USGSCSM_B200_MODEL_LINE_SCANNER: return "USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL";
      case USGSCSM_B200_MODEL_PUSH_FRAME: return "USGS_ASTRO_PUSH_FRAME_SENSOR_MODEL";
      case USGSCSM_B200_MODEL_FRAME: return "USGS_ASTRO_FRAME_SENSOR_MODEL";
      default: return "USGS_ASTRO_SAR_SENSOR_MODEL";
    }
  }
  std::string getModelState() const {  // current parameters / covariance / ellipsoid written back
    size_t need = 0;
    check(usgscsm_b200_model_get_state(m_h, NULL, 0, &need), "getModelState");
    std::string s(need, '\0');
    if (need) check(usgscsm_b200_model_get_state(m_h, &s[0], need, &need), "getModelState");
    return s;
  }
  csm::ImageVector getImageSize() const {  // LS:1845-1852
    int l = 0, s = 0;
    check(usgscsm_b200_model_image_size(m_h, &l, &s), "getImageSize");
    return csm::ImageVector(l, s);
  }
  std::pair<csm::ImageCoord, csm::ImageCoord> getValidImageRange() const {  // LS:1862-1869
    csm::ImageVector sz = getImageSize();
    return std::make_pair(csm::ImageCoord(0.0, 0.0), csm::ImageCoord(sz.line, sz.samp));
  }
  int getNumParameters() const { return usgscsm_b200_model_num_parameters(m_h); }
  double getParameterValue(int index) const {
    double v = 0.0;
    if (usgscsm_b200_model_get_parameter(m_h, index, &v) != 0)
      throw csm::Error(csm::Error::INDEX_OUT_OF_RANGE, usgscsm_b200_last_error(),
                       "getParameterValue");
    return v;
  }
  void setParameterValue(int index, double value) {
    if (usgscsm_b200_model_set_parameter(m_h, index, value) != 0)
      throw csm::Error(csm::Error::INDEX_OUT_OF_RANGE, usgscsm_b200_last_error(),
                       "setParameterValue");
  }
  int device() const { return usgscsm_b200_model_device(m_h); }
  usgscsm_b200_model *handle() const { return m_h; }

  // ---- csm::RasterGM per-point API (a batch of one; pays one kernel launch) ---------------
  csm::ImageCoord groundToImage(const csm::EcefCoord &groundPt, double desiredPrecision = 0.001,
                                double *achievedPrecision = NULL,
                                csm::WarningList *warnings = NULL) const {
    double l = 0, s = 0, a = 0;
    uint32_t st = 0;
    check(usgscsm_b200_ground_to_image(m_h, 1, &groundPt.x, &groundPt.y, &groundPt.z,
                                       desiredPrecision, &l, &s, &a, &st),
          "groundToImage");
    raiseOrWarn(st, warnings, "groundToImage");
    if (achievedPrecision) *achievedPrecision = a;
    return csm::ImageCoord(l, s);
  }
  csm::EcefCoord imageToGround(const csm::ImageCoord &imagePt, double height,
                               double desiredPrecision = 0.001, double *achievedPrecision = NULL,
                               csm::WarningList *warnings = NULL) const {
    double x = 0, y = 0, z = 0, a = 0;
    uint32_t st = 0;
    check(usgscsm_b200_image_to_ground(m_h, 1, &imagePt.line, &imagePt.samp, &height, 1,
                                       desiredPrecision, &x, &y, &z, &a, &st),
          "imageToGround");
    raiseOrWarn(st, warnings, "imageToGround");
    if (achievedPrecision) *achievedPrecision = a;
    return csm::EcefCoord(x, y, z);
  }

  // LS:1254-1282, PF:1060-1082, FR:488-543, SAR:1009-1030
  csm::EcefLocus imageToRemoteImagingLocus(const csm::ImageCoord &imagePt,
                                           double desiredPrecision = 0.001,
                                           double *achievedPrecision = NULL,
                                           csm::WarningList *warnings = NULL) const {
    double o[7] = {0, 0, 0, 0, 0, 0, 0};
    uint32_t st = 0;
    check(usgscsm_b200_image_to_remote_imaging_locus(m_h, 1, &imagePt.line, &imagePt.samp,
                                                     desiredPrecision, o, o + 1, o + 2, o + 3,
                                                     o + 4, o + 5, o + 6, &st),
          "imageToRemoteImagingLocus");
    raiseOrWarn(st, warnings, "imageToRemoteImagingLocus");
    if (achievedPrecision) *achievedPrecision = o[6];
    return csm::EcefLocus(o[0], o[1], o[2], o[3], o[4], o[5]);
  }
  // LS:1175-1235, PF:995-1055, FR:424-462, SAR:959-993
  csm::EcefLocus imageToProximateImagingLocus(const csm::ImageCoord &imagePt,
                                              const csm::EcefCoord &groundPt,
                                              double desiredPrecision = 0.001,
                                              double *achievedPrecision = NULL,
                                              csm::WarningList *warnings = NULL) const {
    double o[7] = {0, 0, 0, 0, 0, 0, 0};
    uint32_t st = 0;
    check(usgscsm_b200_image_to_proximate_imaging_locus(
              m_h, 1, &imagePt.line, &imagePt.samp, &groundPt.x, &groundPt.y, &groundPt.z,
              desiredPrecision, o, o + 1, o + 2, o + 3, o + 4, o + 5, o + 6, &st),
          "imageToProximateImagingLocus");
    raiseOrWarn(st, warnings, "imageToProximateImagingLocus");
    if (achievedPrecision) *achievedPrecision = o[6];
    return csm::EcefLocus(o[0], o[1], o[2], o[3], o[4], o[5]);
  }
  // LS:1299-1325, PF:1091-1116, FR:955-1003, SAR:1384-1406
  std::vector<double> computeGroundPartials(const csm::EcefCoord &groundPt) const {
    std::vector<double> p(6, 0.0);
    uint32_t st = 0;
    check(usgscsm_b200_ground_partials(m_h, 1, &groundPt.x, &groundPt.y, &groundPt.z, &p[0], &st),
          "computeGroundPartials");
    raiseOrWarn(st, NULL, "computeGroundPartials");
    return p;
  }
  // LS:1336-1396, PF:1121-1162, FR:814-873, SAR:1308-1371 (SensorPartials = pair<line, sample>)
  std::pair<double, double> computeSensorPartials(int index, const csm::EcefCoord &groundPt,
                                                  double desiredPrecision = 0.001,
                                                  double *achievedPrecision = NULL,
                                                  csm::WarningList *warnings = NULL) const {
    double dl = 0, ds = 0, a = 0;
    uint32_t st = 0;
    check(usgscsm_b200_sensor_partials(m_h, index, 1, &groundPt.x, &groundPt.y, &groundPt.z, NULL,
                                       NULL, desiredPrecision, &dl, &ds, &a, &st),
          "computeSensorPartials");
    raiseOrWarn(st, warnings, "computeSensorPartials");
    if (achievedPrecision) *achievedPrecision = a;
    return std::make_pair(dl, ds);
  }
  void computeSensorPartialsBatch(int index, size_t n, const double *x, const double *y,
                                  const double *z, double desiredPrecision, double *linePartial,
                                  double *sampPartial, double *achievedPrecision = NULL,
                                  uint32_t *status = NULL, const double *imgLine = NULL,
                                  const double *imgSamp = NULL) const {
    check(usgscsm_b200_sensor_partials(m_h, index, (int64_t)n, x, y, z, imgLine, imgSamp,
                                       desiredPrecision, linePartial, sampPartial,
                                       achievedPrecision, status),
          "computeSensorPartialsBatch");
  }
  // csm/RasterGM.cpp:52-140 via LS:1413-1455, PF:1170-1201, FR:894-940 (pSet = VALID: every
  // parameter of these models)
  std::vector<std::pair<double, double> > computeAllSensorPartials(
      const csm::EcefCoord &groundPt, double desiredPrecision = 0.001,
      double *achievedPrecision = NULL, csm::WarningList *warnings = NULL) const {
    const int np = getNumParameters();
    std::vector<double> p((size_t)np * 2);
    double a = 0;
    uint32_t st = 0;
    check(usgscsm_b200_all_sensor_partials(m_h, np, NULL, 1, &groundPt.x, &groundPt.y, &groundPt.z,
                                           NULL, NULL, desiredPrecision, p.data(), &a, &st),
          "computeAllSensorPartials");
    raiseOrWarn(st, warnings, "computeAllSensorPartials");
    if (achievedPrecision) *achievedPrecision = a;
    std::vector<std::pair<double, double> > out((size_t)np);
    for (int k = 0; k < np; k++) out[(size_t)k] = std::make_pair(p[2 * k], p[2 * k + 1]);
    return out;
  }
  // partials: n * numIndices * 2 doubles, (dline, dsamp) per point per parameter;
  // indices = NULL: all parameters (numIndices = getNumParameters()).
  void computeAllSensorPartialsBatch(size_t n, const double *x, const double *y, const double *z,
                                     double desiredPrecision, double *partials,
                                     double *achievedPrecision = NULL, uint32_t *status = NULL,
                                     const double *imgLine = NULL, const double *imgSamp = NULL,
                                     int numIndices = -1, const int *indices = NULL) const {
    check(usgscsm_b200_all_sensor_partials(m_h, numIndices < 0 ? getNumParameters() : numIndices,
                                           indices, (int64_t)n, x, y, z, imgLine, imgSamp,
                                           desiredPrecision, partials, achievedPrecision, status),
          "computeAllSensorPartialsBatch");
  }
  // ---- image time and sensor ephemeris of an image point (LS:1543-1653, PF:1258-1320,
  // FR:665-765, SAR:1113-1238)
  void imageSensorStateBatch(size_t n, const double *line, const double *samp, double *time,
                             double *px, double *py, double *pz, double *vx, double *vy,
                             double *vz, uint32_t *status = NULL) const {
    check(usgscsm_b200_image_sensor_state(m_h, (int64_t)n, line, samp, time, px, py, pz, vx, vy,
                                          vz, status),
          "imageSensorStateBatch");
  }
  double getImageTime(const csm::ImageCoord &imagePt) const {
    double t = 0;
    imageSensorStateBatch(1, &imagePt.line, &imagePt.samp, &t, NULL, NULL, NULL, NULL, NULL, NULL);
    return t;
  }
  csm::EcefCoord getSensorPosition(const csm::ImageCoord &imagePt) const {
    csm::EcefCoord p;
    imageSensorStateBatch(1, &imagePt.line, &imagePt.samp, NULL, &p.x, &p.y, &p.z, NULL, NULL, NULL);
    return p;
  }
  csm::EcefVector getSensorVelocity(const csm::ImageCoord &imagePt) const {
    csm::EcefVector v;
    uint32_t st = 0;
    imageSensorStateBatch(1, &imagePt.line, &imagePt.samp, NULL, NULL, NULL, NULL, &v.x, &v.y, &v.z,
                          &st);
    raiseOrWarn(st & ~(uint32_t)USGSCSM_B200_ST_NONFINITE, NULL, "getSensorVelocity");
    return v;
  }

  // LS:1878-1899, PF:1477-1498, SAR:1095-1102, FR:643-652
  csm::EcefVector getIlluminationDirection(const csm::EcefCoord &groundPt) const {
    csm::EcefVector v;
    uint32_t st = 0;
    check(usgscsm_b200_illumination_direction(m_h, 1, &groundPt.x, &groundPt.y, &groundPt.z, &v.x,
                                              &v.y, &v.z, &st),
          "getIlluminationDirection");
    raiseOrWarn(st & ~(uint32_t)USGSCSM_B200_ST_NONFINITE, NULL, "getIlluminationDirection");
    return v;
  }
  void getIlluminationDirectionBatch(size_t n, const double *x, const double *y, const double *z,
                                     double *dx, double *dy, double *dz,
                                     uint32_t *status = NULL) const {
    check(usgscsm_b200_illumination_direction(m_h, (int64_t)n, x, y, z, dx, dy, dz, status),
          "getIlluminationDirectionBatch");
  }

  // ---- error propagation (LS:910-964 / 1075-1157, PF:784-840 / 909-990, SAR:709-758 / 864-940)
  double getParameterCovariance(int index1, int index2) const {
    double v = 0.0;
    check(usgscsm_b200_model_get_parameter_covariance(m_h, index1, index2, &v),
          "getParameterCovariance");
    return v;
  }
  void setParameterCovariance(int index1, int index2, double covariance) {
    check(usgscsm_b200_model_set_parameter_covariance(m_h, index1, index2, covariance),
          "setParameterCovariance");
  }
  csm::ImageCoordCovar groundToImage(const csm::EcefCoordCovar &groundPt,
                                     double desiredPrecision = 0.001,
                                     double *achievedPrecision = NULL,
                                     csm::WarningList *warnings = NULL) const {
    csm::ImageCoordCovar r;
    double a = 0;
    uint32_t st = 0;
    check(usgscsm_b200_ground_to_image_covar(m_h, 1, &groundPt.x, &groundPt.y, &groundPt.z,
                                             groundPt.covariance, desiredPrecision, &r.line,
                                             &r.samp, r.covariance, &a, &st),
          "groundToImage");
    raiseOrWarn(st, warnings, "groundToImage");
    if (achievedPrecision) *achievedPrecision = a;
    return r;
  }
  csm::EcefCoordCovar imageToGround(const csm::ImageCoordCovar &imagePt, double height,
                                    double heightVariance, double desiredPrecision = 0.001,
                                    double *achievedPrecision = NULL,
                                    csm::WarningList *warnings = NULL) const {
    csm::EcefCoordCovar r;
    double a = 0;
    uint32_t st = 0;
    check(usgscsm_b200_image_to_ground_covar(m_h, 1, &imagePt.line, &imagePt.samp,
                                             imagePt.covariance, &height, 0, &heightVariance, 0,
                                             desiredPrecision, &r.x, &r.y, &r.z, r.covariance, &a,
                                             &st),
          "imageToGround");
    raiseOrWarn(st, warnings, "imageToGround");
    if (achievedPrecision) *achievedPrecision = a;
    return r;
  }
  // whole arrays: groundCov 9 doubles per point, imageCov 4 doubles per point
  void groundToImageCovarBatch(size_t n, const double *x, const double *y, const double *z,
                               const double *groundCov, double desiredPrecision, double *line,
                               double *samp, double *imageCov, double *achievedPrecision = NULL,
                               uint32_t *status = NULL) const {
    check(usgscsm_b200_ground_to_image_covar(m_h, (int64_t)n, x, y, z, groundCov, desiredPrecision,
                                             line, samp, imageCov, achievedPrecision, status),
          "groundToImageCovarBatch");
  }
  void imageToGroundCovarBatch(size_t n, const double *line, const double *samp,
                               const double *imageCov, const double *height, bool perPointHeight,
                               const double *heightVariance, bool perPointVariance,
                               double desiredPrecision, double *x, double *y, double *z,
                               double *groundCov, double *achievedPrecision = NULL,
                               uint32_t *status = NULL) const {
    check(usgscsm_b200_image_to_ground_covar(m_h, (int64_t)n, line, samp, imageCov, height,
                                             perPointHeight ? 1 : 0, heightVariance,
                                             perPointVariance ? 1 : 0, desiredPrecision, x, y, z,
                                             groundCov, achievedPrecision, status),
          "imageToGroundCovarBatch");
  }
  // csm::SettableEllipsoid (LS:2116-2119)
  void setEllipsoid(double semiMajorRadius, double semiMinorRadius) {
    check(usgscsm_b200_model_set_ellipsoid(m_h, semiMajorRadius, semiMinorRadius), "setEllipsoid");
  }

  // whole arrays: loci (SoA: point px,py,pz and direction dx,dy,dz) and ground partials
  // (6 doubles per point, the layout of the vector the scalar call returns)
  void imageToRemoteImagingLocusBatch(size_t n, const double *line, const double *samp,
                                      double desiredPrecision, double *px, double *py, double *pz,
                                      double *dx, double *dy, double *dz,
                                      double *achievedPrecision = NULL,
                                      uint32_t *status = NULL) const {
    check(usgscsm_b200_image_to_remote_imaging_locus(m_h, (int64_t)n, line, samp, desiredPrecision,
                                                     px, py, pz, dx, dy, dz, achievedPrecision,
                                                     status),
          "imageToRemoteImagingLocusBatch");
  }
  void imageToProximateImagingLocusBatch(size_t n, const double *line, const double *samp,
                                         const double *x, const double *y, const double *z,
                                         double desiredPrecision, double *px, double *py,
                                         double *pz, double *dx, double *dy, double *dz,
                                         double *achievedPrecision = NULL,
                                         uint32_t *status = NULL) const {
    check(usgscsm_b200_image_to_proximate_imaging_locus(m_h, (int64_t)n, line, samp, x, y, z,
                                                        desiredPrecision, px, py, pz, dx, dy, dz,
                                                        achievedPrecision, status),
          "imageToProximateImagingLocusBatch");
  }
  void computeGroundPartialsBatch(size_t n, const double *x, const double *y, const double *z,
                                  double *partials6n, uint32_t *status = NULL) const {
    check(usgscsm_b200_ground_partials(m_h, (int64_t)n, x, y, z, partials6n, status),
          "computeGroundPartialsBatch");
  }

  // ---- batched entry points: SoA host arrays (H2D/D2H inside) ------------------------------
  // status[i] carries the warnings / errors the scalar call would have raised for point i
  // (USGSCSM_B200_ST_*); no exception is thrown for a per-point failure.
  void groundToImageBatch(size_t n, const double *x, const double *y, const double *z,
                          double desiredPrecision, double *line, double *samp,
                          double *achievedPrecision = NULL, uint32_t *status = NULL) const {
    check(usgscsm_b200_ground_to_image(m_h, (int64_t)n, x, y, z, desiredPrecision, line, samp,
                                       achievedPrecision, status),
          "groundToImageBatch");
  }
  void imageToGroundBatch(size_t n, const double *line, const double *samp, const double *height,
                          bool perPointHeight, double desiredPrecision, double *x, double *y,
                          double *z, double *achievedPrecision = NULL,
                          uint32_t *status = NULL) const {
    check(usgscsm_b200_image_to_ground(m_h, (int64_t)n, line, samp, height, perPointHeight ? 1 : 0,
                                       desiredPrecision, x, y, z, achievedPrecision, status),
          "imageToGroundBatch");
  }
  // AoS flavours over the CSM value types (csm/csm.h:212-308).
  void groundToImageBatch(const std::vector<csm::EcefCoord> &groundPts, double desiredPrecision,
                          std::vector<csm::ImageCoord> &imagePts,
                          std::vector<double> *achievedPrecision = NULL,
                          std::vector<uint32_t> *status = NULL) const {
    const size_t n = groundPts.size();
    imagePts.resize(n);
    if (achievedPrecision) achievedPrecision->resize(n);
    if (status) status->resize(n);
    check(usgscsm_b200_ground_to_image_aos(
              m_h, (int64_t)n, n ? &groundPts[0].x : NULL, desiredPrecision,
              n ? &imagePts[0].line : NULL, achievedPrecision && n ? &(*achievedPrecision)[0] : NULL,
              status && n ? &(*status)[0] : NULL),
          "groundToImageBatch");
  }
  void imageToGroundBatch(const std::vector<csm::ImageCoord> &imagePts,
                          const std::vector<double> &heights, double desiredPrecision,
                          std::vector<csm::EcefCoord> &groundPts,
                          std::vector<double> *achievedPrecision = NULL,
                          std::vector<uint32_t> *status = NULL) const {
    const size_t n = imagePts.size();
    if (heights.size() != n && heights.size() != 1)
      throw csm::Error(csm::Error::BOUNDS, "heights must hold 1 or n entries", "imageToGroundBatch");
    groundPts.resize(n);
    if (achievedPrecision) achievedPrecision->resize(n);
    if (status) status->resize(n);
    check(usgscsm_b200_image_to_ground_aos(
              m_h, (int64_t)n, n ? &imagePts[0].line : NULL, &heights[0], heights.size() == n ? 1 : 0,
              desiredPrecision, n ? &groundPts[0].x : NULL,
              achievedPrecision && n ? &(*achievedPrecision)[0] : NULL,
              status && n ? &(*status)[0] : NULL),
          "imageToGroundBatch");
  }

  // ---- device-pointer flavours (enqueue on `stream`, no synchronisation) --------------------
  void groundToImageDevice(size_t n, const double *x, const double *y, const double *z,
                           double desiredPrecision, double *line, double *samp,
                           double *achievedPrecision, uint32_t *status, void *stream) const {
    check(usgscsm_b200_ground_to_image_dev(m_h, (int64_t)n, x, y, z, desiredPrecision, line, samp,
                                           achievedPrecision, status, stream),
          "groundToImageDevice");
  }
  void imageToGroundDevice(size_t n, const double *line, const double *samp, const double *height,
                           bool perPointHeight, double desiredPrecision, double *x, double *y,
                           double *z, double *achievedPrecision, uint32_t *status,
                           void *stream) const {
    check(usgscsm_b200_image_to_ground_dev(m_h, (int64_t)n, line, samp, height,
                                           perPointHeight ? 1 : 0, desiredPrecision, x, y, z,
                                           achievedPrecision, status, stream),
          "imageToGroundDevice");
  }
  // The pixel double loop of bin/usgscsm_cam_test.cc:294-309 with the coordinates generated
  // in the kernel; point k = r * nSamps + c is pixel (line0 + r*dline, samp0 + c*dsamp).
  void imageToGroundGridDevice(double line0, double samp0, double dline, double dsamp,
                               size_t nLines, size_t nSamps, const double *height,
                               bool perPointHeight, double desiredPrecision, double *x, double *y,
                               double *z, double *achievedPrecision, uint32_t *status,
                               void *stream) const {
    check(usgscsm_b200_image_to_ground_grid_dev(m_h, line0, samp0, dline, dsamp, (int64_t)nLines,
                                                (int64_t)nSamps, height, perPointHeight ? 1 : 0,
                                                desiredPrecision, x, y, z, achievedPrecision,
                                                status, stream),
          "imageToGroundGridDevice");
  }

  // The same with host arrays (heights up, ground points down, copies inside the call).
  void imageToGroundGrid(double line0, double samp0, double dline, double dsamp, size_t nLines,
                         size_t nSamps, const double *height, bool perPointHeight,
                         double desiredPrecision, double *x, double *y, double *z,
                         double *achievedPrecision = NULL, uint32_t *status = NULL) const {
    check(usgscsm_b200_image_to_ground_grid(m_h, line0, samp0, dline, dsamp, (int64_t)nLines,
                                            (int64_t)nSamps, height, perPointHeight ? 1 : 0,
                                            desiredPrecision, x, y, z, achievedPrecision, status),
          "imageToGroundGrid");
  }

  // Status word of one point -> what the scalar reference call does: throw csm::Error for
  // the error bits (Distortion.cpp:357-365, SAR:569-575), append csm::Warning otherwise.
  static void raiseOrWarn(uint32_t st, csm::WarningList *warnings, const char *function) {
    if (st & USGSCSM_B200_ST_ERROR_ALGORITHM)
      throw csm::Error(csm::Error::ALGORITHM, "Unable to apply transform.", function);
    if (st & USGSCSM_B200_ST_ERROR_UNKNOWN)
      throw csm::Error(csm::Error::UNKNOWN_ERROR, "Could not calculate the image coordinate.",
                       function);
    if (st & USGSCSM_B200_ST_ERROR_OTHER)
      throw csm::Error(csm::Error::UNKNOWN_ERROR, "per-point failure", function);
    if (!warnings) return;
    if (st & USGSCSM_B200_ST_NO_INTERSECTION)
      warnings->push_back(csm::Warning(csm::Warning::NO_INTERSECTION, "No Intersection", function));
    if (st & USGSCSM_B200_ST_PRECISION_NOT_MET)
      warnings->push_back(csm::Warning(csm::Warning::PRECISION_NOT_MET,
                                       "Desired precision not achieved.", function));
  }

 private:
  RasterGM(const RasterGM &);
  RasterGM &operator=(const RasterGM &);
  static void check(int code, const char *function) {
    if (code == USGSCSM_B200_OK) return;
    throw csm::Error(code == USGSCSM_B200_EINVAL ? csm::Error::BOUNDS : csm::Error::UNKNOWN_ERROR,
                     usgscsm_b200_last_error(), function);
  }
  usgscsm_b200_model *m_h;
  std::string m_state;
};

// src/UsgsAstroPlugin.cpp: the factory.  Models are built from model-state strings or from
// ISDs (converted on the host by usgscsm_b200_state_from_isd).
class UsgsAstroPlugin {
 public:
  explicit UsgsAstroPlugin(int device = 0) : m_device(device) {}
  std::string getPluginName() const { return "UsgsAstroPluginCSM"; }       // :62
  std::string getManufacturer() const { return "UsgsAstrogeology"; }       // :63
  std::string getReleaseDate() const { return "20190222"; }                // :64
  size_t getNumModels() const { return 4; }  // the projected wrapper (PROJ) is out of scope
  std::string getModelName(size_t modelIndex) const {                      // :158-170
    static const char *names[4] = {
        "USGS_ASTRO_FRAME_SENSOR_MODEL", "USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL",
        "USGS_ASTRO_PUSH_FRAME_SENSOR_MODEL", "USGS_ASTRO_SAR_SENSOR_MODEL"};
    if (modelIndex >= 4)
      throw csm::Error(csm::Error::INDEX_OUT_OF_RANGE, "Model index out of range.",
                       "UsgsAstroPlugin::getModelName");
    return names[modelIndex];
  }
  // :176-218 -- the JSON key "m_modelName" of the state (after stateAsJson U:2489-2500).
  std::string getModelNameFromModelState(const std::string &modelState,
                                         csm::WarningList * = NULL) const {
    char name[128];
    int code = usgscsm_b200_model_name_from_state(modelState.data(), modelState.size(), name,
                                                  sizeof(name));
    if (code == USGSCSM_B200_EUNSUPPORTED)
      throw csm::Error(csm::Error::SENSOR_MODEL_NOT_SUPPORTED, "Sensor model not supported.",
                       "UsgsAstroPlugin::getModelNameFromModelState");
    if (code != USGSCSM_B200_OK)
      throw csm::Error(csm::Error::INVALID_SENSOR_MODEL_STATE,
                       "No 'm_modelName' key in the model state object.",
                       "UsgsAstroPlugin::getModelNameFromModelState");
    return name;
  }
  bool canModelBeConstructedFromState(const std::string &modelName, const std::string &modelState,
                                      csm::WarningList * = NULL) const {  // :229-250
    try {
      return getModelNameFromModelState(modelState) == modelName;
    } catch (const csm::Error &) {
      return false;
    }
  }
  // :535-555.  Returns a raw heap pointer the caller owns and deletes (csm/Plugin.h:186-205).
  RasterGM *constructModelFromState(const std::string &modelState,
                                    csm::WarningList * = NULL) const {
    usgscsm_b200_model *h = NULL;
    int code = usgscsm_b200_model_from_state(modelState.data(), modelState.size(), m_device, &h);
    if (code == USGSCSM_B200_EUNSUPPORTED)
      throw csm::Error(csm::Error::SENSOR_MODEL_NOT_SUPPORTED, usgscsm_b200_last_error(),
                       "UsgsAstroPlugin::constructModelFromState");
    if (code != USGSCSM_B200_OK)
      throw csm::Error(csm::Error::SENSOR_MODEL_NOT_CONSTRUCTIBLE, usgscsm_b200_last_error(),
                       "UsgsAstroPlugin::constructModelFromState");
    return new RasterGM(h, modelState);
  }

  // ---- ISD side (include/usgscsm/UsgsAstroPlugin.h:42-72) ------------------------------------
  // loadImageSupportData :324-353: the csm::Isd names the IMAGE; the ISD is the sidecar
  // "<base name>.json", and the image's base name becomes the image identifier.
  static std::string loadImageSupportData(const std::string &imageFilename, std::string *identifier) {
    std::string base = imageFilename.substr(0, imageFilename.find_last_of("."));
    if (identifier) *identifier = base.substr(base.find_last_of("/") + 1);
    const std::string isdFilename = base + ".json";
    FILE *f = std::fopen(isdFilename.c_str(), "rb");
    if (!f)
      throw csm::Error(csm::Error::FILE_READ,
                       "Could not read metadata file associated with image [" + isdFilename + "]",
                       "UsgsAstroPlugin::loadImageSupportData");
    std::string text;
    char buf[1 << 16];
    size_t got;
    while ((got = std::fread(buf, 1, sizeof buf, f)) > 0) text.append(buf, got);
    std::fclose(f);
    return text;
  }
  // <Model>::constructStateFromIsd of the ISD text (host only; modelName "" = the ISD's own)
  std::string constructStateFromIsd(const std::string &isdText, const std::string &modelName,
                                    const std::string &imageIdentifier) const {
    size_t need = 0;
    const char *name = modelName.empty() ? NULL : modelName.c_str();
    int code = usgscsm_b200_state_from_isd(isdText.data(), isdText.size(), name,
                                           imageIdentifier.c_str(), NULL, 0, &need);
    if (code != USGSCSM_B200_OK)
      throw csm::Error(csm::Error::SENSOR_MODEL_NOT_CONSTRUCTIBLE, usgscsm_b200_last_error(),
                       "constructStateFromIsd");
    std::string state(need, '\0');
    usgscsm_b200_state_from_isd(isdText.data(), isdText.size(), name, imageIdentifier.c_str(),
                                &state[0], need, &need);
    return state;
  }
  // :488-519 (the four un-projected models).  imageFilename is csm::Isd::filename().
  RasterGM *constructModelFromISD(const std::string &imageFilename, const std::string &modelName,
                                  csm::WarningList *warnings = NULL) const {
    std::string ident;
    const std::string isd = loadImageSupportData(imageFilename, &ident);
    return constructModelFromState(constructStateFromIsd(isd, modelName, ident), warnings);
  }
  bool canModelBeConstructedFromISD(const std::string &imageFilename, const std::string &modelName,
                                    csm::WarningList *warnings = NULL) const {  // :272-307
    try {
      RasterGM *m = constructModelFromISD(imageFilename, modelName, warnings);
      delete m;
      return true;
    } catch (const std::exception &e) {
      if (warnings)
        warnings->push_back(csm::Warning(
            csm::Warning::UNKNOWN_WARNING,
            "Could not create model [" + modelName + "] with error [" + e.what() + "]",
            "UsgsAstroPlugin::canModelBeConstructedFromISD()"));
      return false;
    }
  }
  std::string convertISDToModelState(const std::string &imageFilename, const std::string &modelName,
                                     csm::WarningList *warnings = NULL) const {  // :465-472
    RasterGM *m = constructModelFromISD(imageFilename, modelName, warnings);
    std::string state = m->getModelState();
    delete m;
    return state;
  }
  std::string getStateFromISD(const std::string &imageFilename) const {  // :459-463
    return convertISDToModelState(imageFilename, "");
  }

 private:
  int m_device;
};

}  // namespace usgscsm_b200
#endif  // USGSCSM_B200_PLUGIN_HPP_
