// csm_plugin.cpp -- libusgscsm_b200_csm.so: the engine behind the real CSM API
// (include/usgscsm_b200_csm.h).  Host C++ only; every per-point virtual ends in a kernel launch
// through the C-ABI of libusgscsm_b200.so.
//
// Reference counterparts: src/UsgsAstroPlugin.cpp (factory, registration :72), the csm::RasterGM
// overrides of src/UsgsAstro{Ls,PushFrame,Frame,Sar}SensorModel.cpp (line numbers at each
// method).  Per-model metadata (pedigree, sensor type / mode, parameter names and units) is
// what those classes hard-code.
#include "../../include/usgscsm_b200_csm.h"

#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <ctime>
#include <fstream>
#include <sstream>

#include "Error.h"
#include "Isd.h"
#include "Warning.h"

#include "../csrc/state_reader.h"

namespace usgscsm_b200 {

namespace {

const char *const kModelNames[4] = {
    "USGS_ASTRO_FRAME_SENSOR_MODEL", "USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL",
    "USGS_ASTRO_PUSH_FRAME_SENSOR_MODEL", "USGS_ASTRO_SAR_SENSOR_MODEL"};

[[noreturn]] void fail(::csm::Error::ErrorType type, const std::string &msg, const char *fn) {
  throw ::csm::Error(type, msg, fn);
}
void check(int code, const char *fn) {
  if (code == USGSCSM_B200_OK) return;
  fail(code == USGSCSM_B200_EINVAL ? ::csm::Error::BOUNDS : ::csm::Error::UNKNOWN_ERROR,
       usgscsm_b200_last_error(), fn);
}
// The per-point status word back into what the scalar reference call does: a thrown csm::Error
// (Distortion.cpp:357-365 ALGORITHM; SAR:569-575 UNKNOWN_ERROR; FR:754-760 BOUNDS) or warnings
// appended to the caller's list (LS:882-887, 2433-2441).
void raiseOrWarn(uint32_t st, ::csm::WarningList *warnings, const char *fn) {
  if (st & USGSCSM_B200_ST_ERROR_ALGORITHM) fail(::csm::Error::ALGORITHM, "Unable to apply transform.", fn);
  if (st & USGSCSM_B200_ST_ERROR_UNKNOWN)
    fail(::csm::Error::UNKNOWN_ERROR, "Could not calculate the image coordinate.", fn);
  if (st & USGSCSM_B200_ST_ERROR_OTHER) fail(::csm::Error::BOUNDS, "Image coordinate or time out of bounds.", fn);
  if (!warnings) return;
  if (st & USGSCSM_B200_ST_NO_INTERSECTION)
    warnings->push_back(::csm::Warning(::csm::Warning::NO_INTERSECTION, "No Intersection", fn));
  if (st & USGSCSM_B200_ST_PRECISION_NOT_MET)
    warnings->push_back(::csm::Warning(::csm::Warning::PRECISION_NOT_MET, "Desired precision not achieved.", fn));
}

// Ephemeris time (TDB seconds past J2000) -> UTC calendar string "YYYY-MM-DDTHH:MM:SS.mmmZ"
// (ephemTimeToCalendarTime, src/Utilities.cpp): ET - 32.184 s = TAI; TAI - leap seconds = UTC,
// the leap-second count taken from the IERS table for the UTC date.
std::string calendarTime(double et) {
  static const int leapDates[][4] = {
      {1972, 1, 1, 10}, {1972, 7, 1, 11}, {1973, 1, 1, 12}, {1974, 1, 1, 13}, {1975, 1, 1, 14},
      {1976, 1, 1, 15}, {1977, 1, 1, 16}, {1978, 1, 1, 17}, {1979, 1, 1, 18}, {1980, 1, 1, 19},
      {1981, 7, 1, 20}, {1982, 7, 1, 21}, {1983, 7, 1, 22}, {1985, 7, 1, 23}, {1988, 1, 1, 24},
      {1990, 1, 1, 25}, {1991, 1, 1, 26}, {1992, 7, 1, 27}, {1993, 7, 1, 28}, {1994, 7, 1, 29},
      {1996, 1, 1, 30}, {1997, 7, 1, 31}, {1999, 1, 1, 32}, {2006, 1, 1, 33}, {2009, 1, 1, 34},
      {2012, 7, 1, 35}, {2015, 7, 1, 36}, {2017, 1, 1, 37}};
  const double taiUnix = et - 32.184 + 946728000.0;  // J2000 (TAI) as Unix seconds
  std::time_t guess = (std::time_t)(taiUnix - 32);
  std::tm *g = std::gmtime(&guess);
  const long date = (g->tm_year + 1900) * 10000L + (g->tm_mon + 1) * 100 + g->tm_mday;
  int leap = 10;
  for (const auto &d : leapDates)
    if (date >= d[0] * 10000L + d[1] * 100 + d[2]) leap = d[3];
  const double utc = taiUnix - leap;
  std::time_t whole = (std::time_t)std::floor(utc);
  int msec = (int)((utc - (double)whole) * 1000.0 + 0.5);
  if (msec >= 1000) {
    whole++;
    msec -= 1000;
  }
  char buf[40];
  std::strftime(buf, sizeof buf, "%Y-%m-%dT%H:%M:%S", std::gmtime(&whole));
  char out[80];
  std::snprintf(out, sizeof out, "%s.%03dZ", buf, msec);
  return out;
}

int deviceFromEnv() {
  const char *v = std::getenv("USGSCSM_B200_DEVICE");
  return v && *v ? std::atoi(v) : 0;
}

}  // namespace

// ---- SensorModel ------------------------------------------------------------------------------------
struct SensorModel::Meta {
  ub200::json::Value state;  // the state as loaded
  std::vector< ::csm::param::Type> parameterType;
  ::csm::EcefCoord referencePoint;
  std::string imageIdentifier;
};

SensorModel::SensorModel(const std::string &state, int device)
    : m_h(NULL), m_device(device), m_meta(new Meta) {
  try {
    load(state);
  } catch (...) {
    delete m_meta;
    throw;
  }
}
SensorModel::~SensorModel() {
  usgscsm_b200_model_free(m_h);
  delete m_meta;
}
int SensorModel::device() const { return m_device; }

void SensorModel::load(const std::string &state) {
  usgscsm_b200_model *h = NULL;
  const int code = usgscsm_b200_model_from_state(state.data(), state.size(), m_device, &h);
  if (code == USGSCSM_B200_EUNSUPPORTED)
    fail(::csm::Error::SENSOR_MODEL_NOT_SUPPORTED, usgscsm_b200_last_error(), "SensorModel::replaceModelState");
  if (code != USGSCSM_B200_OK)
    fail(::csm::Error::SENSOR_MODEL_NOT_CONSTRUCTIBLE, usgscsm_b200_last_error(), "SensorModel::replaceModelState");
  usgscsm_b200_model_free(m_h);
  m_h = h;
  m_meta->state = ub200::parseState(state.data(), state.size());
  m_meta->imageIdentifier = stateString("m_imageIdentifier", "");
  // every parameter of these models is REAL unless the state says otherwise (LS:237-250)
  m_meta->parameterType.assign((size_t)usgscsm_b200_model_num_parameters(m_h), ::csm::param::REAL);
  auto it = m_meta->state.members.find("m_parameterType");
  if (it != m_meta->state.members.end() && it->second.numericArray)
    for (size_t i = 0; i < it->second.nums.size() && i < m_meta->parameterType.size(); i++)
      m_meta->parameterType[i] = (::csm::param::Type)(int)it->second.nums[i];
  // The reference point: line-scan / push-frame derive it in updateState (LS:649-735) -- the
  // engine writes it back into the state; frame: imageToGround of the image centre at height 0,
  // always (FR:1569-1570); SAR: the same when the state holds (0, 0, 0) (SAR:375-390).
  const int kind = usgscsm_b200_model_kind(m_h);
  std::vector<double> ref(3, 0.0);
  if (kind == USGSCSM_B200_MODEL_LINE_SCANNER || kind == USGSCSM_B200_MODEL_PUSH_FRAME) {
    const std::string now = getModelState();
    ub200::json::Value v = ub200::parseState(now.data(), now.size());
    ref = ub200::StateReader(v).getVecDouble("m_referencePointXyz", ref);
  } else {
    ref = ub200::StateReader(m_meta->state).getVecDouble("m_referencePointXyz", ref);
    if (kind == USGSCSM_B200_MODEL_FRAME || (ref[0] == 0 && ref[1] == 0 && ref[2] == 0)) {
      const ::csm::ImageVector sz = getImageSize();
      const ::csm::EcefCoord c = imageToGround(::csm::ImageCoord(sz.line / 2.0, sz.samp / 2.0), 0.0);
      ref[0] = c.x;
      ref[1] = c.y;
      ref[2] = c.z;
    }
  }
  m_meta->referencePoint = ::csm::EcefCoord(ref[0], ref[1], ref[2]);
}

std::string SensorModel::stateString(const char *key, const std::string &fallback) const {
  auto it = m_meta->state.members.find(key);
  return it != m_meta->state.members.end() && it->second.type == ub200::json::Type::String ? it->second.str : fallback;
}
double SensorModel::stateNumber(const char *key, double fallback) const {
  auto it = m_meta->state.members.find(key);
  return it != m_meta->state.members.end() && it->second.isNumber() ? it->second.num : fallback;
}

// csm::Model
::csm::Version SensorModel::getVersion() const { return ::csm::Version(1, 0, 0); }
std::string SensorModel::getModelName() const {
  switch (usgscsm_b200_model_kind(m_h)) {
    case USGSCSM_B200_MODEL_FRAME: return kModelNames[0];
    case USGSCSM_B200_MODEL_LINE_SCANNER: return kModelNames[1];
    case USGSCSM_B200_MODEL_PUSH_FRAME: return kModelNames[2];
    default: return kModelNames[3];
  }
}
std::string SensorModel::getImageIdentifier() const { return m_meta->imageIdentifier; }
void SensorModel::setImageIdentifier(const std::string &imageId, ::csm::WarningList *) {
  m_meta->imageIdentifier = imageId;
}
std::string SensorModel::getSensorIdentifier() const { return stateString("m_sensorIdentifier", stateString("m_sensorName", "")); }
std::string SensorModel::getPlatformIdentifier() const { return stateString("m_platformIdentifier", stateString("m_platformName", "")); }
std::string SensorModel::getCollectionIdentifier() const { return "UNKNOWN"; }  // LS:1813, PF:1603
std::string SensorModel::getTrajectoryIdentifier() const { return "UNKNOWN"; }  // LS:1523, PF:1236
// LS:1527-1537, PF:1243-1254, SAR:1848-1856: the image time of the reference point, added to the
// centre ephemeris time; frame (FR:1186-1190): m_ephemerisTime.  The reference truncates the
// ephemeris time to whole seconds (time_t) before converting.
std::string SensorModel::getReferenceDateAndTime() const {
  if (usgscsm_b200_model_kind(m_h) == USGSCSM_B200_MODEL_FRAME)
    return calendarTime((double)(std::time_t)stateNumber("m_ephemerisTime", 0.0));
  const double relative = getImageTime(groundToImage(getReferencePoint()));
  return calendarTime((double)(std::time_t)(stateNumber("m_centerEphemerisTime", 0.0) + relative));
}
std::string SensorModel::getModelState() const {
  size_t need = 0;
  check(usgscsm_b200_model_get_state(m_h, NULL, 0, &need), "getModelState");
  std::string s(need, '\0');
  if (need) check(usgscsm_b200_model_get_state(m_h, &s[0], need, &need), "getModelState");
  return s;
}
void SensorModel::replaceModelState(const std::string &argState) { load(argState); }

// csm::GeometricModel
::csm::EcefCoord SensorModel::getReferencePoint() const { return m_meta->referencePoint; }
void SensorModel::setReferencePoint(const ::csm::EcefCoord &groundPt) { m_meta->referencePoint = groundPt; }
int SensorModel::getNumParameters() const { return usgscsm_b200_model_num_parameters(m_h); }
std::string SensorModel::getParameterUnits(int) const { return "m"; }  // LS:1575, PF:1346, SAR:1497
bool SensorModel::hasShareableParameters() const { return false; }
bool SensorModel::isParameterShareable(int) const { return false; }
::csm::SharingCriteria SensorModel::getParameterSharingCriteria(int) const { return ::csm::SharingCriteria(); }
double SensorModel::getParameterValue(int index) const {
  double v = 0.0;
  if (usgscsm_b200_model_get_parameter(m_h, index, &v) != 0)
    fail(::csm::Error::INDEX_OUT_OF_RANGE, usgscsm_b200_last_error(), "getParameterValue");
  return v;
}
void SensorModel::setParameterValue(int index, double value) {
  if (usgscsm_b200_model_set_parameter(m_h, index, value) != 0)
    fail(::csm::Error::INDEX_OUT_OF_RANGE, usgscsm_b200_last_error(), "setParameterValue");
}
::csm::param::Type SensorModel::getParameterType(int index) const {
  if (index < 0 || (size_t)index >= m_meta->parameterType.size())
    fail(::csm::Error::INDEX_OUT_OF_RANGE, "Index is out of range.", "getParameterType");
  return m_meta->parameterType[(size_t)index];
}
void SensorModel::setParameterType(int index, ::csm::param::Type pType) {
  if (index < 0 || (size_t)index >= m_meta->parameterType.size())
    fail(::csm::Error::INDEX_OUT_OF_RANGE, "Index is out of range.", "setParameterType");
  m_meta->parameterType[(size_t)index] = pType;
}
double SensorModel::getParameterCovariance(int index1, int index2) const {
  double v = 0.0;
  if (usgscsm_b200_model_get_parameter_covariance(m_h, index1, index2, &v) != 0)
    fail(::csm::Error::INDEX_OUT_OF_RANGE, usgscsm_b200_last_error(), "getParameterCovariance");
  return v;
}
void SensorModel::setParameterCovariance(int index1, int index2, double covariance) {
  if (usgscsm_b200_model_set_parameter_covariance(m_h, index1, index2, covariance) != 0)
    fail(::csm::Error::INDEX_OUT_OF_RANGE, usgscsm_b200_last_error(), "setParameterCovariance");
}
int SensorModel::getNumGeometricCorrectionSwitches() const { return 0; }
std::string SensorModel::getGeometricCorrectionName(int) const {
  fail(::csm::Error::INDEX_OUT_OF_RANGE, "Index is out of range.", "getGeometricCorrectionName");
}
void SensorModel::setGeometricCorrectionSwitch(int, bool, ::csm::param::Type) {
  fail(::csm::Error::INDEX_OUT_OF_RANGE, "Index is out of range.", "setGeometricCorrectionSwitch");
}
bool SensorModel::getGeometricCorrectionSwitch(int) const {
  fail(::csm::Error::INDEX_OUT_OF_RANGE, "Index is out of range.", "getGeometricCorrectionSwitch");
}
// LS:2074-2101: this model's own parameter covariance, zero against any other model.
std::vector<double> SensorModel::getCrossCovarianceMatrix(
    const ::csm::GeometricModel &comparisonModel, ::csm::param::Set pSet,
    const ::csm::GeometricModel::GeometricModelList &) const {
  const std::vector<int> rows = getParameterSetIndices(pSet);
  if (&comparisonModel == this) {
    std::vector<double> c(rows.size() * rows.size(), 0.0);
    for (size_t i = 0; i < rows.size(); i++)
      for (size_t j = 0; j < rows.size(); j++)
        c[i * rows.size() + j] = getParameterCovariance(rows[i], rows[j]);
    return c;
  }
  const std::vector<int> cols = comparisonModel.getParameterSetIndices(pSet);
  return std::vector<double>(rows.size() * cols.size(), 0.0);
}

// ---- csm::RasterGM per-point virtuals: a batch of one ---------------------------------------------
::csm::ImageCoord SensorModel::groundToImage(const ::csm::EcefCoord &g, double prec, double *ach,
                                             ::csm::WarningList *warnings) const {
  double l = 0, s = 0, a = 0;
  uint32_t st = 0;
  check(usgscsm_b200_ground_to_image(m_h, 1, &g.x, &g.y, &g.z, prec, &l, &s, &a, &st), "groundToImage");
  raiseOrWarn(st, warnings, "groundToImage");
  if (ach) *ach = a;
  return ::csm::ImageCoord(l, s);
}
::csm::ImageCoordCovar SensorModel::groundToImage(const ::csm::EcefCoordCovar &g, double prec,
                                                  double *ach, ::csm::WarningList *warnings) const {
  ::csm::ImageCoordCovar r;
  if (usgscsm_b200_model_kind(m_h) == USGSCSM_B200_MODEL_FRAME) {
    // FR:234-278: the covariance is not propagated by the frame model
    const ::csm::ImageCoord p = groundToImage(::csm::EcefCoord(g.x, g.y, g.z), prec, ach, warnings);
    r.line = p.line;
    r.samp = p.samp;
    return r;
  }
  double a = 0;
  uint32_t st = 0;
  check(usgscsm_b200_ground_to_image_covar(m_h, 1, &g.x, &g.y, &g.z, g.covariance, prec, &r.line,
                                           &r.samp, r.covariance, &a, &st),
        "groundToImage");
  raiseOrWarn(st, warnings, "groundToImage");
  if (ach) *ach = a;
  return r;
}
::csm::EcefCoord SensorModel::imageToGround(const ::csm::ImageCoord &p, double height, double prec,
                                            double *ach, ::csm::WarningList *warnings) const {
  double x = 0, y = 0, z = 0, a = 0;
  uint32_t st = 0;
  check(usgscsm_b200_image_to_ground(m_h, 1, &p.line, &p.samp, &height, 1, prec, &x, &y, &z, &a, &st),
        "imageToGround");
  raiseOrWarn(st, warnings, "imageToGround");
  if (ach) *ach = a;
  return ::csm::EcefCoord(x, y, z);
}
::csm::EcefCoordCovar SensorModel::imageToGround(const ::csm::ImageCoordCovar &p, double height,
                                                 double heightVariance, double prec, double *ach,
                                                 ::csm::WarningList *warnings) const {
  ::csm::EcefCoordCovar r;
  if (usgscsm_b200_model_kind(m_h) == USGSCSM_B200_MODEL_FRAME) {  // FR:374-395
    const ::csm::EcefCoord g = imageToGround(::csm::ImageCoord(p.line, p.samp), height, prec, ach, warnings);
    r.x = g.x;
    r.y = g.y;
    r.z = g.z;
    return r;
  }
  double a = 0;
  uint32_t st = 0;
  check(usgscsm_b200_image_to_ground_covar(m_h, 1, &p.line, &p.samp, p.covariance, &height, 1,
                                           &heightVariance, 1, prec, &r.x, &r.y, &r.z, r.covariance,
                                           &a, &st),
        "imageToGround");
  raiseOrWarn(st, warnings, "imageToGround");
  if (ach) *ach = a;
  return r;
}
::csm::EcefLocus SensorModel::imageToProximateImagingLocus(const ::csm::ImageCoord &p,
                                                           const ::csm::EcefCoord &g, double prec,
                                                           double *ach, ::csm::WarningList *warnings) const {
  double o[7] = {0, 0, 0, 0, 0, 0, 0};
  uint32_t st = 0;
  check(usgscsm_b200_image_to_proximate_imaging_locus(m_h, 1, &p.line, &p.samp, &g.x, &g.y, &g.z, prec,
                                                      o, o + 1, o + 2, o + 3, o + 4, o + 5, o + 6, &st),
        "imageToProximateImagingLocus");
  raiseOrWarn(st, warnings, "imageToProximateImagingLocus");
  if (ach) *ach = o[6];
  return ::csm::EcefLocus(o[0], o[1], o[2], o[3], o[4], o[5]);
}
::csm::EcefLocus SensorModel::imageToRemoteImagingLocus(const ::csm::ImageCoord &p, double prec,
                                                        double *ach, ::csm::WarningList *warnings) const {
  double o[7] = {0, 0, 0, 0, 0, 0, 0};
  uint32_t st = 0;
  check(usgscsm_b200_image_to_remote_imaging_locus(m_h, 1, &p.line, &p.samp, prec, o, o + 1, o + 2,
                                                   o + 3, o + 4, o + 5, o + 6, &st),
        "imageToRemoteImagingLocus");
  raiseOrWarn(st, warnings, "imageToRemoteImagingLocus");
  if (ach) *ach = o[6];
  return ::csm::EcefLocus(o[0], o[1], o[2], o[3], o[4], o[5]);
}
::csm::ImageCoord SensorModel::getImageStart() const { return ::csm::ImageCoord(0.0, 0.0); }
::csm::ImageVector SensorModel::getImageSize() const {
  int l = 0, s = 0;
  check(usgscsm_b200_model_image_size(m_h, &l, &s), "getImageSize");
  return ::csm::ImageVector(l, s);
}
std::pair< ::csm::ImageCoord, ::csm::ImageCoord> SensorModel::getValidImageRange() const {
  const ::csm::ImageVector sz = getImageSize();
  return std::make_pair(::csm::ImageCoord(0.0, 0.0), ::csm::ImageCoord(sz.line, sz.samp));
}
std::pair<double, double> SensorModel::getValidHeightRange() const {
  return std::make_pair(stateNumber("m_minElevation", -1000.0), stateNumber("m_maxElevation", 1000.0));
}
::csm::EcefVector SensorModel::getIlluminationDirection(const ::csm::EcefCoord &g) const {
  ::csm::EcefVector d;
  uint32_t st = 0;
  check(usgscsm_b200_illumination_direction(m_h, 1, &g.x, &g.y, &g.z, &d.x, &d.y, &d.z, &st),
        "getIlluminationDirection");
  raiseOrWarn(st, NULL, "getIlluminationDirection");
  return d;
}
double SensorModel::getImageTime(const ::csm::ImageCoord &p) const {
  double t = 0;
  check(usgscsm_b200_image_sensor_state(m_h, 1, &p.line, &p.samp, &t, NULL, NULL, NULL, NULL, NULL, NULL, NULL),
        "getImageTime");
  return t;
}
::csm::EcefCoord SensorModel::getSensorPosition(const ::csm::ImageCoord &p) const {
  ::csm::EcefCoord c;
  check(usgscsm_b200_image_sensor_state(m_h, 1, &p.line, &p.samp, NULL, &c.x, &c.y, &c.z, NULL, NULL, NULL, NULL),
        "getSensorPosition");
  return c;
}
::csm::EcefCoord SensorModel::getSensorPosition(double time) const {
  ::csm::EcefCoord c;
  uint32_t st = 0;
  check(usgscsm_b200_sensor_state_at_time(m_h, 1, &time, &c.x, &c.y, &c.z, NULL, NULL, NULL, &st), "getSensorPosition");
  if (st & USGSCSM_B200_ST_ERROR_OTHER) fail(::csm::Error::BOUNDS, "Valid image time is 0.0", "getSensorPosition");
  return c;
}
::csm::EcefVector SensorModel::getSensorVelocity(const ::csm::ImageCoord &p) const {
  ::csm::EcefVector v;
  uint32_t st = 0;
  check(usgscsm_b200_image_sensor_state(m_h, 1, &p.line, &p.samp, NULL, NULL, NULL, NULL, &v.x, &v.y, &v.z, &st),
        "getSensorVelocity");
  if (st & USGSCSM_B200_ST_ERROR_OTHER) fail(::csm::Error::BOUNDS, "Image coordinate out of bounds.", "getSensorVelocity");
  return v;
}
::csm::EcefVector SensorModel::getSensorVelocity(double time) const {
  ::csm::EcefVector v;
  uint32_t st = 0;
  check(usgscsm_b200_sensor_state_at_time(m_h, 1, &time, NULL, NULL, NULL, &v.x, &v.y, &v.z, &st), "getSensorVelocity");
  if (st & USGSCSM_B200_ST_ERROR_OTHER) fail(::csm::Error::BOUNDS, "Valid image time is 0.0", "getSensorVelocity");
  return v;
}
::csm::RasterGM::SensorPartials SensorModel::computeSensorPartials(int index, const ::csm::EcefCoord &g,
                                                                   double prec, double *ach,
                                                                   ::csm::WarningList *warnings) const {
  double dl = 0, ds = 0, a = 0;
  uint32_t st = 0;
  check(usgscsm_b200_sensor_partials(m_h, index, 1, &g.x, &g.y, &g.z, NULL, NULL, prec, &dl, &ds, &a, &st),
        "computeSensorPartials");
  raiseOrWarn(st, warnings, "computeSensorPartials");
  if (ach) *ach = a;
  return std::make_pair(dl, ds);
}
::csm::RasterGM::SensorPartials SensorModel::computeSensorPartials(int index, const ::csm::ImageCoord &p,
                                                                   const ::csm::EcefCoord &g, double prec,
                                                                   double *ach, ::csm::WarningList *warnings) const {
  double dl = 0, ds = 0, a = 0;
  uint32_t st = 0;
  check(usgscsm_b200_sensor_partials(m_h, index, 1, &g.x, &g.y, &g.z, &p.line, &p.samp, prec, &dl, &ds, &a, &st),
        "computeSensorPartials");
  raiseOrWarn(st, warnings, "computeSensorPartials");
  if (ach) *ach = a;
  return std::make_pair(dl, ds);
}
namespace {
std::vector< ::csm::RasterGM::SensorPartials> allPartials(usgscsm_b200_model *h, const std::vector<int> &idx,
                                                          const ::csm::EcefCoord &g, const ::csm::ImageCoord *p,
                                                          double prec, double *ach, ::csm::WarningList *warnings) {
  std::vector<double> raw(idx.size() * 2 + 2);
  double a = 0;
  uint32_t st = 0;
  if (!idx.empty())
    check(usgscsm_b200_all_sensor_partials(h, (int)idx.size(), idx.data(), 1, &g.x, &g.y, &g.z,
                                           p ? &p->line : NULL, p ? &p->samp : NULL, prec, raw.data(), &a, &st),
          "computeAllSensorPartials");
  raiseOrWarn(st, warnings, "computeAllSensorPartials");
  if (ach) *ach = a;
  std::vector< ::csm::RasterGM::SensorPartials> out(idx.size());
  for (size_t k = 0; k < idx.size(); k++) out[k] = std::make_pair(raw[2 * k], raw[2 * k + 1]);
  return out;
}
}  // namespace
std::vector< ::csm::RasterGM::SensorPartials> SensorModel::computeAllSensorPartials(
    const ::csm::EcefCoord &g, ::csm::param::Set pSet, double prec, double *ach, ::csm::WarningList *warnings) const {
  return allPartials(m_h, getParameterSetIndices(pSet), g, NULL, prec, ach, warnings);
}
std::vector< ::csm::RasterGM::SensorPartials> SensorModel::computeAllSensorPartials(
    const ::csm::ImageCoord &p, const ::csm::EcefCoord &g, ::csm::param::Set pSet, double prec, double *ach,
    ::csm::WarningList *warnings) const {
  return allPartials(m_h, getParameterSetIndices(pSet), g, &p, prec, ach, warnings);
}
std::vector<double> SensorModel::computeGroundPartials(const ::csm::EcefCoord &g) const {
  std::vector<double> p(6, 0.0);
  uint32_t st = 0;
  check(usgscsm_b200_ground_partials(m_h, 1, &g.x, &g.y, &g.z, p.data(), &st), "computeGroundPartials");
  raiseOrWarn(st, NULL, "computeGroundPartials");
  return p;
}
const ::csm::CorrelationModel &SensorModel::getCorrelationModel() const { return m_noCorrelation; }
std::vector<double> SensorModel::getUnmodeledCrossCovariance(const ::csm::ImageCoord &, const ::csm::ImageCoord &) const {
  return std::vector<double>(4, 0.0);  // no unmodeled error (LS:2022-2026)
}
::csm::Ellipsoid SensorModel::getEllipsoid() const {
  double a = 0, b = 0;
  check(usgscsm_b200_model_get_ellipsoid(m_h, &a, &b), "getEllipsoid");
  return ::csm::Ellipsoid(a, b);
}
void SensorModel::setEllipsoid(const ::csm::Ellipsoid &e) {
  check(usgscsm_b200_model_set_ellipsoid(m_h, e.getSemiMajorRadius(), e.getSemiMinorRadius()), "setEllipsoid");
}

// ---- batched calls ------------------------------------------------------------------------------------
void SensorModel::groundToImageBatch(size_t n, const double *x, const double *y, const double *z, double prec,
                                     double *line, double *samp, double *ach, uint32_t *status) const {
  check(usgscsm_b200_ground_to_image(m_h, (int64_t)n, x, y, z, prec, line, samp, ach, status), "groundToImageBatch");
}
void SensorModel::imageToGroundBatch(size_t n, const double *line, const double *samp, const double *height,
                                     bool heightPerPoint, double prec, double *x, double *y, double *z,
                                     double *ach, uint32_t *status) const {
  check(usgscsm_b200_image_to_ground(m_h, (int64_t)n, line, samp, height, heightPerPoint ? 1 : 0, prec, x, y, z,
                                     ach, status),
        "imageToGroundBatch");
}
void SensorModel::groundToImageBatch(const std::vector< ::csm::EcefCoord> &g, double prec,
                                     std::vector< ::csm::ImageCoord> &out, std::vector<uint32_t> *status) const {
  static_assert(sizeof(::csm::EcefCoord) == 3 * sizeof(double) && sizeof(::csm::ImageCoord) == 2 * sizeof(double),
                "CSM value types are plain arrays of doubles");
  out.resize(g.size());
  if (status) status->resize(g.size());
  if (g.empty()) return;
  check(usgscsm_b200_ground_to_image_aos(m_h, (int64_t)g.size(), &g[0].x, prec, &out[0].line, NULL,
                                         status ? status->data() : NULL),
        "groundToImageBatch");
}
void SensorModel::imageToGroundBatch(const std::vector< ::csm::ImageCoord> &p, double height, double prec,
                                     std::vector< ::csm::EcefCoord> &out, std::vector<uint32_t> *status) const {
  out.resize(p.size());
  if (status) status->resize(p.size());
  if (p.empty()) return;
  check(usgscsm_b200_image_to_ground_aos(m_h, (int64_t)p.size(), &p[0].line, &height, 0, prec, &out[0].x, NULL,
                                         status ? status->data() : NULL),
        "imageToGroundBatch");
}

// ---- per-model metadata -----------------------------------------------------------------------------------
namespace {
const char *const kOrbitParams[16] = {  // LS:47-64, PF:46-63
    "IT Pos. Bias   ", "CT Pos. Bias   ", "Rad Pos. Bias  ", "IT Vel. Bias   ", "CT Vel. Bias   ", "Rad Vel. Bias  ",
    "Omega Bias     ", "Phi Bias       ", "Kappa Bias     ", "Omega Rate     ", "Phi Rate       ", "Kappa Rate     ",
    "Omega Accl     ", "Phi Accl       ", "Kappa Accl     ", "Focal Bias     "};
const char *const kFrameParams[7] = {"X Sensor Position (m)", "Y Sensor Position (m)", "Z Sensor Position (m)",
                                     "w", "v1", "v2", "v3"};  // FR:21-29
const char *const kSarParams[6] = {"X Pos. Bias   ", "Y Pos. Bias   ", "Z Pos. Bias   ",
                                   "X Vel. Bias   ", "Y Vel. Bias   ", "Z Vel. Bias   "};  // SAR:20-27
std::string nameOf(const char *const *names, int count, int index, const char *fn) {
  if (index < 0 || index >= count) fail(::csm::Error::INDEX_OUT_OF_RANGE, "Index is out of range.", fn);
  return names[index];
}
}  // namespace
std::string LineScannerModel::getPedigree() const { return "USGS_LINE_SCANNER"; }  // LS:1783
std::string LineScannerModel::getSensorType() const { return CSM_SENSOR_TYPE_EO; }
std::string LineScannerModel::getSensorMode() const { return CSM_SENSOR_MODE_PB; }
std::string LineScannerModel::getParameterName(int i) const { return nameOf(kOrbitParams, 16, i, "getParameterName"); }
std::string PushFrameModel::getPedigree() const { return "USGS_LINE_SCANNER"; }  // PF:1380 (sic)
std::string PushFrameModel::getSensorType() const { return CSM_SENSOR_TYPE_EO; }
std::string PushFrameModel::getSensorMode() const { return CSM_SENSOR_MODE_PB; }
std::string PushFrameModel::getParameterName(int i) const { return nameOf(kOrbitParams, 16, i, "getParameterName"); }
std::string FrameModel::getPedigree() const { return "USGS_FRAMER"; }  // FR:1074
std::string FrameModel::getSensorType() const { return CSM_SENSOR_TYPE_EO; }
std::string FrameModel::getSensorMode() const { return CSM_SENSOR_MODE_FRAME; }
std::string FrameModel::getParameterName(int i) const { return nameOf(kFrameParams, 7, i, "getParameterName"); }
std::string FrameModel::getParameterUnits(int i) const { return i < 3 ? "m" : "radians"; }  // FR:1960-1966
std::string FrameModel::getCollectionIdentifier() const { return stateString("m_collectionIdentifier", ""); }
std::string FrameModel::getTrajectoryIdentifier() const { return ""; }
std::string SarModel::getPedigree() const { return "USGS_SAR"; }  // SAR:1754
std::string SarModel::getSensorType() const { return "SAR"; }
std::string SarModel::getSensorMode() const { return "STRIP"; }
std::string SarModel::getParameterName(int i) const { return nameOf(kSarParams, 6, i, "getParameterName"); }
std::string SarModel::getCollectionIdentifier() const { return stateString("m_collectionIdentifier", ""); }
std::string SarModel::getTrajectoryIdentifier() const { return stateString("m_trajectoryIdentifier", ""); }

// ---- the plugin ---------------------------------------------------------------------------------------------
// The static instance registers the plugin with csm::Plugin's list when the library is loaded
// (csm/Plugin.cpp:36-117), exactly like src/UsgsAstroPlugin.cpp:72.
const UsgsAstroPlugin UsgsAstroPlugin::m_registeredPlugin;

UsgsAstroPlugin::UsgsAstroPlugin() : m_device(deviceFromEnv()) {}
UsgsAstroPlugin::~UsgsAstroPlugin() {}
std::string UsgsAstroPlugin::getPluginName() const { return "UsgsAstroPluginCSM"; }
std::string UsgsAstroPlugin::getManufacturer() const { return "UsgsAstrogeology"; }
std::string UsgsAstroPlugin::getReleaseDate() const { return "20190222"; }
::csm::Version UsgsAstroPlugin::getCsmVersion() const { return ::csm::Version(3, 0, 3); }
size_t UsgsAstroPlugin::getNumModels() const { return 4; }  // the PROJ-backed projected wrapper is out of scope
std::string UsgsAstroPlugin::getModelName(size_t i) const {
  if (i >= 4) fail(::csm::Error::INDEX_OUT_OF_RANGE, "Model index out of range.", "UsgsAstroPlugin::getModelName");
  return kModelNames[i];
}
std::string UsgsAstroPlugin::getModelFamily(size_t) const { return CSM_RASTER_FAMILY; }
::csm::Version UsgsAstroPlugin::getModelVersion(const std::string &) const { return ::csm::Version(1, 0, 0); }

std::string UsgsAstroPlugin::getModelNameFromModelState(const std::string &modelState, ::csm::WarningList *) const {
  char name[128];
  const int code = usgscsm_b200_model_name_from_state(modelState.data(), modelState.size(), name, sizeof name);
  if (code == USGSCSM_B200_EUNSUPPORTED)
    fail(::csm::Error::SENSOR_MODEL_NOT_SUPPORTED, "Sensor model not supported.", "UsgsAstroPlugin::getModelNameFromModelState");
  if (code != USGSCSM_B200_OK)
    fail(::csm::Error::INVALID_SENSOR_MODEL_STATE, "No 'm_modelName' key in the model state object.",
         "UsgsAstroPlugin::getModelNameFromModelState");
  return name;
}
bool UsgsAstroPlugin::canModelBeConstructedFromState(const std::string &modelName, const std::string &modelState,
                                                     ::csm::WarningList *warnings) const {
  try {
    ::csm::Model *m = constructModelFromState(modelState, warnings);
    const bool same = m->getModelName() == modelName;
    delete m;
    return same;
  } catch (const std::exception &e) {
    if (warnings)
      warnings->push_back(::csm::Warning(::csm::Warning::UNKNOWN_WARNING,
                                         "Could not create model [" + modelName + "] with error [" + e.what() + "]",
                                         "UsgsAstroPlugin::canModelBeConstructedFromState()"));
    return false;
  }
}
::csm::Model *UsgsAstroPlugin::constructModelFromState(const std::string &modelState, ::csm::WarningList *) const {
  const std::string name = getModelNameFromModelState(modelState);
  if (name == kModelNames[0]) return new FrameModel(modelState, m_device);
  if (name == kModelNames[1]) return new LineScannerModel(modelState, m_device);
  if (name == kModelNames[2]) return new PushFrameModel(modelState, m_device);
  return new SarModel(modelState, m_device);
}
// :324-353: the csm::Isd names the IMAGE; the ISD is the sidecar "<base name>.json".
std::string UsgsAstroPlugin::loadImageSupportData(const ::csm::Isd &isd, std::string *identifier) const {
  const std::string image = isd.filename();
  std::string base = image.substr(0, image.find_last_of("."));
  if (identifier) *identifier = base.substr(base.find_last_of("/") + 1);
  std::ifstream f((base + ".json").c_str(), std::ios::binary);
  if (!f)
    fail(::csm::Error::FILE_READ, "Could not read metadata file associated with image [" + base + ".json]",
         "UsgsAstroPlugin::loadImageSupportData");
  std::stringstream ss;
  ss << f.rdbuf();
  return ss.str();
}
std::string UsgsAstroPlugin::convertISDToModelState(const ::csm::Isd &isd, const std::string &modelName,
                                                    ::csm::WarningList *warnings) const {
  ::csm::Model *m = constructModelFromISD(isd, modelName, warnings);
  const std::string state = m->getModelState();
  delete m;
  return state;
}
std::string UsgsAstroPlugin::getStateFromISD(::csm::Isd isd) const { return convertISDToModelState(isd, ""); }
::csm::Model *UsgsAstroPlugin::constructModelFromISD(const ::csm::Isd &isd, const std::string &modelName,
                                                     ::csm::WarningList *warnings) const {
  std::string ident;
  const std::string text = loadImageSupportData(isd, &ident);
  size_t need = 0;
  const char *name = modelName.empty() ? NULL : modelName.c_str();
  if (usgscsm_b200_state_from_isd(text.data(), text.size(), name, ident.c_str(), NULL, 0, &need) != USGSCSM_B200_OK)
    fail(::csm::Error::SENSOR_MODEL_NOT_CONSTRUCTIBLE,
         "Could not construct model [" + modelName + "] with error [" + usgscsm_b200_last_error() + "]",
         "UsgsAstroPlugin::constructModelFromISD");
  std::string state(need, '\0');
  usgscsm_b200_state_from_isd(text.data(), text.size(), name, ident.c_str(), &state[0], need, &need);
  return constructModelFromState(state, warnings);
}
bool UsgsAstroPlugin::canModelBeConstructedFromISD(const ::csm::Isd &isd, const std::string &modelName,
                                                   ::csm::WarningList *warnings) const {
  try {
    delete constructModelFromISD(isd, modelName, warnings);
    return true;
  } catch (const std::exception &e) {
    if (warnings)
      warnings->push_back(::csm::Warning(::csm::Warning::UNKNOWN_WARNING,
                                         "Could not create model [" + modelName + "] with error [" + e.what() + "]",
                                         "UsgsAstroPlugin::canModelBeConstructedFromISD()"));
    return false;
  }
}
bool UsgsAstroPlugin::canISDBeConvertedToModelState(const ::csm::Isd &isd, const std::string &modelName,
                                                    ::csm::WarningList *warnings) const {
  try {
    convertISDToModelState(isd, modelName, warnings);
    return true;
  } catch (const std::exception &e) {
    if (warnings)
      warnings->push_back(::csm::Warning(::csm::Warning::UNKNOWN_WARNING,
                                         "Could not create model [" + modelName + "] state with error [" + e.what() + "]",
                                         "UsgsAstroPlugin::canISDBeConvertedToModelState()"));
    return false;
  }
}

}  // namespace usgscsm_b200
