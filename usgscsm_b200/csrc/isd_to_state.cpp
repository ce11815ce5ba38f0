// isd_to_state.cpp -- ISD (ALE-format image support data, JSON) -> model state (host only).
//
// Replaces <Model>::constructStateFromIsd of the four sensor models:
//   line-scan  src/UsgsAstroLsSensorModel.cpp:2818-3152
//   push-frame src/UsgsAstroPushFrameSensorModel.cpp:2114-2463
//   frame      src/UsgsAstroFrameSensorModel.cpp:1660-1903
//   SAR        src/UsgsAstroSarSensorModel.cpp:52-230
// and the ISD accessors they call (ALE's Util.h; src/Utilities.cpp:1266-1400, 2317-2480 for the
// SAR keys).  The reference builds a JSON object and hands it to populateModel through
// variantMapFromJson; here the same object is built as a json::Value and returned as text, so
// that usgscsm_b200_model_from_state (or the reference itself) loads it.  The ephemeris /
// attitude resampling lives in isd_math.h.
//
// What ends up in the state, per model (units: the ISD is in km, the state in m):
//   LS / PF  positions + velocities resampled to >= 25 uniformly spaced nodes over the ISD's
//            ephemeris span (cubic Hermite), rotated J2000 -> body; sensor -> body quaternions on
//            the merged attitude grid, re-sampled uniformly (the model assumes a uniform table:
//            dtQuat = span / (n - 1)); sun table on the ephemeris grid;
//   frame    one position / velocity / quaternion (the exposure), sun positions as given;
//   SAR      positions + velocities on the ISD's own node count, uniformly re-spaced.
#include <cmath>
#include <cstring>
#include <iterator>
#include <string>
#include <vector>

#include "isd_math.h"
#include "json_min.h"
#include "model.h"
#include "state_reader.h"

namespace ub200 {
namespace isd {

namespace {

[[noreturn]] void bad(const std::string &what) { throw StateError("ISD: " + what); }

// ---- typed access to the ISD -------------------------------------------------------------------
const json::Value &member(const json::Value &obj, const char *key, const char *what) {
  if (obj.type != json::Type::Object) bad(std::string("could not parse the ") + what);
  auto it = obj.members.find(key);
  if (it == obj.members.end()) bad(std::string("could not parse the ") + what + " (no \"" + key + "\")");
  return it->second;
}
double number(const json::Value &v, const char *what) {
  if (!v.isNumber()) bad(std::string("could not parse the ") + what);
  return v.num;
}
int integer(const json::Value &v, const char *what) { return (int)number(v, what); }  // get<int>()
std::string text(const json::Value &v, const char *what) {
  if (v.type != json::Type::String) bad(std::string("could not parse the ") + what);
  return v.str;
}
std::vector<double> numbers(const json::Value &v, const char *what) {
  if (v.type != json::Type::Array || (v.arraySize() && !v.numericArray))
    bad(std::string("could not parse the ") + what);
  return v.nums;
}
std::vector<std::vector<double>> rows(const json::Value &v, const char *what) {
  if (v.type != json::Type::Array) bad(std::string("could not parse the ") + what);
  std::vector<std::vector<double>> out;
  if (v.numericArray && v.arraySize()) bad(std::string("could not parse the ") + what);
  for (const json::Value &r : v.items) out.push_back(numbers(r, what));
  return out;
}
std::vector<V3> vectors(const json::Value &v, const char *what) {
  std::vector<V3> out;
  for (const std::vector<double> &r : rows(v, what)) {
    if (r.size() != 3) bad(std::string("could not parse the ") + what);
    out.push_back({r[0], r[1], r[2]});
  }
  return out;
}

Ephemeris ephemeris(const json::Value &isd, const char *key, const char *what) {
  const json::Value &s = member(isd, key, what);
  Ephemeris e;
  e.pos = vectors(member(s, "positions", what), what);
  e.times = numbers(member(s, "ephemeris_times", what), what);
  number(member(s, "reference_frame", what), what);
  if (s.members.count("velocities")) e.vel = vectors(s.members.at("velocities"), what);
  if (e.times.size() != e.pos.size() || (!e.vel.empty() && e.vel.size() != e.pos.size()))
    bad(std::string("the ") + what + " has ragged tables");
  return e;
}

Attitude attitude(const json::Value &isd, const char *key, const char *what) {
  const json::Value &o = member(isd, key, what);
  Attitude a;
  for (const std::vector<double> &q : rows(member(o, "quaternions", what), what)) {
    if (q.size() != 4) bad(std::string("could not parse the ") + what);
    a.rot.push_back(unit(q[0], q[1], q[2], q[3]));  // scalar first
  }
  a.times = numbers(member(o, "ephemeris_times", what), what);
  if (o.members.count("angular_velocities"))
    a.av = vectors(o.members.at("angular_velocities"), what);
  if (o.members.count("constant_rotation"))
    a.constant = fromMatrix(numbers(o.members.at("constant_rotation"), what));
  if (a.rot.empty() || a.rot.size() != a.times.size() || (!a.av.empty() && a.av.size() != a.times.size()))
    bad(std::string("the ") + what + " needs as many rotations (>= 1) as times");
  return a;
}

// "optical_distortion": the model is named by the (alphabetically first) key; the coefficient
// layout per model is the one applyDistortion / removeDistortion expect (src/Distortion.cpp).
void distortion(const json::Value &isd, int &type, std::vector<double> &coeffs) {
  const json::Value &od = member(isd, "optical_distortion", "distortion model");
  if (od.type != json::Type::Object || od.members.empty()) bad("could not parse the distortion model");
  const std::string name = od.members.begin()->first;
  const json::Value &m = od.members.begin()->second;
  const char *what = "distortion model coefficients";
  auto padded = [&](const char *key, size_t n) {
    std::vector<double> v = numbers(member(m, key, what), what);
    v.resize(n, 0.0);
    return v;
  };
  if (name == "transverse") {
    type = DIST_TRANSVERSE;
    coeffs = padded("x", 10);
    std::vector<double> y = padded("y", 10);
    coeffs.insert(coeffs.end(), y.begin(), y.end());
  } else if (name == "kaguyalism") {
    type = DIST_KAGUYALISM;
    coeffs = padded("x", 4);
    std::vector<double> y = padded("y", 4);
    coeffs.insert(coeffs.begin(), number(member(m, "boresight_x", what), what));
    y.insert(y.begin(), number(member(m, "boresight_y", what), what));
    coeffs.insert(coeffs.end(), y.begin(), y.end());
  } else if (name == "lunarorbiter") {
    type = DIST_LUNARORBITER;
    for (const char *k : {"perspective_x", "perspective_y", "center_point_x", "center_point_y"})
      coeffs.push_back(number(member(m, k, what), what));
  } else {
    static const struct { const char *name; int type; } kinds[] = {
        {"radial", DIST_RADIAL},         {"dawnfc", DIST_DAWNFC}, {"lrolrocnac", DIST_LROLROCNAC},
        {"cahvor", DIST_CAHVOR},         {"radtan", DIST_RADTAN}, {"kploshadowcam", DIST_KPLOSHADOWCAM}};
    type = DIST_TRANSVERSE;  // what the reference falls back to for an unknown name (U:1897)
    for (const auto &k : kinds)
      if (name == k.name) {
        type = k.type;
        coeffs = numbers(member(m, "coefficients", what), what);
      }
  }
}

// ---- building the state object -------------------------------------------------------------------------
struct State {
  json::Value v;
  State() { v.type = json::Type::Object; }
  void num(const char *k, double d) { v.members[k] = json::number(d); }
  void integer(const char *k, long i) {
    json::Value n;
    n.type = json::Type::Int;
    n.num = (double)i;
    v.members[k] = n;
  }
  void str(const char *k, const std::string &s) {
    json::Value n;
    n.type = json::Type::String;
    n.str = s;
    v.members[k] = n;
  }
  void arr(const char *k, const std::vector<double> &a) { v.members[k] = json::numberArray(a); }
  void raw(const char *k, const json::Value &x) { v.members[k] = x; }
};

void push(std::vector<double> &out, const V3 &v, double scale) {
  out.push_back(v.x * scale);
  out.push_back(v.y * scale);
  out.push_back(v.z * scale);
}

// Fields every model copies the same way.
void identity(State &st, const json::Value &isd) {
  st.str("m_modelName", text(member(isd, "name_model", "sensor model name"), "sensor model name"));
  st.str("m_imageIdentifier", text(member(isd, "image_identifier", "image identifier"), "image identifier"));
  st.str("m_sensorName", text(member(isd, "name_sensor", "sensor name"), "sensor name"));
  st.str("m_platformName", text(member(isd, "name_platform", "platform name"), "platform name"));
  st.integer("m_nLines", integer(member(isd, "image_lines", "number of lines"), "number of lines"));
  st.integer("m_nSamples", integer(member(isd, "image_samples", "number of samples"), "number of samples"));
}
void bodyAndHeights(State &st, const json::Value &isd) {
  const json::Value &radii = member(isd, "radii", "reference ellipsoid radii");
  st.num("m_minorAxis", number(member(radii, "semiminor", "semiminor radius"), "semiminor radius") * 1000);
  st.num("m_majorAxis", number(member(radii, "semimajor", "semimajor radius"), "semimajor radius") * 1000);
  const json::Value &h = member(isd, "reference_height", "reference height");
  st.num("m_minElevation", number(member(h, "minheight", "minimum height"), "minimum height"));
  st.num("m_maxElevation", number(member(h, "maxheight", "maximum height"), "maximum height"));
}
void detector(State &st, const json::Value &isd) {
  st.integer("m_detectorSampleSumming", integer(member(isd, "detector_sample_summing", "sample summing"), "sample summing"));
  st.integer("m_detectorLineSumming", integer(member(isd, "detector_line_summing", "line summing"), "line summing"));
  st.num("m_startingDetectorSample", number(member(isd, "starting_detector_sample", "detector starting sample"), "detector starting sample"));
  st.num("m_startingDetectorLine", number(member(isd, "starting_detector_line", "detector starting line"), "detector starting line"));
  st.num("m_focalLength", number(member(member(isd, "focal_length_model", "focal length"), "focal_length", "focal length"), "focal length"));
  st.arr("m_iTransS", numbers(member(isd, "focal2pixel_samples", "focal plane to detector samples transformation"), "focal2pixel_samples"));
  st.arr("m_iTransL", numbers(member(isd, "focal2pixel_lines", "focal plane to detector lines transformation"), "focal2pixel_lines"));
  int type = 0;
  std::vector<double> coeffs;
  distortion(isd, type, coeffs);
  st.integer("m_distortionType", type);
  st.arr("m_opticalDistCoeffs", coeffs);
}
// integerOnes: the frame model writes the diagonal as the integer literal 1 (FR:1879-1882), which
// makes the reference type the whole array as vector<int> (U:2760-2768); harmless for 0 / 1.
void identityCovariance(State &st, const char *key, int n, bool integerOnes = false) {
  std::vector<double> c((size_t)n * n, 0.0);
  for (int i = 0; i < n; i++) c[(size_t)i * n + i] = 1.0;
  st.arr(key, c);
  if (integerOnes) {
    json::Value &v = st.v.members[key];
    for (int i = 0; i < n; i++) v.numIsInt[(size_t)i * n + i] = 1;
    v.firstIsInt = true;
  }
}

// Uniform grid over [first, last] with `count` nodes: first + i * step.
std::vector<double> uniformTimes(double first, double last, size_t count, double &step) {
  step = (last - first) / (double)(count - 1);
  std::vector<double> t(count);
  for (size_t i = 0; i < count; i++) t[i] = first + (double)i * step;
  return t;
}

// Instrument (or sun) states at `times`: interpolate in J2000, rotate into the body frame, km -> m.
void bodyFixedStates(const Ephemeris &eph, const Attitude &j2000ToTarget,
                     const std::vector<double> &times, std::vector<double> &positions,
                     std::vector<double> &velocities) {
  for (double t : times) {
    V3 p, v, rp, rv;
    eph.at(t, p, v);
    j2000ToTarget.rotateState(t, p, v, rp, rv);
    push(positions, rp, 1000);
    push(velocities, rv, 1000);
  }
}

// Sensor -> body quaternions in the state's (x, y, z, w) order at the given times.
std::vector<double> stateQuaternions(const Attitude &s2t, const std::vector<double> &times) {
  std::vector<double> out;
  for (double t : times) {
    const Quat q = s2t.full(t);
    out.push_back(q.x);
    out.push_back(q.y);
    out.push_back(q.z);
    out.push_back(q.w);
  }
  return out;
}

// ---- line-scan and push-frame (they share everything but the timing keys) ------------------------------------
State scannerState(const json::Value &isd, bool pushFrame) {
  State st;
  identity(st, isd);
  detector(st, isd);
  st.integer("m_platformFlag", 1);
  st.integer("m_ikCode", 0);
  st.integer("m_zDirection", 1);
  st.arr("m_referencePointXyz", {0.0, 0.0, 0.0});
  const double center = number(member(isd, "center_ephemeris_time", "center image time"), "center image time");
  st.num("m_centerEphemerisTime", center);
  st.num("m_startingEphemerisTime", number(member(isd, "starting_ephemeris_time", "image start time"), "image start time"));
  const json::Value &dc = member(isd, "detector_center", "detector center");
  st.num("m_detectorSampleOrigin", number(member(dc, "sample", "detector center sample"), "detector center sample"));
  st.num("m_detectorLineOrigin", number(member(dc, "line", "detector center line"), "detector center line"));

  // ephemeris: >= 25 uniformly spaced nodes over the ISD's span (Lagrange order 8 needs them)
  const Ephemeris inst = ephemeris(isd, "instrument_position", "instrument position");
  const Ephemeris sun = ephemeris(isd, "sun_position", "sun position");
  const Attitude j2000ToTarget = attitude(isd, "body_rotation", "body rotation");
  const Attitude j2000ToSensor = attitude(isd, "instrument_pointing", "instrument pointing");
  if (inst.times.empty()) bad("the instrument position is empty");
  double dtEphem;
  const std::vector<double> ephemTimes = uniformTimes(
      inst.times.front(), inst.times.back(), inst.times.size() < 25 ? 25 : inst.times.size(), dtEphem);
  st.num("m_dtEphem", dtEphem);
  st.num("m_t0Ephem", inst.times.front() - center);
  std::vector<double> positions, velocities, sunPositions, sunVelocities;
  bodyFixedStates(sun, j2000ToTarget, ephemTimes, sunPositions, sunVelocities);
  bodyFixedStates(inst, j2000ToTarget, ephemTimes, positions, velocities);
  st.arr("m_sunPosition", sunPositions);
  st.arr("m_sunVelocity", sunVelocities);
  st.arr("m_positions", positions);
  st.integer("m_numPositions", (long)positions.size());
  st.arr("m_velocities", velocities);

  // attitude: sensor -> body on the merged grid, then uniformly re-sampled with the same count
  const Attitude s2t = sensorToTarget(j2000ToTarget, j2000ToSensor);
  double dtQuat;
  const std::vector<double> quatTimes =
      uniformTimes(s2t.times.front(), s2t.times.back(), s2t.times.size(), dtQuat);
  st.num("m_dtQuat", dtQuat);
  st.num("m_t0Quat", s2t.times.front() - center);
  const std::vector<double> quats = stateQuaternions(s2t, quatTimes);
  st.arr("m_quaternions", quats);
  st.integer("m_numQuaternions", (long)quats.size());

  // derived by the model at load (updateState LS:649-735); these values ask for it (LS:283-285)
  st.num("m_gsd", 1.0);
  st.num("m_flyingHeight", 1000.0);
  st.num("m_halfSwath", 1000.0);
  st.num("m_halfTime", 10.0);
  st.arr("m_currentParameterValue", std::vector<double>(16, 0.0));
  identityCovariance(st, "m_covariance", 16);
  json::Value types;
  types.type = json::Type::Array;
  for (int i = 0; i < 16; i++) {
    json::Value s;
    s.type = json::Type::String;
    s.str = "REAL";
    types.items.push_back(s);
  }
  st.raw("m_parameterType", types);
  bodyAndHeights(st, isd);
  st.str("m_platformIdentifier", st.v.members["m_platformName"].str);
  st.str("m_sensorIdentifier", st.v.members["m_sensorName"].str);

  if (!pushFrame) {
    // line_scan_rate rows: [first line, start time relative to the centre, exposure per line]
    std::vector<double> lines, starts, rates;
    for (const std::vector<double> &r : rows(member(isd, "line_scan_rate", "integration time table"), "integration time table")) {
      if (r.size() != 3) bad("a line_scan_rate entry does not hold 3 values");
      lines.push_back(r[0]);
      starts.push_back(r[1]);
      rates.push_back(r[2]);
    }
    st.arr("m_intTimeLines", lines);
    st.arr("m_intTimeStartTimes", starts);
    st.arr("m_intTimes", rates);
    return st;
  }
  // framelet timing and geometry (PF:2230-2260, 2432-2444)
  st.raw("m_exposureDuration", member(isd, "exposure_duration", "framelet timing information"));
  st.raw("m_interframeDelay", member(isd, "interframe_delay", "framelet timing information"));
  st.raw("m_frameletHeight", member(isd, "framelet_height", "framelet geometry"));
  st.raw("m_frameletsFlipped", member(isd, "framelets_flipped", "framelet geometry"));
  st.raw("m_frameletsOrderReversed", member(isd, "framelet_order_reversed", "framelet geometry"));
  if (isd.members.count("num_lines_overlap")) st.raw("m_numLinesOverlap", isd.members.at("num_lines_overlap"));
  else st.integer("m_numLinesOverlap", 0);
  const int frameletHeight = integer(st.v.members["m_frameletHeight"], "framelet height");
  const int overlap = integer(st.v.members["m_numLinesOverlap"], "framelet overlap");
  if (frameletHeight <= 0) bad("framelet_height must be positive");
  const int nLines = (int)st.v.members["m_nLines"].num;
  st.integer("m_reducedframeletHeight", frameletHeight - overlap);
  st.integer("m_nReducedLines", (nLines / frameletHeight) * (frameletHeight - overlap));
  return st;
}

// ---- frame ---------------------------------------------------------------------------------------------------
State frameState(const json::Value &isd) {
  State st;
  identity(st, isd);
  detector(st, isd);
  const json::Value &flm = member(isd, "focal_length_model", "focal length");
  st.num("m_focalLengthEpsilon",
         flm.members.count("focal_uncertainty") ? number(flm.members.at("focal_uncertainty"), "focal length uncertainty") : 0.0);
  const Ephemeris inst = ephemeris(isd, "instrument_position", "instrument position");
  const Ephemeris sun = ephemeris(isd, "sun_position", "sun position");
  const Attitude j2000ToTarget = attitude(isd, "body_rotation", "body rotation");
  const Attitude j2000ToSensor = attitude(isd, "instrument_pointing", "instrument pointing");
  // one exposure: exactly one position, velocity and quaternion
  std::vector<double> positions, velocities, sunPositions;
  for (size_t i = 0; i < inst.times.size(); i++) {
    V3 rp, rv;
    const V3 v = inst.vel.empty() ? V3{std::nan(""), std::nan(""), std::nan("")} : inst.vel[i];
    j2000ToTarget.rotateState(inst.times[i], inst.pos[i], v, rp, rv);
    push(positions, rp, 1000);
    push(velocities, rv, 1000);
  }
  if (positions.size() != 3) bad("Sensor position does not have 3 values.");
  if (velocities.size() != 3) bad("Sensor velocity does not have 3 values.");
  for (size_t i = 0; i < sun.times.size(); i++) {
    V3 rp, rv;
    const V3 v = sun.vel.empty() ? V3{} : sun.vel[i];
    j2000ToTarget.rotateState(sun.times[i], sun.pos[i], v, rp, rv);
    push(sunPositions, rp, 1000);
  }
  st.arr("m_spacecraftVelocity", velocities);
  st.arr("m_sunPosition", sunPositions);
  const Attitude s2t = sensorToTarget(j2000ToTarget, j2000ToSensor);
  const std::vector<double> quats = stateQuaternions(s2t, s2t.times);
  if (quats.size() != 4) bad("Sensor orientation quaternion does not have 4 values.");
  std::vector<double> params(positions);
  params.insert(params.end(), quats.begin(), quats.end());
  st.arr("m_currentParameterValue", params);

  const json::Value &dc = member(isd, "detector_center", "detector center");
  st.arr("m_ccdCenter", {number(member(dc, "line", "detector center line"), "detector center line"),
                         number(member(dc, "sample", "detector center sample"), "detector center sample")});
  bodyAndHeights(st, isd);
  st.num("m_ephemerisTime", number(member(isd, "center_ephemeris_time", "center image time"), "center image time"));
  if (isd.members.count("jitter")) {  // rolling-shutter jitter polynomials, copied as they are
    const json::Value &j = isd.members.at("jitter");
    st.raw("m_lineJitter", member(j, "lineJitterCoefficients", "jitter"));
    st.raw("m_sampleJitter", member(j, "sampleJitterCoefficients", "jitter"));
    st.raw("m_lineTimes", member(j, "lineExposureTimes", "jitter"));
  }
  // pixel -> focal plane: inverse of the focal plane -> pixel affine (FR:1856-1871)
  const std::vector<double> &L = st.v.members["m_iTransL"].nums, &S = st.v.members["m_iTransS"].nums;
  if (L.size() < 3 || S.size() < 3) bad("Could not compute detector pixel to focal plane coordinate transformation.");
  const double det = L[1] * S[2] - L[2] * S[1];
  std::vector<double> tx(3), ty(3);
  tx[1] = L[1] / det;
  tx[2] = -S[1] / det;
  tx[0] = -(tx[1] * L[0] + tx[2] * S[0]);
  ty[1] = -L[2] / det;
  ty[2] = S[2] / det;
  ty[0] = -(ty[1] * L[0] + ty[2] * S[0]);
  st.arr("m_transX", tx);
  st.arr("m_transY", ty);
  st.arr("m_referencePointXyz", {0.0, 0.0, 0.0});
  identityCovariance(st, "m_currentParameterCovariance", 7, true);
  st.str("m_collectionIdentifier", "");
  return st;
}

// ---- SAR ---------------------------------------------------------------------------------------------------
State sarState(const json::Value &isd) {
  State st;
  identity(st, isd);
  st.arr("m_referencePointXyz", {0.0, 0.0, 0.0});
  const Ephemeris inst = ephemeris(isd, "instrument_position", "instrument position");
  const Ephemeris sun = ephemeris(isd, "sun_position", "sun position");
  const Attitude j2000ToTarget = attitude(isd, "body_rotation", "body rotation");
  if (inst.times.size() < 2) bad("Instrument position has less than 2 states.");
  double dtEphem;
  const std::vector<double> ephemTimes =
      uniformTimes(inst.times.front(), inst.times.back(), inst.times.size(), dtEphem);
  st.num("m_t0Ephem", inst.times.front());  // absolute here, unlike the optical models (SAR:95)
  st.num("m_dtEphem", dtEphem);
  std::vector<double> positions, velocities, sunPositions, sunVelocities;
  bodyFixedStates(sun, j2000ToTarget, ephemTimes, sunPositions, sunVelocities);
  bodyFixedStates(inst, j2000ToTarget, ephemTimes, positions, velocities);
  st.arr("m_sunPosition", sunPositions);
  st.arr("m_sunVelocity", sunVelocities);
  st.arr("m_positions", positions);
  st.arr("m_velocities", velocities);
  st.num("m_startingEphemerisTime", number(member(isd, "starting_ephemeris_time", "image start time"), "image start time"));
  st.num("m_centerEphemerisTime", number(member(isd, "center_ephemeris_time", "center image time"), "center image time"));
  st.num("m_endingEphemerisTime", number(member(isd, "ending_ephemeris_time", "ending image time"), "ending image time"));
  st.num("m_exposureDuration", number(member(isd, "line_exposure_duration", "line exposure duration"), "line exposure duration"));
  st.num("m_scaledPixelWidth", number(member(isd, "scaled_pixel_width", "scaled pixel width"), "scaled pixel width"));
  st.num("m_wavelength", number(member(isd, "wavelength", "wavelength"), "wavelength"));
  st.str("m_lookDirection", text(member(isd, "look_direction", "look direction"), "look direction"));
  std::vector<double> coeffs;
  for (const std::vector<double> &r : rows(member(isd, "range_conversion_coefficients", "range conversion coefficients"), "range conversion coefficients")) {
    if (r.size() < 4) bad("Could not parse the range conversion coefficients.");
    coeffs.insert(coeffs.end(), r.begin(), r.begin() + 4);
  }
  st.arr("m_scaleConversionCoefficients", coeffs);
  st.arr("m_scaleConversionTimes", numbers(member(isd, "range_conversion_times", "range conversion times"), "range conversion times"));
  st.arr("m_currentParameterValue", std::vector<double>(6, 0.0));
  identityCovariance(st, "m_covariance", 6);
  bodyAndHeights(st, isd);
  st.str("m_platformIdentifier", st.v.members["m_platformName"].str);
  st.str("m_sensorIdentifier", st.v.members["m_sensorName"].str);
  return st;
}

}  // namespace

// ISD text -> "<model name>\n<state JSON>".  modelName empty = the ISD's own "name_model";
// otherwise it must match (the plugin's constructModelFromISD dispatches on its argument,
// src/Utilities.cpp:2814-2860, and each model's converter rejects a foreign ISD).
std::string stateFromIsd(const char *isdText, size_t len, const std::string &modelName,
                         const std::string &imageIdentifier) {
  json::Value isd;
  try {
    isd = json::parse(stateJsonText(isdText, len));  // stateAsJson, as PF / FR / SAR read the ISD
  } catch (const std::exception &e) {
    bad(std::string("not JSON: ") + e.what());
  }
  if (isd.type != json::Type::Object) bad("not a JSON object");
  if (!imageIdentifier.empty() || !isd.members.count("image_identifier")) {
    // UsgsAstroPlugin::loadImageSupportData adds the image's base name (src/UsgsAstroPlugin.cpp:341)
    json::Value id;
    id.type = json::Type::String;
    id.str = imageIdentifier;
    isd.members["image_identifier"] = id;
  }
  const std::string name = modelName.empty() ? text(member(isd, "name_model", "sensor model name"), "sensor model name") : modelName;
  State st;
  if (name == "USGS_ASTRO_LINE_SCANNER_SENSOR_MODEL") st = scannerState(isd, false);
  else if (name == "USGS_ASTRO_PUSH_FRAME_SENSOR_MODEL") st = scannerState(isd, true);
  else if (name == "USGS_ASTRO_FRAME_SENSOR_MODEL") st = frameState(isd);
  else if (name == "USGS_ASTRO_SAR_SENSOR_MODEL") st = sarState(isd);
  else bad("unsupported model name: " + name);
  return st.v.members["m_modelName"].str + "\n" + json::dump(st.v);
}

}  // namespace isd
}  // namespace ub200
