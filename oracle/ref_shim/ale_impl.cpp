// ALE for the oracle build (TEST INFRASTRUCTURE -- not part of the product).
//
// The reference's constructStateFromIsd routines (e.g. /root/reference/src/
// UsgsAstroLsSensorModel.cpp:2818-3152) call into DOI-USGS/ale, pinned at commit
// 99c4af43e09c1b953592b5581ecf8b4cbc1e55eb.  The `ale/` submodule is EMPTY in the reference
// checkout and there is no network, so the subset those routines use is restated here from
// ALE's published algorithms (src/Util.cpp, States.cpp, Orientations.cpp, Rotation.cpp,
// InterpUtils.cpp of that project; Eigen's quaternion formulas where ALE delegates to Eigen):
//   * ISD accessors over nlohmann::json;
//   * States::getState(t, SPLINE): cubic Hermite on (position, velocity) in the bracketing
//     interval, linear when the table holds no velocities, constant for a single node;
//   * Orientations: SLERP between bracketing nodes (single node: constant, advanced by its
//     angular velocity), constant rotation applied on the left, inverse(), operator* on the
//     merged time grid, rotateStateAt (velocity picks up -R * (av x p));
//   * Rotation: Hamilton product, matrix <-> quaternion as Eigen does them.
// "Parity partly pinned": this file is NOT ALE's source.  It is pinned by the values the
// reference's own tests assert for models built from the ISD fixtures through this code
// (tests/test_oracle_isd_golden.py replays tests/LineScanCameraTests.cpp,
// FrameCameraTests.cpp, PushFrameCameraTests.cpp, SarTests.cpp).
#include <algorithm>
#include <cmath>
#include <iterator>
#include <stdexcept>
#include <string>

#include "ale/Util.h"

namespace ale {

// ---- InterpUtils ------------------------------------------------------------------------------
static int interpolationIndex(const std::vector<double> &times, double t) {
  if (times.size() < 2) throw std::invalid_argument("There must be at least two times.");
  auto next = std::upper_bound(times.begin(), times.end(), t);
  if (next == times.end()) --next;
  if (next != times.begin()) --next;
  return (int)std::distance(times.begin(), next);
}

static double linearInterpolate(double a, double b, double t) { return a + t * (b - a); }

// Cubic Hermite through (x0, y0, m0), (x1, y1, m1) and its first derivative.
static double cubicHermite(double t, double x0, double x1, double y0, double y1, double m0,
                           double m1) {
  const double h = x1 - x0;
  const double s = (t - x0) / h;
  return (2 * s * s * s - 3 * s * s + 1) * y0 + (s * s * s - 2 * s * s + s) * h * m0 +
         (-2 * s * s * s + 3 * s * s) * y1 + (s * s * s - s * s) * h * m1;
}
static double cubicHermiteDeriv(double t, double x0, double x1, double y0, double y1, double m0,
                                double m1) {
  const double h = x1 - x0;
  if (h == 0.0) throw std::invalid_argument("cubic hermite: equal node times");
  const double s = (t - x0) / h;
  return ((6 * s * s - 6 * s) * y0 + (3 * s * s - 4 * s + 1) * h * m0 +
          (-6 * s * s + 6 * s) * y1 + (3 * s * s - 2 * s) * h * m1) / h;
}

// ---- Rotation -----------------------------------------------------------------------------------
Rotation::Rotation() : m_w(1), m_x(0), m_y(0), m_z(0) {}
Rotation::Rotation(double w, double x, double y, double z) {
  const double n = std::sqrt(w * w + x * x + y * y + z * z);
  m_w = w / n;
  m_x = x / n;
  m_y = y / n;
  m_z = z / n;
}
// 3x3 row-major matrix -> quaternion (Eigen's QuaternionBase::operator=(MatrixBase)).
Rotation::Rotation(const std::vector<double> &m) {
  if (m.size() != 9) throw std::invalid_argument("Rotation matrix must be 3 by 3.");
  auto M = [&](int r, int c) { return m[3 * r + c]; };
  double q[4];  // x y z w
  double t = M(0, 0) + M(1, 1) + M(2, 2);
  if (t > 0.0) {
    t = std::sqrt(t + 1.0);
    q[3] = 0.5 * t;
    t = 0.5 / t;
    q[0] = (M(2, 1) - M(1, 2)) * t;
    q[1] = (M(0, 2) - M(2, 0)) * t;
    q[2] = (M(1, 0) - M(0, 1)) * t;
  } else {
    int i = 0;
    if (M(1, 1) > M(0, 0)) i = 1;
    if (M(2, 2) > M(i, i)) i = 2;
    int j = (i + 1) % 3, k = (j + 1) % 3;
    t = std::sqrt(M(i, i) - M(j, j) - M(k, k) + 1.0);
    q[i] = 0.5 * t;
    t = 0.5 / t;
    q[3] = (M(k, j) - M(j, k)) * t;
    q[j] = (M(j, i) + M(i, j)) * t;
    q[k] = (M(k, i) + M(i, k)) * t;
  }
  m_w = q[3];
  m_x = q[0];
  m_y = q[1];
  m_z = q[2];
}
std::vector<double> Rotation::toQuaternion() const { return {m_w, m_x, m_y, m_z}; }
std::vector<double> Rotation::toRotationMatrix() const {
  const double tx = 2 * m_x, ty = 2 * m_y, tz = 2 * m_z;
  const double twx = tx * m_w, twy = ty * m_w, twz = tz * m_w;
  const double txx = tx * m_x, txy = ty * m_x, txz = tz * m_x;
  const double tyy = ty * m_y, tyz = tz * m_y, tzz = tz * m_z;
  return {1 - (tyy + tzz), txy - twz, txz + twy, txy + twz, 1 - (txx + tzz),
          tyz - twx,       txz - twy, tyz + twx, 1 - (txx + tyy)};
}
Rotation Rotation::inverse() const {
  Rotation r;
  r.m_w = m_w;
  r.m_x = -m_x;
  r.m_y = -m_y;
  r.m_z = -m_z;
  return r;
}
Rotation Rotation::operator*(const Rotation &b) const {
  Rotation r;
  r.m_w = m_w * b.m_w - m_x * b.m_x - m_y * b.m_y - m_z * b.m_z;
  r.m_x = m_w * b.m_x + m_x * b.m_w + m_y * b.m_z - m_z * b.m_y;
  r.m_y = m_w * b.m_y + m_y * b.m_w + m_z * b.m_x - m_x * b.m_z;
  r.m_z = m_w * b.m_z + m_z * b.m_w + m_x * b.m_y - m_y * b.m_x;
  return r;
}
Vec3d Rotation::operator()(const Vec3d &v) const {
  const std::vector<double> R = toRotationMatrix();
  return Vec3d(R[0] * v.x + R[1] * v.y + R[2] * v.z, R[3] * v.x + R[4] * v.y + R[5] * v.z,
               R[6] * v.x + R[7] * v.y + R[8] * v.z);
}
Rotation Rotation::interpolate(const Rotation &next, double t, RotationInterpolation interp) const {
  const double d = m_w * next.m_w + m_x * next.m_x + m_y * next.m_y + m_z * next.m_z;
  double s0, s1;
  if (interp == NLERP) {
    s0 = 1.0 - t;
    s1 = t;
  } else {  // Eigen's slerp
    const double one = 1.0 - 2.220446049250313e-16;
    const double absD = std::fabs(d);
    if (absD >= one) {
      s0 = 1.0 - t;
      s1 = t;
    } else {
      const double theta = std::acos(absD);
      const double sinTheta = std::sin(theta);
      s0 = std::sin((1.0 - t) * theta) / sinTheta;
      s1 = std::sin(t * theta) / sinTheta;
    }
    if (d < 0.0) s1 = -s1;
  }
  Rotation r;
  r.m_w = s0 * m_w + s1 * next.m_w;
  r.m_x = s0 * m_x + s1 * next.m_x;
  r.m_y = s0 * m_y + s1 * next.m_y;
  r.m_z = s0 * m_z + s1 * next.m_z;
  if (interp == NLERP) return Rotation(r.m_w, r.m_x, r.m_y, r.m_z);
  return r;
}

// Rotation by `angle` about `axis` (Eigen::AngleAxisd; a zero axis gives the identity).
static Rotation axisAngle(const Vec3d &axis, double angle) {
  const double n = axis.norm();
  const double s = std::sin(0.5 * angle), c = std::cos(0.5 * angle);
  if (!(n > 0.0)) return Rotation(c, 0.0, 0.0, 0.0);
  return Rotation(c, s * axis.x / n, s * axis.y / n, s * axis.z / n);
}

// ---- States -------------------------------------------------------------------------------------
States::States() : m_refFrame(0) {}
States::States(const std::vector<double> &times, const std::vector<Vec3d> &positions, int refFrame)
    : m_ephemTimes(times), m_refFrame(refFrame) {
  if (positions.size() != times.size())
    throw std::invalid_argument("Length of times must match number of positions");
  for (const Vec3d &p : positions) m_states.push_back(State(p));
}
States::States(const std::vector<double> &times, const std::vector<Vec3d> &positions,
               const std::vector<Vec3d> &velocities, int refFrame)
    : m_ephemTimes(times), m_refFrame(refFrame) {
  if (positions.size() != times.size() || positions.size() != velocities.size())
    throw std::invalid_argument("Length of times must match number of positions and velocities");
  for (size_t i = 0; i < positions.size(); i++) m_states.push_back(State(positions[i], velocities[i]));
}
std::vector<State> States::getStates() const { return m_states; }
std::vector<double> States::getTimes() const { return m_ephemTimes; }
int States::getReferenceFrame() const { return m_refFrame; }
bool States::hasVelocity() const {
  for (const State &s : m_states)
    if (!s.hasVelocity()) return false;
  return !m_states.empty();
}
State States::getState(double time, PositionInterpolation interp) const {
  if (m_states.empty()) throw std::invalid_argument("States: no data");
  if (m_states.size() == 1) return m_states.front();
  const int lo = interpolationIndex(m_ephemTimes, time), hi = lo + 1;
  const State &a = m_states[lo], &b = m_states[hi];
  const double t0 = m_ephemTimes[lo], t1 = m_ephemTimes[hi];
  if (interp == LINEAR || !hasVelocity()) {
    const double s = (time - t0) / (t1 - t0);
    Vec3d p(linearInterpolate(a.position.x, b.position.x, s),
            linearInterpolate(a.position.y, b.position.y, s),
            linearInterpolate(a.position.z, b.position.z, s));
    if (!hasVelocity()) return State(p);
    return State(p, Vec3d(linearInterpolate(a.velocity.x, b.velocity.x, s),
                          linearInterpolate(a.velocity.y, b.velocity.y, s),
                          linearInterpolate(a.velocity.z, b.velocity.z, s)));
  }
  // SPLINE (and LAGRANGE, which the reference never asks for): cubic Hermite
  Vec3d p(cubicHermite(time, t0, t1, a.position.x, b.position.x, a.velocity.x, b.velocity.x),
          cubicHermite(time, t0, t1, a.position.y, b.position.y, a.velocity.y, b.velocity.y),
          cubicHermite(time, t0, t1, a.position.z, b.position.z, a.velocity.z, b.velocity.z));
  Vec3d v(cubicHermiteDeriv(time, t0, t1, a.position.x, b.position.x, a.velocity.x, b.velocity.x),
          cubicHermiteDeriv(time, t0, t1, a.position.y, b.position.y, a.velocity.y, b.velocity.y),
          cubicHermiteDeriv(time, t0, t1, a.position.z, b.position.z, a.velocity.z, b.velocity.z));
  return State(p, v);
}

// ---- Orientations ---------------------------------------------------------------------------------
Orientations::Orientations(const std::vector<Rotation> &rotations, const std::vector<double> &times,
                           const std::vector<Vec3d> &avs, const Rotation &constRot,
                           const std::vector<int> const_frames,
                           const std::vector<int> time_dependent_frames)
    : m_rotations(rotations), m_avs(avs), m_times(times), m_timeDepFrames(time_dependent_frames),
      m_constFrames(const_frames), m_constRotation(constRot) {
  if (m_rotations.size() < 1 || m_times.size() < 1)
    throw std::invalid_argument("There must be at least one rotation and time.");
  if (m_rotations.size() != m_times.size())
    throw std::invalid_argument("The number of rotations and times must be the same.");
  if (!m_avs.empty() && m_avs.size() != m_times.size())
    throw std::invalid_argument("The number of angular velocities and times must be the same.");
}
std::vector<Rotation> Orientations::getRotations() const { return m_rotations; }
std::vector<Vec3d> Orientations::getAngularVelocities() const { return m_avs; }
std::vector<double> Orientations::getTimes() const { return m_times; }
std::vector<int> Orientations::getConstantFrames() const { return m_constFrames; }
std::vector<int> Orientations::getTimeDependentFrames() const { return m_timeDepFrames; }
Rotation Orientations::getConstantRotation() const { return m_constRotation; }

Rotation Orientations::interpolateTimeDep(double time, RotationInterpolation interpType) const {
  if (m_times.size() > 1) {
    const int i = interpolationIndex(m_times, time);
    const double t = (time - m_times[i]) / (m_times[i + 1] - m_times[i]);
    return m_rotations[i].interpolate(m_rotations[i + 1], t, interpType);
  }
  if (m_avs.empty()) return m_rotations.front();
  const double dt = time - m_times.front();
  return axisAngle(m_avs.front(), dt * m_avs.front().norm()) * m_rotations.front();
}
Rotation Orientations::interpolate(double time, RotationInterpolation interpType) const {
  return m_constRotation * interpolateTimeDep(time, interpType);
}
static Vec3d interpolateAv(const std::vector<double> &times, const std::vector<Vec3d> &avs,
                           double time) {
  if (avs.empty()) return Vec3d(0.0, 0.0, 0.0);
  if (times.size() < 2) return avs.front();
  const int i = interpolationIndex(times, time);
  const double t = (time - times[i]) / (times[i + 1] - times[i]);
  return Vec3d(linearInterpolate(avs[i].x, avs[i + 1].x, t), linearInterpolate(avs[i].y, avs[i + 1].y, t),
               linearInterpolate(avs[i].z, avs[i + 1].z, t));
}
Vec3d Orientations::rotateVectorAt(double time, const Vec3d &vector, RotationInterpolation interpType,
                                   bool invert) const {
  Rotation r = interpolate(time, interpType);
  return (invert ? r.inverse() : r)(vector);
}
State Orientations::rotateStateAt(double time, const State &state, RotationInterpolation interpType,
                                  bool invert) const {
  Rotation r = interpolate(time, interpType);
  Vec3d av = interpolateAv(m_times, m_avs, time);
  if (invert) {
    const Vec3d neg = r(av);
    av = Vec3d(-neg.x, -neg.y, -neg.z);
    r = r.inverse();
  }
  // position' = R p ; velocity' = R v - R (av x p): the sign of the angular-velocity term is the
  // one the reference's own golden value pins (tests/SarTests.cpp:133-138 spacecraftVelocity
  // matches to 1e-12 relative with this sign and is off by 4e-3 with the other)
  const Vec3d &p = state.position, &v = state.velocity;
  const Vec3d cross(av.y * p.z - av.z * p.y, av.z * p.x - av.x * p.z, av.x * p.y - av.y * p.x);
  const Vec3d rp = r(p), rv = r(v), rc = r(cross);
  return State(rp, Vec3d(rv.x - rc.x, rv.y - rc.y, rv.z - rc.z));
}
Orientations &Orientations::addConstantRotation(const Rotation &addedConst) {
  m_constRotation = addedConst * m_constRotation;
  return *this;
}
Orientations Orientations::inverse() const {
  // (C R_i)^-1 per node; the constant part is folded in, so the result has none
  std::vector<Rotation> rot;
  std::vector<Vec3d> avs;
  for (size_t i = 0; i < m_rotations.size(); i++) {
    const Rotation full = m_constRotation * m_rotations[i];
    rot.push_back(full.inverse());
    if (!m_avs.empty()) {
      const Vec3d a = full(m_avs[i]);
      avs.push_back(Vec3d(-a.x, -a.y, -a.z));
    }
  }
  return Orientations(rot, m_times, avs, Rotation(1, 0, 0, 0), m_timeDepFrames, m_constFrames);
}
Orientations &Orientations::operator*=(const Rotation &rhs) {
  for (Rotation &r : m_rotations) r = r * rhs;
  return *this;
}
Orientations &Orientations::operator*=(const Orientations &rhs) {
  std::vector<double> merged;
  std::set_union(m_times.begin(), m_times.end(), rhs.m_times.begin(), rhs.m_times.end(),
                 std::back_inserter(merged));
  std::vector<Rotation> rot;
  std::vector<Vec3d> avs;
  for (double t : merged) {
    const Rotation r = rhs.interpolate(t, SLERP);
    rot.push_back(interpolateTimeDep(t, SLERP) * r);
    Vec3d a = r.inverse()(interpolateAv(m_times, m_avs, t));
    const Vec3d b = interpolateAv(rhs.m_times, rhs.m_avs, t);
    avs.push_back(Vec3d(a.x + b.x, a.y + b.y, a.z + b.z));
  }
  m_times = merged;
  m_rotations = rot;
  m_avs = avs;
  return *this;
}
Orientations operator*(Orientations lhs, const Rotation &rhs) { return lhs *= rhs; }
Orientations operator*(const Rotation &lhs, Orientations rhs) { return rhs.addConstantRotation(lhs); }
Orientations operator*(Orientations lhs, const Orientations &rhs) { return lhs *= rhs; }

// ---- ISD accessors (ale/Util.h) ----------------------------------------------------------------------
template <class T>
static T field(const json &isd, const char *key, const char *what) {
  try {
    return isd.at(key).get<T>();
  } catch (...) {
    throw std::runtime_error(std::string("Could not parse the ") + what + ".");
  }
}
template <class T>
static T field2(const json &isd, const char *k1, const char *k2, const char *what) {
  try {
    return isd.at(k1).at(k2).get<T>();
  } catch (...) {
    throw std::runtime_error(std::string("Could not parse the ") + what + ".");
  }
}
std::string getSensorModelName(json isd) { return field<std::string>(isd, "name_model", "sensor model name"); }
std::string getImageId(json isd) { return field<std::string>(isd, "image_identifier", "image identifier"); }
std::string getSensorName(json isd) { return field<std::string>(isd, "name_sensor", "sensor name"); }
std::string getPlatformName(json isd) { return field<std::string>(isd, "name_platform", "platform name"); }
std::string getIsisCameraVersion(json isd) {
  return field<std::string>(isd, "IsisCameraVersion", "isis camera version");
}
int getTotalLines(json isd) { return field<int>(isd, "image_lines", "number of lines in the image"); }
int getTotalSamples(json isd) { return field<int>(isd, "image_samples", "number of samples in the image"); }
double getStartingTime(json isd) { return field<double>(isd, "starting_ephemeris_time", "image start time"); }
double getCenterTime(json isd) { return field<double>(isd, "center_ephemeris_time", "center image time"); }
std::vector<std::vector<double>> getLineScanRate(json isd) {
  return field<std::vector<std::vector<double>>>(isd, "line_scan_rate", "integration start lines in the integration time table");
}
int getSampleSumming(json isd) { return field<int>(isd, "detector_sample_summing", "sample direction detector pixel summing"); }
int getLineSumming(json isd) { return field<int>(isd, "detector_line_summing", "line direction detector pixel summing"); }
double getFocalLength(json isd) { return field2<double>(isd, "focal_length_model", "focal_length", "focal length"); }
double getFocalLengthUncertainty(json isd) {
  try {
    return isd.at("focal_length_model").value("focal_uncertainty", 0.0);
  } catch (...) {
    throw std::runtime_error("Could not parse the focal length uncertainty.");
  }
}
std::vector<double> getFocal2PixelLines(json isd) {
  return field<std::vector<double>>(isd, "focal2pixel_lines", "focal plane coordinate to detector lines transformation");
}
std::vector<double> getFocal2PixelSamples(json isd) {
  return field<std::vector<double>>(isd, "focal2pixel_samples", "focal plane coordinate to detector samples transformation");
}
double getDetectorCenterLine(json isd) { return field2<double>(isd, "detector_center", "line", "detector center line"); }
double getDetectorCenterSample(json isd) { return field2<double>(isd, "detector_center", "sample", "detector center sample"); }
double getDetectorStartingLine(json isd) { return field<double>(isd, "starting_detector_line", "detector starting line"); }
double getDetectorStartingSample(json isd) { return field<double>(isd, "starting_detector_sample", "detector starting sample"); }
double getMinHeight(json isd) { return field2<double>(isd, "reference_height", "minheight", "minimum height above the reference ellipsoid"); }
double getMaxHeight(json isd) { return field2<double>(isd, "reference_height", "maxheight", "maximum height above the reference ellipsoid"); }
double getSemiMajorRadius(json isd) { return field2<double>(isd, "radii", "semimajor", "reference ellipsoid semimajor radius"); }
double getSemiMinorRadius(json isd) { return field2<double>(isd, "radii", "semiminor", "reference ellipsoid semiminor radius"); }

DistortionType getDistortionModel(json isd) {
  try {
    json sub = isd.at("optical_distortion");
    const std::string name = sub.begin().key();
    if (name == "transverse") return TRANSVERSE;
    if (name == "radial") return RADIAL;
    if (name == "kaguyalism") return KAGUYALISM;
    if (name == "dawnfc") return DAWNFC;
    if (name == "lrolrocnac") return LROLROCNAC;
    if (name == "cahvor") return CAHVOR;
    if (name == "lunarorbiter") return LUNARORBITER;
    if (name == "radtan") return RADTAN;
    if (name == "kploshadowcam") return KPLOSHADOWCAM;
  } catch (...) {
    throw std::runtime_error("Could not parse the distortion model.");
  }
  return TRANSVERSE;
}

std::vector<double> getDistortionCoeffs(json isd) {
  const DistortionType type = getDistortionModel(isd);
  const json &od = isd.at("optical_distortion");
  try {
    switch (type) {
      case TRANSVERSE: {
        std::vector<double> x = od.at("transverse").at("x").get<std::vector<double>>();
        std::vector<double> y = od.at("transverse").at("y").get<std::vector<double>>();
        x.resize(10, 0.0);
        y.resize(10, 0.0);
        x.insert(x.end(), y.begin(), y.end());
        return x;
      }
      case RADIAL: return od.at("radial").at("coefficients").get<std::vector<double>>();
      case KAGUYALISM: {
        std::vector<double> x = od.at("kaguyalism").at("x").get<std::vector<double>>();
        std::vector<double> y = od.at("kaguyalism").at("y").get<std::vector<double>>();
        x.resize(4, 0.0);
        y.resize(4, 0.0);
        x.insert(x.begin(), od.at("kaguyalism").at("boresight_x").get<double>());
        y.insert(y.begin(), od.at("kaguyalism").at("boresight_y").get<double>());
        x.insert(x.end(), y.begin(), y.end());
        return x;
      }
      case DAWNFC: return od.at("dawnfc").at("coefficients").get<std::vector<double>>();
      case LROLROCNAC: return od.at("lrolrocnac").at("coefficients").get<std::vector<double>>();
      case CAHVOR: return od.at("cahvor").at("coefficients").get<std::vector<double>>();
      case RADTAN: return od.at("radtan").at("coefficients").get<std::vector<double>>();
      case KPLOSHADOWCAM: return od.at("kploshadowcam").at("coefficients").get<std::vector<double>>();
      case LUNARORBITER: {
        const json &lo = od.at("lunarorbiter");
        return {lo.at("perspective_x").get<double>(), lo.at("perspective_y").get<double>(),
                lo.at("center_point_x").get<double>(), lo.at("center_point_y").get<double>()};
      }
    }
  } catch (...) {
    throw std::runtime_error("Could not parse the distortion model coefficients.");
  }
  return {};
}

static std::vector<Vec3d> vec3Array(const json &a) {
  std::vector<Vec3d> out;
  for (const json &v : a) out.push_back(Vec3d(v.at(0).get<double>(), v.at(1).get<double>(), v.at(2).get<double>()));
  return out;
}
static std::vector<Rotation> quatArray(const json &a) {
  std::vector<Rotation> out;  // ISD quaternions are scalar-first (w, x, y, z)
  for (const json &q : a)
    out.push_back(Rotation(q.at(0).get<double>(), q.at(1).get<double>(), q.at(2).get<double>(),
                           q.at(3).get<double>()));
  return out;
}
std::vector<double> getSunPositions(json isd) {
  std::vector<double> out;
  try {
    for (const json &p : isd.at("sun_position").at("positions"))
      for (int k = 0; k < 3; k++) out.push_back(p.at(k).get<double>());
  } catch (...) {
    throw std::runtime_error("Could not parse the sun positions.");
  }
  return out;
}
static States statesOf(const json &isd, const char *key, const char *what) {
  try {
    const json &s = isd.at(key);
    std::vector<Vec3d> p = vec3Array(s.at("positions"));
    std::vector<double> t = s.at("ephemeris_times").get<std::vector<double>>();
    const int frame = s.at("reference_frame").get<int>();
    if (s.find("velocities") != s.end()) return States(t, p, vec3Array(s.at("velocities")), frame);
    return States(t, p, frame);
  } catch (...) {
    throw std::runtime_error(std::string("Could not parse the ") + what + ".");
  }
}
States getInstrumentPosition(json isd) { return statesOf(isd, "instrument_position", "instrument position"); }
States getSunPosition(json isd) { return statesOf(isd, "sun_position", "sun position"); }
static Orientations orientationsOf(const json &isd, const char *key, const char *what) {
  try {
    const json &o = isd.at(key);
    std::vector<Rotation> rot = quatArray(o.at("quaternions"));
    std::vector<double> t = o.at("ephemeris_times").get<std::vector<double>>();
    std::vector<Vec3d> av;
    if (o.find("angular_velocities") != o.end()) av = vec3Array(o.at("angular_velocities"));
    std::vector<int> cf, tf;
    if (o.find("constant_frames") != o.end()) cf = o.at("constant_frames").get<std::vector<int>>();
    if (o.find("time_dependent_frames") != o.end()) tf = o.at("time_dependent_frames").get<std::vector<int>>();
    std::vector<double> c = {1, 0, 0, 0, 1, 0, 0, 0, 1};
    if (o.find("constant_rotation") != o.end()) c = o.at("constant_rotation").get<std::vector<double>>();
    return Orientations(rot, t, av, Rotation(c), cf, tf);
  } catch (...) {
    throw std::runtime_error(std::string("Could not parse the ") + what + ".");
  }
}
Orientations getBodyRotation(json isd) { return orientationsOf(isd, "body_rotation", "body rotation"); }
Orientations getInstrumentPointing(json isd) { return orientationsOf(isd, "instrument_pointing", "instrument pointing"); }
std::vector<double> getGeoTransform(json isd) {
  return field<std::vector<double>>(isd, "geotransform", "geotransform");
}
std::string getProjection(json isd) { return field<std::string>(isd, "projection", "projection string"); }

}  // namespace ale
