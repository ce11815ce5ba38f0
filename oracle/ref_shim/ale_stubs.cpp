// ALE link shim for the oracle build (TEST INFRASTRUCTURE -- not part of the product).
//
// The reference's sensor-model sources call into DOI-USGS/ale (pinned commit
// 99c4af43e09c1b953592b5581ecf8b4cbc1e55eb) from their constructStateFromIsd routines
// only (e.g. /root/reference/src/UsgsAstroLsSensorModel.cpp:2827-3123).  The `ale/`
// submodule is empty in the reference checkout, so these entry points are provided
// here.  The oracle is always fed *model states* (never ISDs), so every ISD accessor
// simply throws: reaching one means a test tried to use the ISD path.
#include <stdexcept>
#include <string>

#include "ale/Util.h"

namespace ale {

static void absent(const char *what) {
  throw std::runtime_error(std::string("oracle ALE shim: ISD path not available (") + what + ")");
}

#define ALE_ABSENT(ret, name)          \
  ret name(json) {                     \
    absent(#name);                     \
    return ret();                      \
  }

ALE_ABSENT(std::string, getSensorModelName)
ALE_ABSENT(std::string, getImageId)
ALE_ABSENT(std::string, getSensorName)
ALE_ABSENT(std::string, getPlatformName)
ALE_ABSENT(std::string, getIsisCameraVersion)
ALE_ABSENT(int, getTotalLines)
ALE_ABSENT(int, getTotalSamples)
ALE_ABSENT(double, getStartingTime)
ALE_ABSENT(double, getCenterTime)
ALE_ABSENT(std::vector<std::vector<double>>, getLineScanRate)
ALE_ABSENT(int, getSampleSumming)
ALE_ABSENT(int, getLineSumming)
ALE_ABSENT(double, getFocalLength)
ALE_ABSENT(double, getFocalLengthUncertainty)
ALE_ABSENT(std::vector<double>, getFocal2PixelLines)
ALE_ABSENT(std::vector<double>, getFocal2PixelSamples)
ALE_ABSENT(double, getDetectorCenterLine)
ALE_ABSENT(double, getDetectorCenterSample)
ALE_ABSENT(double, getDetectorStartingLine)
ALE_ABSENT(double, getDetectorStartingSample)
ALE_ABSENT(double, getMinHeight)
ALE_ABSENT(double, getMaxHeight)
ALE_ABSENT(double, getSemiMajorRadius)
ALE_ABSENT(double, getSemiMinorRadius)
ALE_ABSENT(std::vector<double>, getDistortionCoeffs)
ALE_ABSENT(std::vector<double>, getSunPositions)
ALE_ABSENT(States, getInstrumentPosition)
ALE_ABSENT(States, getSunPosition)
ALE_ABSENT(Orientations, getBodyRotation)
ALE_ABSENT(Orientations, getInstrumentPointing)
ALE_ABSENT(std::vector<double>, getGeoTransform)
ALE_ABSENT(std::string, getProjection)

DistortionType getDistortionModel(json) {
  absent("getDistortionModel");
  return RADIAL;
}

// ---- Rotation ---------------------------------------------------------------------
Rotation::Rotation() : m_w(1), m_x(0), m_y(0), m_z(0) {}
Rotation::Rotation(double w, double x, double y, double z) : m_w(w), m_x(x), m_y(y), m_z(z) {}
Rotation::Rotation(const std::vector<double> &) : m_w(1), m_x(0), m_y(0), m_z(0) {
  absent("Rotation(matrix)");
}
std::vector<double> Rotation::toQuaternion() const { return {m_w, m_x, m_y, m_z}; }
std::vector<double> Rotation::toRotationMatrix() const {
  absent("Rotation::toRotationMatrix");
  return {};
}
Rotation Rotation::inverse() const { return Rotation(m_w, -m_x, -m_y, -m_z); }
Rotation Rotation::operator*(const Rotation &) const {
  absent("Rotation::operator*");
  return Rotation();
}
Vec3d Rotation::operator()(const Vec3d &) const {
  absent("Rotation::operator()");
  return Vec3d();
}
Rotation Rotation::interpolate(const Rotation &, double, RotationInterpolation) const {
  absent("Rotation::interpolate");
  return Rotation();
}

// ---- States -----------------------------------------------------------------------
States::States() : m_refFrame(1) {}
States::States(const std::vector<double> &t, const std::vector<Vec3d> &p, int f)
    : m_ephemTimes(t), m_refFrame(f) {
  for (const Vec3d &v : p) m_states.push_back(State(v));
}
States::States(const std::vector<double> &t, const std::vector<Vec3d> &p,
               const std::vector<Vec3d> &v, int f)
    : m_ephemTimes(t), m_refFrame(f) {
  for (size_t i = 0; i < p.size(); i++) m_states.push_back(State(p[i], v[i]));
}
std::vector<State> States::getStates() const { return m_states; }
std::vector<double> States::getTimes() const { return m_ephemTimes; }
int States::getReferenceFrame() const { return m_refFrame; }
bool States::hasVelocity() const { return !m_states.empty() && m_states[0].hasVelocity(); }
State States::getState(double, PositionInterpolation) const {
  absent("States::getState");
  return State();
}

// ---- Orientations -----------------------------------------------------------------
Orientations::Orientations(const std::vector<Rotation> &r, const std::vector<double> &t,
                           const std::vector<Vec3d> &avs, const Rotation &c,
                           const std::vector<int> cf, const std::vector<int> tf)
    : m_rotations(r), m_avs(avs), m_times(t), m_timeDepFrames(tf), m_constFrames(cf),
      m_constRotation(c) {}
std::vector<Rotation> Orientations::getRotations() const { return m_rotations; }
std::vector<Vec3d> Orientations::getAngularVelocities() const { return m_avs; }
std::vector<double> Orientations::getTimes() const { return m_times; }
std::vector<int> Orientations::getConstantFrames() const { return m_constFrames; }
std::vector<int> Orientations::getTimeDependentFrames() const { return m_timeDepFrames; }
Rotation Orientations::getConstantRotation() const { return m_constRotation; }
Rotation Orientations::interpolate(double, RotationInterpolation) const {
  absent("Orientations::interpolate");
  return Rotation();
}
Rotation Orientations::interpolateTimeDep(double, RotationInterpolation) const {
  absent("Orientations::interpolateTimeDep");
  return Rotation();
}
Vec3d Orientations::rotateVectorAt(double, const Vec3d &, RotationInterpolation, bool) const {
  absent("Orientations::rotateVectorAt");
  return Vec3d();
}
State Orientations::rotateStateAt(double, const State &, RotationInterpolation, bool) const {
  absent("Orientations::rotateStateAt");
  return State();
}
Orientations &Orientations::addConstantRotation(const Rotation &) {
  absent("Orientations::addConstantRotation");
  return *this;
}
Orientations &Orientations::operator*=(const Orientations &) {
  absent("Orientations::operator*=");
  return *this;
}
Orientations &Orientations::operator*=(const Rotation &) {
  absent("Orientations::operator*=");
  return *this;
}
Orientations Orientations::inverse() const {
  absent("Orientations::inverse");
  return Orientations();
}
Orientations operator*(Orientations lhs, const Rotation &) {
  absent("operator*(Orientations, Rotation)");
  return lhs;
}
Orientations operator*(const Rotation &, Orientations rhs) {
  absent("operator*(Rotation, Orientations)");
  return rhs;
}
Orientations operator*(Orientations lhs, const Orientations &) {
  absent("operator*(Orientations, Orientations)");
  return lhs;
}

}  // namespace ale
