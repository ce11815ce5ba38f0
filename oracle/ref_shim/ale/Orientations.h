// ALE shim (see Vectors.h): time-tagged rotation table with a constant rotation.
#ifndef ORACLE_SHIM_ALE_ORIENTATIONS_H
#define ORACLE_SHIM_ALE_ORIENTATIONS_H
#include <vector>
#include "ale/Rotation.h"
#include "ale/States.h"
namespace ale {
class Orientations {
 public:
  Orientations() {}
  Orientations(const std::vector<Rotation> &rotations, const std::vector<double> &times,
               const std::vector<Vec3d> &avs = std::vector<Vec3d>(),
               const Rotation &constRot = Rotation(1, 0, 0, 0),
               const std::vector<int> const_frames = std::vector<int>(),
               const std::vector<int> time_dependent_frames = std::vector<int>());
  std::vector<Rotation> getRotations() const;
  std::vector<Vec3d> getAngularVelocities() const;
  std::vector<double> getTimes() const;
  std::vector<int> getConstantFrames() const;
  std::vector<int> getTimeDependentFrames() const;
  Rotation getConstantRotation() const;
  Rotation interpolate(double time, RotationInterpolation interpType = SLERP) const;
  Rotation interpolateTimeDep(double time, RotationInterpolation interpType = SLERP) const;
  Vec3d rotateVectorAt(double time, const Vec3d &vector, RotationInterpolation interpType = SLERP,
                       bool invert = false) const;
  State rotateStateAt(double time, const State &state, RotationInterpolation interpType = SLERP,
                      bool invert = false) const;
  Orientations &addConstantRotation(const Rotation &addedConst);
  Orientations &operator*=(const Orientations &rhs);
  Orientations &operator*=(const Rotation &rhs);
  Orientations inverse() const;
 private:
  std::vector<Rotation> m_rotations;
  std::vector<Vec3d> m_avs;
  std::vector<double> m_times;
  std::vector<int> m_timeDepFrames;
  std::vector<int> m_constFrames;
  Rotation m_constRotation;
};
Orientations operator*(Orientations lhs, const Rotation &rhs);
Orientations operator*(const Rotation &lhs, Orientations rhs);
Orientations operator*(Orientations lhs, const Orientations &rhs);
}  // namespace ale
#endif
