// ALE shim (see Vectors.h): unit-quaternion rotation, (w,x,y,z) order.
#ifndef ORACLE_SHIM_ALE_ROTATION_H
#define ORACLE_SHIM_ALE_ROTATION_H
#include <vector>
#include "ale/Vectors.h"
namespace ale {
enum RotationInterpolation { SLERP, NLERP };
class Rotation {
 public:
  Rotation();                                        // identity
  Rotation(double w, double x, double y, double z);  // normalised on construction
  Rotation(const std::vector<double> &matrix);       // 3x3 row-major
  std::vector<double> toQuaternion() const;          // w,x,y,z
  std::vector<double> toRotationMatrix() const;      // 3x3 row-major
  Rotation inverse() const;
  Rotation operator*(const Rotation &rightRotation) const;
  Vec3d operator()(const Vec3d &vector) const;
  Rotation interpolate(const Rotation &next, double t, RotationInterpolation interp) const;
 private:
  double m_w, m_x, m_y, m_z;
};
}  // namespace ale
#endif
