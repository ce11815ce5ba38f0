// Declaration shim for the ALE library (DOI-USGS/ale @ 99c4af43, absent from the
// reference checkout: the `ale/` submodule is empty).  Written from scratch for the
// oracle build only; it exists so that the reference's *unmodified* sensor-model
// sources compile here.  TEST INFRASTRUCTURE -- not part of the product.
#ifndef ORACLE_SHIM_ALE_VECTORS_H
#define ORACLE_SHIM_ALE_VECTORS_H
#include <vector>
#include <stdexcept>
#include <cmath>
namespace ale {
struct Vec3d {
  double x, y, z;
  Vec3d(double x_ = 0.0, double y_ = 0.0, double z_ = 0.0) : x(x_), y(y_), z(z_) {}
  Vec3d(const std::vector<double> &v) {
    if (v.size() != 3) throw std::invalid_argument("Vec3d needs 3 values");
    x = v[0]; y = v[1]; z = v[2];
  }
  Vec3d &operator*=(double s) { x *= s; y *= s; z *= s; return *this; }
  Vec3d operator*(double s) const { return Vec3d(x * s, y * s, z * s); }
  Vec3d &operator+=(const Vec3d &o) { x += o.x; y += o.y; z += o.z; return *this; }
  Vec3d operator+(const Vec3d &o) const { return Vec3d(x + o.x, y + o.y, z + o.z); }
  Vec3d &operator-=(const Vec3d &o) { x -= o.x; y -= o.y; z -= o.z; return *this; }
  Vec3d operator-(const Vec3d &o) const { return Vec3d(x - o.x, y - o.y, z - o.z); }
  double norm() const { return std::sqrt(x * x + y * y + z * z); }
};
}  // namespace ale
#endif
