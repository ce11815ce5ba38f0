// isd_math.h -- the ephemeris / attitude resampling an ISD needs on its way to a model state
// (host only).
//
// The reference delegates this to the ALE library (DOI-USGS/ale @ 99c4af43, an un-vendored
// submodule): States::getState(t, SPLINE), Orientations::rotateStateAt / inverse / operator* /
// interpolate(SLERP), Rotation::toQuaternion, called from the four constructStateFromIsd
// routines (src/UsgsAstroLsSensorModel.cpp:2883-3090, src/UsgsAstroPushFrameSensorModel.cpp:
// 2160-2400, src/UsgsAstroFrameSensorModel.cpp:1687-1806, src/UsgsAstroSarSensorModel.cpp:84-168).
// ALE's source is not available here, so its published algorithms are restated: cubic Hermite
// position interpolation on (position, velocity) nodes, SLERP between attitude nodes (Eigen's
// formula), frame chains as quaternion products on the merged time grid, and the transport
// term -R (w x p) on rotated velocities.  Pinned by the values the reference's own tests
// assert for its ISD fixtures (tests/test_isd_golden_gpu.py, tests/test_isd_state.py); the
// bits of a converted state may differ from ALE's in the last place.
#ifndef USGSCSM_B200_ISD_MATH_H_
#define USGSCSM_B200_ISD_MATH_H_

#include <algorithm>
#include <cmath>
#include <iterator>
#include <stdexcept>
#include <vector>

namespace ub200 {
namespace isd {

struct V3 {
  double x = 0.0, y = 0.0, z = 0.0;
};
inline V3 cross(const V3 &a, const V3 &b) {
  return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
inline double norm(const V3 &a) { return std::sqrt(a.x * a.x + a.y * a.y + a.z * a.z); }

// Unit quaternion, scalar first (the ISD's order).
struct Quat {
  double w = 1.0, x = 0.0, y = 0.0, z = 0.0;
};
inline Quat unit(double w, double x, double y, double z) {
  const double n = std::sqrt(w * w + x * x + y * y + z * z);
  return {w / n, x / n, y / n, z / n};
}
inline Quat conj(const Quat &q) { return {q.w, -q.x, -q.y, -q.z}; }
inline Quat mul(const Quat &a, const Quat &b) {  // Hamilton product: rotate by b, then by a
  return {a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z,
          a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y,
          a.w * b.y + a.y * b.w + a.z * b.x - a.x * b.z,
          a.w * b.z + a.z * b.w + a.x * b.y - a.y * b.x};
}
inline void toMatrix(const Quat &q, double R[9]) {
  const double tx = 2 * q.x, ty = 2 * q.y, tz = 2 * q.z;
  const double twx = tx * q.w, twy = ty * q.w, twz = tz * q.w;
  const double txx = tx * q.x, txy = ty * q.x, txz = tz * q.x;
  const double tyy = ty * q.y, tyz = tz * q.y, tzz = tz * q.z;
  R[0] = 1 - (tyy + tzz); R[1] = txy - twz;       R[2] = txz + twy;
  R[3] = txy + twz;       R[4] = 1 - (txx + tzz); R[5] = tyz - twx;
  R[6] = txz - twy;       R[7] = tyz + twx;       R[8] = 1 - (txx + tyy);
}
inline V3 rotate(const Quat &q, const V3 &v) {
  double R[9];
  toMatrix(q, R);
  return {R[0] * v.x + R[1] * v.y + R[2] * v.z, R[3] * v.x + R[4] * v.y + R[5] * v.z,
          R[6] * v.x + R[7] * v.y + R[8] * v.z};
}
// Row-major 3x3 rotation matrix -> quaternion (largest-pivot form, as Eigen converts).
inline Quat fromMatrix(const std::vector<double> &m) {
  if (m.size() != 9) throw std::runtime_error("constant_rotation must hold 9 values");
  const double tr = m[0] + m[4] + m[8];
  double q[4];  // x y z w
  if (tr > 0.0) {
    double t = std::sqrt(tr + 1.0);
    q[3] = 0.5 * t;
    t = 0.5 / t;
    q[0] = (m[7] - m[5]) * t;
    q[1] = (m[2] - m[6]) * t;
    q[2] = (m[3] - m[1]) * t;
  } else {
    int i = 0;
    if (m[4] > m[0]) i = 1;
    if (m[8] > m[4 * i]) i = 2;
    const int j = (i + 1) % 3, k = (j + 1) % 3;
    double t = std::sqrt(m[4 * i] - m[4 * j] - m[4 * k] + 1.0);
    q[i] = 0.5 * t;
    t = 0.5 / t;
    q[3] = (m[3 * k + j] - m[3 * j + k]) * t;
    q[j] = (m[3 * j + i] + m[3 * i + j]) * t;
    q[k] = (m[3 * k + i] + m[3 * i + k]) * t;
  }
  return {q[3], q[0], q[1], q[2]};
}
// Spherical linear interpolation a -> b at fraction t (shortest arc; linear when the two are
// within one epsilon of parallel).
inline Quat slerp(const Quat &a, const Quat &b, double t) {
  const double d = a.w * b.w + a.x * b.x + a.y * b.y + a.z * b.z;
  const double ad = std::fabs(d);
  double s0 = 1.0 - t, s1 = t;
  if (ad < 1.0 - 2.220446049250313e-16) {
    const double theta = std::acos(ad), st = std::sin(theta);
    s0 = std::sin((1.0 - t) * theta) / st;
    s1 = std::sin(t * theta) / st;
  }
  if (d < 0.0) s1 = -s1;
  return {s0 * a.w + s1 * b.w, s0 * a.x + s1 * b.x, s0 * a.y + s1 * b.y, s0 * a.z + s1 * b.z};
}

// Index of the node interval [i, i+1] used for time t: the one containing t, the first / last
// interval outside the table.  Needs >= 2 nodes.
inline int bracket(const std::vector<double> &times, double t) {
  if (times.size() < 2) throw std::runtime_error("interpolation needs at least two times");
  size_t hi = std::upper_bound(times.begin(), times.end(), t) - times.begin();
  if (hi == times.size()) --hi;
  if (hi > 0) --hi;
  return (int)hi;
}

// ---- position / velocity table ("instrument_position", "sun_position") ---------------------------
struct Ephemeris {
  std::vector<double> times;
  std::vector<V3> pos, vel;  // vel empty when the ISD has none

  // ALE States::getState(t, SPLINE)
  void at(double t, V3 &p, V3 &v) const {
    if (pos.empty()) throw std::runtime_error("empty position table");
    const bool hasVel = vel.size() == pos.size();
    const double nan = std::nan("");
    if (pos.size() == 1) {
      p = pos[0];
      v = hasVel ? vel[0] : V3{nan, nan, nan};
      return;
    }
    const int i = bracket(times, t);
    const double h = times[i + 1] - times[i], s = (t - times[i]) / h;
    if (!hasVel) {
      p = {pos[i].x + s * (pos[i + 1].x - pos[i].x), pos[i].y + s * (pos[i + 1].y - pos[i].y),
           pos[i].z + s * (pos[i + 1].z - pos[i].z)};
      v = {nan, nan, nan};
      return;
    }
    // cubic Hermite basis and its derivative
    const double s2 = s * s, s3 = s2 * s;
    const double h00 = 2 * s3 - 3 * s2 + 1, h10 = s3 - 2 * s2 + s, h01 = -2 * s3 + 3 * s2,
                 h11 = s3 - s2;
    const double d00 = 6 * s2 - 6 * s, d10 = 3 * s2 - 4 * s + 1, d01 = -6 * s2 + 6 * s,
                 d11 = 3 * s2 - 2 * s;
    auto val = [&](double y0, double y1, double m0, double m1) {
      return h00 * y0 + h10 * h * m0 + h01 * y1 + h11 * h * m1;
    };
    auto der = [&](double y0, double y1, double m0, double m1) {
      return (d00 * y0 + d10 * h * m0 + d01 * y1 + d11 * h * m1) / h;
    };
    p = {val(pos[i].x, pos[i + 1].x, vel[i].x, vel[i + 1].x),
         val(pos[i].y, pos[i + 1].y, vel[i].y, vel[i + 1].y),
         val(pos[i].z, pos[i + 1].z, vel[i].z, vel[i + 1].z)};
    v = {der(pos[i].x, pos[i + 1].x, vel[i].x, vel[i + 1].x),
         der(pos[i].y, pos[i + 1].y, vel[i].y, vel[i + 1].y),
         der(pos[i].z, pos[i + 1].z, vel[i].z, vel[i + 1].z)};
  }
};

// ---- attitude table ("instrument_pointing", "body_rotation") ------------------------------------------
// full(t) = constant * slerp(nodes)(t)
struct Attitude {
  std::vector<double> times;
  std::vector<Quat> rot;
  std::vector<V3> av;  // angular velocities; may be empty
  Quat constant;       // identity unless the ISD has "constant_rotation"

  Quat timeDependent(double t) const {
    if (rot.empty()) throw std::runtime_error("empty attitude table");
    if (rot.size() > 1) {
      const int i = bracket(times, t);
      return slerp(rot[i], rot[i + 1], (t - times[i]) / (times[i + 1] - times[i]));
    }
    if (av.empty()) return rot[0];
    // a single node is carried along by its angular velocity
    const double rate = norm(av[0]), angle = (t - times[0]) * rate;
    const double c = std::cos(0.5 * angle), s = std::sin(0.5 * angle);
    const Quat step = rate > 0.0 ? unit(c, s * av[0].x / rate, s * av[0].y / rate, s * av[0].z / rate)
                                 : unit(c, 0.0, 0.0, 0.0);
    return mul(step, rot[0]);
  }
  Quat full(double t) const { return mul(constant, timeDependent(t)); }
  V3 angularVelocity(double t) const {
    if (av.empty()) return {};
    if (times.size() < 2) return av[0];
    const int i = bracket(times, t);
    const double s = (t - times[i]) / (times[i + 1] - times[i]);
    return {av[i].x + s * (av[i + 1].x - av[i].x), av[i].y + s * (av[i + 1].y - av[i].y),
            av[i].z + s * (av[i + 1].z - av[i].z)};
  }
  // ALE Orientations::rotateStateAt(t, state, SLERP): position R p, velocity R v - R (w x p).
  void rotateState(double t, const V3 &p, const V3 &v, V3 &pOut, V3 &vOut) const {
    const Quat q = full(t);
    const V3 w = angularVelocity(t);
    const V3 rp = rotate(q, p), rv = rotate(q, v), rc = rotate(q, cross(w, p));
    pOut = rp;
    vOut = {rv.x - rc.x, rv.y - rc.y, rv.z - rc.z};
  }
};

// Nodes of (target <- J2000) * (sensor <- J2000)^-1, i.e. sensor -> target, on the union of the
// two time grids (ALE: j2000_to_target * j2000_to_sensor.inverse()).  Returned as an Attitude
// whose constant part is the target chain's.
inline Attitude sensorToTarget(const Attitude &j2000ToTarget, const Attitude &j2000ToSensor) {
  // inverse of the sensor chain, node by node, constant part folded in
  Attitude inv;
  inv.times = j2000ToSensor.times;
  for (size_t i = 0; i < j2000ToSensor.rot.size(); i++) {
    const Quat whole = mul(j2000ToSensor.constant, j2000ToSensor.rot[i]);
    inv.rot.push_back(conj(whole));
    if (!j2000ToSensor.av.empty()) {  // the inverse turns the other way, seen from the far frame
      const V3 w = rotate(whole, j2000ToSensor.av[i]);
      inv.av.push_back({-w.x, -w.y, -w.z});
    }
  }
  Attitude out;
  out.constant = j2000ToTarget.constant;
  std::set_union(j2000ToTarget.times.begin(), j2000ToTarget.times.end(), inv.times.begin(),
                 inv.times.end(), std::back_inserter(out.times));
  for (double t : out.times) out.rot.push_back(mul(j2000ToTarget.timeDependent(t), inv.full(t)));
  return out;
}

}  // namespace isd
}  // namespace ub200
#endif
