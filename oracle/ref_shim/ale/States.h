// ALE shim (see Vectors.h): time-tagged position/velocity table.
#ifndef ORACLE_SHIM_ALE_STATES_H
#define ORACLE_SHIM_ALE_STATES_H
#include <vector>
#include <limits>
#include "ale/Vectors.h"
namespace ale {
enum PositionInterpolation { LINEAR, SPLINE, LAGRANGE };
struct State {
  Vec3d position;
  Vec3d velocity;
  State() : position(0, 0, 0),
            velocity(std::numeric_limits<double>::quiet_NaN(),
                     std::numeric_limits<double>::quiet_NaN(),
                     std::numeric_limits<double>::quiet_NaN()) {}
  State(Vec3d p) : position(p),
            velocity(std::numeric_limits<double>::quiet_NaN(),
                     std::numeric_limits<double>::quiet_NaN(),
                     std::numeric_limits<double>::quiet_NaN()) {}
  State(Vec3d p, Vec3d v) : position(p), velocity(v) {}
  bool hasVelocity() const { return !(std::isnan(velocity.x) || std::isnan(velocity.y) || std::isnan(velocity.z)); }
};
class States {
 public:
  States();
  States(const std::vector<double> &ephemTimes, const std::vector<Vec3d> &positions,
         int refFrame = 1);
  States(const std::vector<double> &ephemTimes, const std::vector<Vec3d> &positions,
         const std::vector<Vec3d> &velocities, int refFrame = 1);
  std::vector<State> getStates() const;
  std::vector<double> getTimes() const;
  int getReferenceFrame() const;
  bool hasVelocity() const;
  State getState(double time, PositionInterpolation interp = LINEAR) const;
 private:
  std::vector<State> m_states;
  std::vector<double> m_ephemTimes;
  int m_refFrame;
};
}  // namespace ale
#endif
