// utils.h — the comparisons and the distance the hot path depends on, with the reference's exact
// semantics (reference include/sspp/utils.h): Dist :351-359, isEqual :221-239, isPositionEqual /
// isOrientationEqual :241-259, samePosition :386-390, Pose :392-413, Tree :493-497.
#pragma once
#include <cmath>
#include <vector>
#include "geometry_msgs.h"

// sqrt of the sum of squared differences, evaluated left to right in double (build with
// -ffp-contract=off: the A* tie-breaks hinge on the last ulp of g + h)
inline double Dist(const geometry_msgs::Point& a, const geometry_msgs::Point& b) {
  return std::sqrt((a.x - b.x) * (a.x - b.x) + (a.y - b.y) * (a.y - b.y) + (a.z - b.z) * (a.z - b.z));
}
inline double Dist(const geometry_msgs::Pose& a, const geometry_msgs::Pose& b) { return Dist(a.position, b.position); }

inline bool isEqual(double a, double b) { return std::fabs(a - b) < 1E-12; }
inline bool isPositionEqual(const geometry_msgs::Point& a, const geometry_msgs::Point& b) {
  return std::fabs(a.x - b.x) < 1E-12 && std::fabs(a.y - b.y) < 1E-12 && std::fabs(a.z - b.z) < 1E-12;
}
inline bool isOrientationEqual(const geometry_msgs::Quaternion& a, const geometry_msgs::Quaternion& b) {
  return std::fabs(a.x - b.x) < 1E-12 && std::fabs(a.y - b.y) < 1E-12 && std::fabs(a.z - b.z) < 1E-12 &&
         std::fabs(a.w - b.w) < 1E-12;
}
// all seven pose components within 1e-12 (the search space's identity test)
inline bool samePosition(const geometry_msgs::Pose& a, const geometry_msgs::Pose& b) {
  return isPositionEqual(a.position, b.position) && isOrientationEqual(a.orientation, b.orientation);
}

class Pose {
 public:
  geometry_msgs::Pose p;
  double phi = 0;
  Pose() {}
  Pose(double x, double y, double z, double theta) : phi(theta) { p.position.x = x; p.position.y = y; p.position.z = z; }
  bool operator==(const Pose& o) const { return samePosition(p, o.p); }
  bool operator!=(const Pose& o) const { return !samePosition(p, o.p); }
};

// one expansion of the search tree, kept for Heuristic::displayProgress
struct Tree {
  geometry_msgs::Pose location;
  std::vector<geometry_msgs::Pose> children;
};

#define DTOR(d) ((d) * M_PI / 180)
