// Robot — footprint parameters (reference include/sspp/robot.h).  Kept as a type of the planner API;
// the collision test never consults it (the swept box is set on the heuristic).
#pragma once
#include <string>
#include "utils.h"
class Robot {
 public:
  double robotLength = 0, robotWidth = 0, narrowestPathDist = 0, safetyTolerance = 0.05, robotRadius = 0;
  std::string robotName;
  geometry_msgs::Point robotCenter;
  Pose robotLocation;
  Robot() {}
  Robot(std::string name, double length, double width, double narrowDist, geometry_msgs::Point center, double tol = 0.05)
      : robotLength(length), robotWidth(width), narrowestPathDist(narrowDist), safetyTolerance(tol), robotName(name),
        robotCenter(center) {
    robotRadius = std::sqrt(length * length + width * width) / 2.0;
  }
};
