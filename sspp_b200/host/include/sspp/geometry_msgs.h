// geometry_msgs.h — the handful of ROS message types the planner API is written in, as plain
// structs, so the host mirror builds without ROS.  Inside a catkin build define SSPP_WITH_ROS and
// the real headers are used instead (same member names, same zero initialisation).
#pragma once
#ifdef SSPP_WITH_ROS
#include <geometry_msgs/Pose.h>
#include <geometry_msgs/PoseArray.h>
#include <geometry_msgs/Vector3.h>
#else
#include <vector>
namespace geometry_msgs {
struct Point { double x = 0, y = 0, z = 0; };
struct Vector3 { double x = 0, y = 0, z = 0; };
struct Quaternion { double x = 0, y = 0, z = 0, w = 0; };
struct Pose { Point position; Quaternion orientation; };
struct PoseArray { std::vector<Pose> poses; };
}  // namespace geometry_msgs
#endif
