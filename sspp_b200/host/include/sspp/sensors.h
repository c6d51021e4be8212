// Sensors — camera intrinsics and mount pose, and the robot-to-sensor pose that defines a
// viewpoint for the visibility pass (reference include/sspp/sensors.h:57-68, src/sensors.cpp:47-92).
#pragma once
#include "geometry_msgs.h"
class Sensors {
 public:
  double hFOV = 0, vFOV = 0, focalLength = 0, near = 0, far = 0, resWidth = 0, resHeight = 0;
  float sensorPose[3] = {0, 0, 0}, sensorRPY[3] = {0, 0, 0};
  Sensors() {}
  Sensors(double hfov, double vfov, double fl, double nearR, double farR, double width, double height, const float sP[3],
          const float sRPY[3]);
  // T_robot * T_mount * Rx(+90deg): the trailing rotation turns the optical frame into the
  // X-forward / Y-up / Z-right camera frame the frustum test expects
  geometry_msgs::Pose robot2sensorTransformation(const geometry_msgs::Pose& pose) const;
};
