// Sensors (reference src/sensors.cpp:47-92): sensor pose = T_robot * T_mount * Rx(+90deg), rotations
// from quaternions the way tf does it, all in double.
#include "sspp/sensors.h"

#include <cmath>

namespace {
struct Mat3 { double m[3][3]; };
struct Quat { double x, y, z, w; };

Quat quat_from_rpy(double roll, double pitch, double yaw) {  // tf::Quaternion::setRPY
  double hy = yaw * 0.5, hp = pitch * 0.5, hr = roll * 0.5;
  double cy = std::cos(hy), sy = std::sin(hy), cp = std::cos(hp), sp = std::sin(hp), cr = std::cos(hr), sr = std::sin(hr);
  Quat q;
  q.x = sr * cp * cy - cr * sp * sy;
  q.y = cr * sp * cy + sr * cp * sy;
  q.z = cr * cp * sy - sr * sp * cy;
  q.w = cr * cp * cy + sr * sp * sy;
  return q;
}
Mat3 mat_from_quat(const Quat& q) {  // tf::Matrix3x3::setRotation
  double d = q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w;
  double s = 2.0 / d;
  double xs = q.x * s, ys = q.y * s, zs = q.z * s;
  double wx = q.w * xs, wy = q.w * ys, wz = q.w * zs;
  double xx = q.x * xs, xy = q.x * ys, xz = q.x * zs;
  double yy = q.y * ys, yz = q.y * zs, zz = q.z * zs;
  Mat3 r = {{{1.0 - (yy + zz), xy - wz, xz + wy}, {xy + wz, 1.0 - (xx + zz), yz - wx}, {xz - wy, yz + wx, 1.0 - (xx + yy)}}};
  return r;
}
Quat quat_from_mat(const Mat3& r) {  // tf::Matrix3x3::getRotation
  const double(*m)[3] = r.m;
  double t[4];
  double trace = m[0][0] + m[1][1] + m[2][2];
  if (trace > 0.0) {
    double s = std::sqrt(trace + 1.0);
    t[3] = s * 0.5;
    s = 0.5 / s;
    t[0] = (m[2][1] - m[1][2]) * s;
    t[1] = (m[0][2] - m[2][0]) * s;
    t[2] = (m[1][0] - m[0][1]) * s;
  } else {
    int i = m[0][0] < m[1][1] ? (m[1][1] < m[2][2] ? 2 : 1) : (m[0][0] < m[2][2] ? 2 : 0);
    int j = (i + 1) % 3, k = (i + 2) % 3;
    double s = std::sqrt(m[i][i] - m[j][j] - m[k][k] + 1.0);
    t[i] = s * 0.5;
    s = 0.5 / s;
    t[3] = (m[k][j] - m[j][k]) * s;
    t[j] = (m[j][i] + m[i][j]) * s;
    t[k] = (m[k][i] + m[i][k]) * s;
  }
  Quat q = {t[0], t[1], t[2], t[3]};
  return q;
}
// homogeneous 4x4 product, each coefficient accumulated over k = 0..3 in order
void mul4(const double A[4][4], const double B[4][4], double C[4][4]) {
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      double acc = A[i][0] * B[0][j];
      for (int k = 1; k < 4; ++k) acc = acc + A[i][k] * B[k][j];
      C[i][j] = acc;
    }
}
}  // namespace

Sensors::Sensors(double hfov, double vfov, double fl, double nearR, double farR, double width, double height, const float sP[3],
                 const float sRPY[3])
    : hFOV(hfov), vFOV(vfov), focalLength(fl), near(nearR), far(farR), resWidth(width), resHeight(height) {
  for (int i = 0; i < 3; ++i) { sensorPose[i] = sP[i]; sensorRPY[i] = sRPY[i]; }
}

geometry_msgs::Pose Sensors::robot2sensorTransformation(const geometry_msgs::Pose& pose) const {
  Quat qr = {pose.orientation.x, pose.orientation.y, pose.orientation.z, pose.orientation.w};
  Mat3 R1 = mat_from_quat(qr);
  Mat3 R2 = mat_from_quat(quat_from_rpy(sensorRPY[0], sensorRPY[1], sensorRPY[2]));
  double A[4][4] = {{0}}, B[4][4] = {{0}}, C[4][4], D[4][4];
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) { A[i][j] = R1.m[i][j]; B[i][j] = R2.m[i][j]; }
  A[0][3] = pose.position.x; A[1][3] = pose.position.y; A[2][3] = pose.position.z; A[3][3] = 1;
  B[0][3] = sensorPose[0]; B[1][3] = sensorPose[1]; B[2][3] = sensorPose[2]; B[3][3] = 1;
  mul4(A, B, C);
  const double S[4][4] = {{1, 0, 0, 0}, {0, 0, -1, 0}, {0, 1, 0, 0}, {0, 0, 0, 1}};  // Rx(+90deg)
  mul4(C, S, D);
  Mat3 R3;
  for (int i = 0; i < 3; ++i)
    for (int j = 0; j < 3; ++j) R3.m[i][j] = D[i][j];
  Quat q = quat_from_mat(R3);
  geometry_msgs::Pose p;
  p.position.x = D[0][3]; p.position.y = D[1][3]; p.position.z = D[2][3];
  p.orientation.x = q.x; p.orientation.y = q.y; p.orientation.z = q.z; p.orientation.w = q.w;
  return p;
}
