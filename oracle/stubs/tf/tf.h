#pragma once
// stand-in for tf (LinearMath): only what the planner sources call.  Formulas are the published
// tf::Quaternion::setRPY / tf::Matrix3x3::setRotation / tf::getYaw ones, in double (tfScalar).
#include <cmath>
namespace tf {
typedef double tfScalar;
struct Vector3 { tfScalar v[3]; Vector3(tfScalar x = 0, tfScalar y = 0, tfScalar z = 0) { v[0] = x; v[1] = y; v[2] = z; }
  tfScalar x() const { return v[0]; } tfScalar y() const { return v[1]; } tfScalar z() const { return v[2]; } };
struct Quaternion {
  tfScalar q[4];
  Quaternion() { q[0] = q[1] = q[2] = 0; q[3] = 1; }
  Quaternion(tfScalar x, tfScalar y, tfScalar z, tfScalar w) { q[0] = x; q[1] = y; q[2] = z; q[3] = w; }
  void setRPY(tfScalar roll, tfScalar pitch, tfScalar yaw) {
    tfScalar halfYaw = yaw * 0.5, halfPitch = pitch * 0.5, halfRoll = roll * 0.5;
    tfScalar cosYaw = std::cos(halfYaw), sinYaw = std::sin(halfYaw);
    tfScalar cosPitch = std::cos(halfPitch), sinPitch = std::sin(halfPitch);
    tfScalar cosRoll = std::cos(halfRoll), sinRoll = std::sin(halfRoll);
    q[0] = sinRoll * cosPitch * cosYaw - cosRoll * sinPitch * sinYaw;
    q[1] = cosRoll * sinPitch * cosYaw + sinRoll * cosPitch * sinYaw;
    q[2] = cosRoll * cosPitch * sinYaw - sinRoll * sinPitch * cosYaw;
    q[3] = cosRoll * cosPitch * cosYaw + sinRoll * sinPitch * sinYaw;
  }
  tfScalar getX() const { return q[0]; } tfScalar getY() const { return q[1]; }
  tfScalar getZ() const { return q[2]; } tfScalar getW() const { return q[3]; }
  tfScalar x() const { return q[0]; } tfScalar y() const { return q[1]; }
  tfScalar z() const { return q[2]; } tfScalar w() const { return q[3]; }
};
inline Quaternion createQuaternionFromYaw(double yaw) { Quaternion q; q.setRPY(0.0, 0.0, yaw); return q; }
inline Quaternion createQuaternionFromRPY(double r, double p, double y) { Quaternion q; q.setRPY(r, p, y); return q; }
struct Matrix3x3 {
  tfScalar m[3][3];
  Matrix3x3() { for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) m[i][j] = (i == j); }
  explicit Matrix3x3(const Quaternion& q) { setRotation(q); }
  void setRotation(const Quaternion& q) {
    tfScalar d = q.q[0] * q.q[0] + q.q[1] * q.q[1] + q.q[2] * q.q[2] + q.q[3] * q.q[3];
    tfScalar s = 2.0 / d;
    tfScalar xs = q.x() * s, ys = q.y() * s, zs = q.z() * s;
    tfScalar wx = q.w() * xs, wy = q.w() * ys, wz = q.w() * zs;
    tfScalar xx = q.x() * xs, xy = q.x() * ys, xz = q.x() * zs;
    tfScalar yy = q.y() * ys, yz = q.y() * zs, zz = q.z() * zs;
    m[0][0] = 1.0 - (yy + zz); m[0][1] = xy - wz; m[0][2] = xz + wy;
    m[1][0] = xy + wz; m[1][1] = 1.0 - (xx + zz); m[1][2] = yz - wx;
    m[2][0] = xz - wy; m[2][1] = yz + wx; m[2][2] = 1.0 - (xx + yy);
  }
  // tf::Matrix3x3::getRotation (trace / largest-diagonal branches as published)
  void getRotation(Quaternion& q) const {
    tfScalar trace = m[0][0] + m[1][1] + m[2][2];
    tfScalar t[4];
    if (trace > 0.0) {
      tfScalar s = std::sqrt(trace + 1.0);
      t[3] = s * 0.5;
      s = 0.5 / s;
      t[0] = (m[2][1] - m[1][2]) * s;
      t[1] = (m[0][2] - m[2][0]) * s;
      t[2] = (m[1][0] - m[0][1]) * s;
    } else {
      int i = m[0][0] < m[1][1] ? (m[1][1] < m[2][2] ? 2 : 1) : (m[0][0] < m[2][2] ? 2 : 0);
      int j = (i + 1) % 3, k = (i + 2) % 3;
      tfScalar s = std::sqrt(m[i][i] - m[j][j] - m[k][k] + 1.0);
      t[i] = s * 0.5;
      s = 0.5 / s;
      t[3] = (m[k][j] - m[j][k]) * s;
      t[j] = (m[j][i] + m[i][j]) * s;
      t[k] = (m[k][i] + m[i][k]) * s;
    }
    q = Quaternion(t[0], t[1], t[2], t[3]);
  }
  void getRPY(tfScalar& roll, tfScalar& pitch, tfScalar& yaw) const {
    // single-solution branch of tf::Matrix3x3::getEulerYPR, enough for getYaw
    if (std::fabs(m[2][0]) >= 1) { yaw = 0; pitch = (m[2][0] < 0 ? 1 : -1) * M_PI / 2; roll = std::atan2(m[0][1], m[0][2]); return; }
    pitch = -std::asin(m[2][0]);
    roll = std::atan2(m[2][1] / std::cos(pitch), m[2][2] / std::cos(pitch));
    yaw = std::atan2(m[1][0] / std::cos(pitch), m[0][0] / std::cos(pitch));
  }
};
inline double getYaw(const Quaternion& q) { tfScalar r, p, y; Matrix3x3(q).getRPY(r, p, y); return y; }
}
