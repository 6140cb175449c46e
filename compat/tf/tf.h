// Stand-in for tf/tf.h (ROS geometry; its LinearMath is a copy of Bullet's), used ONLY when ROS is absent.
// Quaternion, Matrix3x3 and Vector3 with the formulas of tf's LinearMath headers, in double like tfScalar.
#pragma once
#include <cmath>
#include <geometry_msgs/Pose.h>
namespace tf {
typedef double tfScalar;
class Vector3 {
 public:
  Vector3() { m[0] = m[1] = m[2] = 0; }
  Vector3(double x, double y, double z) { m[0] = x; m[1] = y; m[2] = z; }
  double x() const { return m[0]; } double y() const { return m[1]; } double z() const { return m[2]; }
  double getX() const { return m[0]; } double getY() const { return m[1]; } double getZ() const { return m[2]; }
  void setValue(double x, double y, double z) { m[0] = x; m[1] = y; m[2] = z; }
  void setX(double v) { m[0] = v; } void setY(double v) { m[1] = v; } void setZ(double v) { m[2] = v; }
  double& operator[](int i) { return m[i]; }
  double operator[](int i) const { return m[i]; }
  double dot(const Vector3& v) const { return m[0] * v.m[0] + m[1] * v.m[1] + m[2] * v.m[2]; }
  double m[3];
};
class Quaternion {
 public:
  Quaternion() { q[0] = q[1] = q[2] = 0; q[3] = 1; }
  Quaternion(double x, double y, double z, double w) { q[0] = x; q[1] = y; q[2] = z; q[3] = w; }
  double x() const { return q[0]; } double y() const { return q[1]; } double z() const { return q[2]; } double w() const { return q[3]; }
  double getX() const { return q[0]; } double getY() const { return q[1]; } double getZ() const { return q[2]; } double getW() const { return q[3]; }
  void setValue(double x, double y, double z, double w) { q[0] = x; q[1] = y; q[2] = z; q[3] = w; }
  void setX(double v) { q[0] = v; } void setY(double v) { q[1] = v; } void setZ(double v) { q[2] = v; } void setW(double v) { q[3] = v; }
  // LinearMath/Quaternion.h setRPY
  void setRPY(double roll, double pitch, double yaw) {
    double halfYaw = yaw * 0.5, halfPitch = pitch * 0.5, halfRoll = roll * 0.5;
    double cosYaw = std::cos(halfYaw), sinYaw = std::sin(halfYaw), cosPitch = std::cos(halfPitch),
           sinPitch = std::sin(halfPitch), cosRoll = std::cos(halfRoll), sinRoll = std::sin(halfRoll);
    setValue(sinRoll * cosPitch * cosYaw - cosRoll * sinPitch * sinYaw,
             cosRoll * sinPitch * cosYaw + sinRoll * cosPitch * sinYaw,
             cosRoll * cosPitch * sinYaw - sinRoll * sinPitch * cosYaw,
             cosRoll * cosPitch * cosYaw + sinRoll * sinPitch * sinYaw);
  }
  double dot(const Quaternion& o) const { return q[0] * o.q[0] + q[1] * o.q[1] + q[2] * o.q[2] + q[3] * o.q[3]; }
  double length2() const { return dot(*this); }
  double length() const { return std::sqrt(length2()); }
  Quaternion& normalize() { double l = length(); for (int i = 0; i < 4; ++i) q[i] /= l; return *this; }
  // LinearMath/Quaternion.h angleShortestPath
  double angleShortestPath(const Quaternion& o) const {
    double s = std::sqrt(length2() * o.length2());
    if (dot(o) < 0) return std::acos(dot(Quaternion(-o.q[0], -o.q[1], -o.q[2], -o.q[3])) / s) * 2.0;
    return std::acos(dot(o) / s) * 2.0;
  }
  double q[4];
};
class Matrix3x3 {
 public:
  Matrix3x3() { for (int i = 0; i < 3; ++i) for (int j = 0; j < 3; ++j) m[i][j] = i == j; }
  explicit Matrix3x3(const Quaternion& q) { setRotation(q); }
  // LinearMath/Matrix3x3.h setRotation
  void setRotation(const Quaternion& q) {
    double d = q.length2();
    double s = 2.0 / d;
    double xs = q.x() * s, ys = q.y() * s, zs = q.z() * s;
    double wx = q.w() * xs, wy = q.w() * ys, wz = q.w() * zs;
    double xx = q.x() * xs, xy = q.x() * ys, xz = q.x() * zs;
    double yy = q.y() * ys, yz = q.y() * zs, zz = q.z() * zs;
    m[0][0] = 1.0 - (yy + zz); m[0][1] = xy - wz; m[0][2] = xz + wy;
    m[1][0] = xy + wz; m[1][1] = 1.0 - (xx + zz); m[1][2] = yz - wx;
    m[2][0] = xz - wy; m[2][1] = yz + wx; m[2][2] = 1.0 - (xx + yy);
  }
  // LinearMath/Matrix3x3.h getRotation
  void getRotation(Quaternion& q) const {
    double trace = m[0][0] + m[1][1] + m[2][2];
    double temp[4];
    if (trace > 0.0) {
      double s = std::sqrt(trace + 1.0);
      temp[3] = s * 0.5;
      s = 0.5 / s;
      temp[0] = (m[2][1] - m[1][2]) * s;
      temp[1] = (m[0][2] - m[2][0]) * s;
      temp[2] = (m[1][0] - m[0][1]) * s;
    } else {
      int i = m[0][0] < m[1][1] ? (m[1][1] < m[2][2] ? 2 : 1) : (m[0][0] < m[2][2] ? 2 : 0);
      int j = (i + 1) % 3, k = (i + 2) % 3;
      double s = std::sqrt(m[i][i] - m[j][j] - m[k][k] + 1.0);
      temp[i] = s * 0.5;
      s = 0.5 / s;
      temp[3] = (m[k][j] - m[j][k]) * s;
      temp[j] = (m[j][i] + m[i][j]) * s;
      temp[k] = (m[k][i] + m[i][k]) * s;
    }
    q.setValue(temp[0], temp[1], temp[2], temp[3]);
  }
  // LinearMath/Matrix3x3.h getEulerYPR, solution 1
  void getEulerYPR(double& yaw, double& pitch, double& roll) const {
    if (std::fabs(m[2][0]) >= 1) {
      yaw = 0;
      if (m[2][0] < 0) { pitch = M_PI / 2.0; roll = std::atan2(m[0][1], m[0][2]); }
      else { pitch = -M_PI / 2.0; roll = std::atan2(-m[0][1], -m[0][2]); }
    } else {
      pitch = -std::asin(m[2][0]);
      roll = std::atan2(m[2][1] / std::cos(pitch), m[2][2] / std::cos(pitch));
      yaw = std::atan2(m[1][0] / std::cos(pitch), m[0][0] / std::cos(pitch));
    }
  }
  void getRPY(double& roll, double& pitch, double& yaw) const { getEulerYPR(yaw, pitch, roll); }
  void setValue(double a, double b, double c, double d, double e, double f, double g, double h, double i) {
    m[0][0] = a; m[0][1] = b; m[0][2] = c; m[1][0] = d; m[1][1] = e; m[1][2] = f; m[2][0] = g; m[2][1] = h; m[2][2] = i;
  }
  void setRPY(double roll, double pitch, double yaw) { setEulerYPR(yaw, pitch, roll); }
  void setEulerYPR(double eulerZ, double eulerY, double eulerX) {
    double ci = std::cos(eulerX), cj = std::cos(eulerY), ch = std::cos(eulerZ);
    double si = std::sin(eulerX), sj = std::sin(eulerY), sh = std::sin(eulerZ);
    double cc = ci * ch, cs = ci * sh, sc = si * ch, ss = si * sh;
    setValue(cj * ch, sj * sc - cs, sj * cc + ss, cj * sh, sj * ss + cc, sj * cs - sc, -sj, cj * si, cj * ci);
  }
  const Vector3& operator[](int i) const { return m[i]; }
  Vector3& operator[](int i) { return m[i]; }
  const Vector3& getRow(int i) const { return m[i]; }
  Vector3 m[3];  // rows, as tf's m_el
};
inline Vector3 operator*(const Matrix3x3& M, const Vector3& v) {
  return Vector3(M[0][0] * v[0] + M[0][1] * v[1] + M[0][2] * v[2], M[1][0] * v[0] + M[1][1] * v[1] + M[1][2] * v[2],
                 M[2][0] * v[0] + M[2][1] * v[1] + M[2][2] * v[2]);
}
inline Vector3 operator+(const Vector3& a, const Vector3& b) { return Vector3(a[0] + b[0], a[1] + b[1], a[2] + b[2]); }
inline Quaternion createQuaternionFromRPY(double roll, double pitch, double yaw) { Quaternion q; q.setRPY(roll, pitch, yaw); return q; }
inline Quaternion createQuaternionFromYaw(double yaw) { Quaternion q; q.setRPY(0.0, 0.0, yaw); return q; }
inline geometry_msgs::Quaternion createQuaternionMsgFromYaw(double yaw) {
  Quaternion q = createQuaternionFromYaw(yaw);
  geometry_msgs::Quaternion m; m.x = q.x(); m.y = q.y(); m.z = q.z(); m.w = q.w();
  return m;
}
inline double getYaw(const Quaternion& q) { double y, p, r; Matrix3x3(q).getEulerYPR(y, p, r); return y; }
inline double getYaw(const geometry_msgs::Quaternion& m) { return getYaw(Quaternion(m.x, m.y, m.z, m.w)); }
inline void quaternionMsgToTF(const geometry_msgs::Quaternion& m, Quaternion& q) { q = Quaternion(m.x, m.y, m.z, m.w); }
inline void quaternionTFToMsg(const Quaternion& q, geometry_msgs::Quaternion& m) { m.x = q.x(); m.y = q.y(); m.z = q.z(); m.w = q.w(); }
class Transform {};
class TransformListener {};
}  // namespace tf
