// oracle/shim: TEST INFRASTRUCTURE ONLY. tf2::Vector3 / Quaternion / Matrix3x3 / Transform look-alikes (f64, tf2Scalar =
// double), restating the published arithmetic of ros/geometry2 tf2/LinearMath (version unpinned by the reference's
// package.xml:33-35, so this sequence is "parity unpinned"): Quaternion::setRPY, Matrix3x3::setRotation /
// getRotation (Shepperd branches) / getEulerYPR / operator*, Quaternion product, Transform::operator* / inverse. It exists
// so that the reference's sensors/point_cloud_scanner.cpp and node/*.cpp compile UNMODIFIED into oracle/_ref; everything
// those files compute themselves is then the reference's own code.
#pragma once
#include <cmath>
namespace tf2
{
typedef double tf2Scalar;
class Vector3
{
public:
  double m[3];
  Vector3() : m{0.0, 0.0, 0.0} {}
  Vector3(double x, double y, double z) : m{x, y, z} {}
  double x() const { return m[0]; }
  double y() const { return m[1]; }
  double z() const { return m[2]; }
  double getX() const { return m[0]; }
  double getY() const { return m[1]; }
  double getZ() const { return m[2]; }
  Vector3 operator-() const { return Vector3(-m[0], -m[1], -m[2]); }
  double dot(const Vector3& v) const { return m[0] * v.m[0] + m[1] * v.m[1] + m[2] * v.m[2]; }
  double operator[](int i) const { return m[i]; }
};
class Quaternion
{
public:
  double q[4];  // x, y, z, w
  Quaternion() : q{0.0, 0.0, 0.0, 1.0} {}
  Quaternion(double x, double y, double z, double w) : q{x, y, z, w} {}
  double x() const { return q[0]; }
  double y() const { return q[1]; }
  double z() const { return q[2]; }
  double w() const { return q[3]; }
  double length2() const { return q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]; }
  Quaternion operator*(const Quaternion& o) const  // tf2::operator*(Quaternion, Quaternion)
  {
    return Quaternion(w() * o.x() + x() * o.w() + y() * o.z() - z() * o.y(), w() * o.y() + y() * o.w() + z() * o.x() - x() * o.z(),
                      w() * o.z() + z() * o.w() + x() * o.y() - y() * o.x(), w() * o.w() - x() * o.x() - y() * o.y() - z() * o.z());
  }
  void setRPY(double roll, double pitch, double yaw)
  {
    const double halfYaw = yaw * 0.5, halfPitch = pitch * 0.5, halfRoll = roll * 0.5;
    const double cosYaw = std::cos(halfYaw), sinYaw = std::sin(halfYaw);
    const double cosPitch = std::cos(halfPitch), sinPitch = std::sin(halfPitch);
    const double cosRoll = std::cos(halfRoll), sinRoll = std::sin(halfRoll);
    q[0] = sinRoll * cosPitch * cosYaw - cosRoll * sinPitch * sinYaw;
    q[1] = cosRoll * sinPitch * cosYaw + sinRoll * cosPitch * sinYaw;
    q[2] = cosRoll * cosPitch * sinYaw - sinRoll * sinPitch * cosYaw;
    q[3] = cosRoll * cosPitch * cosYaw + sinRoll * sinPitch * sinYaw;
  }
};
class Matrix3x3
{
public:
  Vector3 el[3];  // rows
  Matrix3x3() { el[0] = Vector3(1, 0, 0); el[1] = Vector3(0, 1, 0); el[2] = Vector3(0, 0, 1); }
  explicit Matrix3x3(const Quaternion& q) { setRotation(q); }
  Matrix3x3(double xx, double xy, double xz, double yx, double yy, double yz, double zx, double zy, double zz)
  {
    el[0] = Vector3(xx, xy, xz);
    el[1] = Vector3(yx, yy, yz);
    el[2] = Vector3(zx, zy, zz);
  }
  const Vector3& operator[](int i) const { return el[i]; }
  void setRotation(const Quaternion& q)
  {
    const double d = q.length2();
    const double s = 2.0 / d;
    const double xs = q.x() * s, ys = q.y() * s, zs = q.z() * s;
    const double wx = q.w() * xs, wy = q.w() * ys, wz = q.w() * zs;
    const double xx = q.x() * xs, xy = q.x() * ys, xz = q.x() * zs;
    const double yy = q.y() * ys, yz = q.y() * zs, zz = q.z() * zs;
    el[0] = Vector3(1.0 - (yy + zz), xy - wz, xz + wy);
    el[1] = Vector3(xy + wz, 1.0 - (xx + zz), yz - wx);
    el[2] = Vector3(xz - wy, yz + wx, 1.0 - (xx + yy));
  }
  void getRotation(Quaternion& out) const
  {
    const double trace = el[0].x() + el[1].y() + el[2].z();
    double temp[4];
    if (trace > 0.0)
    {
      double s = std::sqrt(trace + 1.0);
      temp[3] = s * 0.5;
      s = 0.5 / s;
      temp[0] = (el[2].y() - el[1].z()) * s;
      temp[1] = (el[0].z() - el[2].x()) * s;
      temp[2] = (el[1].x() - el[0].y()) * s;
    }
    else
    {
      const int i = el[0].x() < el[1].y() ? (el[1].y() < el[2].z() ? 2 : 1) : (el[0].x() < el[2].z() ? 2 : 0);
      const int j = (i + 1) % 3, k = (i + 2) % 3;
      double s = std::sqrt(el[i][i] - el[j][j] - el[k][k] + 1.0);
      temp[i] = s * 0.5;
      s = 0.5 / s;
      temp[3] = (el[k][j] - el[j][k]) * s;
      temp[j] = (el[j][i] + el[i][j]) * s;
      temp[k] = (el[k][i] + el[i][k]) * s;
    }
    out = Quaternion(temp[0], temp[1], temp[2], temp[3]);
  }
  // Matrix3x3::getEulerYPR, solution 1 (the default), and getRPY = the same with the outputs reordered
  void getEulerYPR(double& yaw, double& pitch, double& roll) const
  {
    const double pi = 3.1415926535897932384626433832795029;  // TF2SIMD_PI
    if (std::fabs(el[2].x()) >= 1.0)
    {
      yaw = 0.0;
      const double delta = std::atan2(el[2].y(), el[2].z());
      pitch = el[2].x() < 0.0 ? pi / 2.0 : -pi / 2.0;
      roll = delta;
    }
    else
    {
      double a = el[2].x();  // tf2Asin clamps to [-1, 1]
      if (a < -1.0) a = -1.0;
      if (a > 1.0) a = 1.0;
      pitch = -std::asin(a);
      roll = std::atan2(el[2].y() / std::cos(pitch), el[2].z() / std::cos(pitch));
      yaw = std::atan2(el[1].x() / std::cos(pitch), el[0].x() / std::cos(pitch));
    }
  }
  void getRPY(double& roll, double& pitch, double& yaw) const { getEulerYPR(yaw, pitch, roll); }
  Matrix3x3 transpose() const
  {
    return Matrix3x3(el[0].x(), el[1].x(), el[2].x(), el[0].y(), el[1].y(), el[2].y(), el[0].z(), el[1].z(), el[2].z());
  }
  Vector3 operator*(const Vector3& v) const { return Vector3(el[0].dot(v), el[1].dot(v), el[2].dot(v)); }
  double tdotx(const Vector3& v) const { return el[0].x() * v.x() + el[1].x() * v.y() + el[2].x() * v.z(); }
  double tdoty(const Vector3& v) const { return el[0].y() * v.x() + el[1].y() * v.y() + el[2].y() * v.z(); }
  double tdotz(const Vector3& v) const { return el[0].z() * v.x() + el[1].z() * v.y() + el[2].z() * v.z(); }
};
inline Matrix3x3 operator*(const Matrix3x3& m1, const Matrix3x3& m2)
{
  return Matrix3x3(m2.tdotx(m1[0]), m2.tdoty(m1[0]), m2.tdotz(m1[0]), m2.tdotx(m1[1]), m2.tdoty(m1[1]), m2.tdotz(m1[1]),
                   m2.tdotx(m1[2]), m2.tdoty(m1[2]), m2.tdotz(m1[2]));
}
class Transform
{
public:
  Matrix3x3 basis;
  Vector3 origin;
  Transform() {}
  Transform(const Quaternion& q, const Vector3& c = Vector3(0, 0, 0)) : basis(q), origin(c) {}
  Transform(const Matrix3x3& b, const Vector3& c) : basis(b), origin(c) {}
  Vector3 operator()(const Vector3& x) const
  {
    return Vector3(basis[0].dot(x) + origin.x(), basis[1].dot(x) + origin.y(), basis[2].dot(x) + origin.z());
  }
  Transform operator*(const Transform& t) const { return Transform(basis * t.basis, (*this)(t.origin)); }
  const Matrix3x3& getBasis() const { return basis; }
  const Vector3& getOrigin() const { return origin; }
  void setOrigin(const Vector3& o) { origin = o; }
  void setRotation(const Quaternion& q) { basis.setRotation(q); }
  void setIdentity()
  {
    basis = Matrix3x3();
    origin = Vector3(0.0, 0.0, 0.0);
  }
  Transform inverse() const
  {
    const Matrix3x3 inv = basis.transpose();
    return Transform(inv, inv * -origin);
  }
  Quaternion getRotation() const
  {
    Quaternion q;
    basis.getRotation(q);
    return q;
  }
};
}  // namespace tf2
