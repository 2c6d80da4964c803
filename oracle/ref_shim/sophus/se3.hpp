// Minimal stand-in for <sophus/se3.hpp>: the SE3d surface the reference's alignment path uses (exp, log, inverse,
// composition, action on a point, rotationMatrix, translation), so its translation units compile unchanged
// (oracle/build_ref.py).  Test infrastructure.  Closed forms as Sophus publishes them: unit quaternion + translation,
// small-angle series below 1e-10, the rotation of a point through the quaternion (v + w * 2(q x v) + q x 2(q x v)).
// The reference pins no Sophus version; the product of two rotations is renormalised here.
#ifndef REF_SHIM_SOPHUS_SE3
#define REF_SHIM_SOPHUS_SE3
#include <Eigen/Dense>
#include <cmath>

namespace Sophus
{
template <class S>
class SE3
{
public:
  typedef Eigen::Matrix<S, 6, 1> Tangent;
  typedef Eigen::Matrix<S, 3, 1> Point;
  typedef Eigen::Matrix<S, 3, 3> Mat3;
  static constexpr S kEps = S(1e-10);

  SE3() : _w(1), _x(0), _y(0), _z(0) { _t.setZero(); }

  static SE3 exp(const Tangent & a)
  {
    const Point ups(a(0), a(1), a(2)), om(a(3), a(4), a(5));
    const S th2 = om.squaredNorm();
    S theta, imag, real;
    if (th2 < kEps * kEps) {
      theta = S(0);
      const S th4 = th2 * th2;
      imag = S(0.5) - S(1.0 / 48.0) * th2 + S(1.0 / 3840.0) * th4;
      real = S(1) - S(1.0 / 8.0) * th2 + S(1.0 / 384.0) * th4;
    } else {
      theta = std::sqrt(th2);
      const S half = S(0.5) * theta;
      imag = std::sin(half) / theta;
      real = std::cos(half);
    }
    SE3 T;
    T._w = real;
    T._x = imag * om.x();
    T._y = imag * om.y();
    T._z = imag * om.z();
    const Mat3 Om = hat(om), Om2 = Om * Om;
    Mat3 V;
    if (theta < kEps) V = T.rotationMatrix();
    else V = Mat3::Identity() + Om * ((S(1) - std::cos(theta)) / th2) + Om2 * ((theta - std::sin(theta)) / (th2 * theta));
    T._t = V * ups;
    return T;
  }

  Tangent log() const
  {
    const S n2 = _x * _x + _y * _y + _z * _z;
    S k, theta;  // omega = k * (x, y, z)
    if (n2 < kEps * kEps) {
      k = S(2) / _w - S(2.0 / 3.0) * n2 / (_w * _w * _w);
      theta = S(2) * n2 / _w;
    } else {
      const S n = std::sqrt(n2);
      if (std::abs(_w) < kEps) k = (_w > S(0) ? S(M_PI) : -S(M_PI)) / n;
      else k = S(2) * std::atan(n / _w) / n;
      theta = k * n;
    }
    const Point om(k * _x, k * _y, k * _z);
    const Mat3 Om = hat(om), Om2 = Om * Om;
    Mat3 Vinv;
    if (std::abs(theta) < kEps) {
      Vinv = Mat3::Identity() - Om * S(0.5) + Om2 * S(1.0 / 12.0);
    } else {
      const S half = S(0.5) * theta;
      Vinv = Mat3::Identity() - Om * S(0.5) + Om2 * ((S(1) - S(0.5) * theta * std::cos(half) / std::sin(half)) / (theta * theta));
    }
    const Point ups = Vinv * _t;
    Tangent out;
    out << ups.x(), ups.y(), ups.z(), om.x(), om.y(), om.z();
    return out;
  }

  SE3 inverse() const
  {
    SE3 I;
    I._w = _w;
    I._x = -_x;
    I._y = -_y;
    I._z = -_z;
    I._t = I.rotate(_t * S(-1));
    return I;
  }

  SE3 operator*(const SE3 & o) const
  {
    SE3 P;
    P._w = _w * o._w - _x * o._x - _y * o._y - _z * o._z;
    P._x = _w * o._x + _x * o._w + _y * o._z - _z * o._y;
    P._y = _w * o._y + _y * o._w + _z * o._x - _x * o._z;
    P._z = _w * o._z + _z * o._w + _x * o._y - _y * o._x;
    const S n = std::sqrt(P._w * P._w + P._x * P._x + P._y * P._y + P._z * P._z);
    P._w /= n;
    P._x /= n;
    P._y /= n;
    P._z /= n;
    P._t = _t + rotate(o._t);
    return P;
  }
  Point operator*(const Point & p) const { return rotate(p) + _t; }

  Mat3 rotationMatrix() const
  {
    const S tx = S(2) * _x, ty = S(2) * _y, tz = S(2) * _z;
    const S twx = tx * _w, twy = ty * _w, twz = tz * _w, txx = tx * _x, txy = ty * _x, txz = tz * _x, tyy = ty * _y,
            tyz = tz * _y, tzz = tz * _z;
    Mat3 R;
    R << S(1) - (tyy + tzz), txy - twz, txz + twy, txy + twz, S(1) - (txx + tzz), tyz - twx, txz - twy, tyz + twx,
      S(1) - (txx + tyy);
    return R;
  }
  const Point & translation() const { return _t; }
  Point & translation() { return _t; }

private:
  static Mat3 hat(const Point & o)
  {
    Mat3 H;
    H << S(0), -o.z(), o.y(), o.z(), S(0), -o.x(), -o.y(), o.x(), S(0);
    return H;
  }
  static Point cross(const Point & a, const Point & b)
  {
    return Point(a.y() * b.z() - a.z() * b.y(), a.z() * b.x() - a.x() * b.z(), a.x() * b.y() - a.y() * b.x());
  }
  Point rotate(const Point & v) const
  {
    const Point q(_x, _y, _z);
    Point uv = cross(q, v);
    uv += uv;
    return v + uv * _w + cross(q, uv);
  }
  S _w, _x, _y, _z;
  Point _t;
};
typedef SE3<double> SE3d;
}  // namespace Sophus
#endif
