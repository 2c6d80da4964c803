// vslam_b200.hpp — C++17 host mirror of the reference's aligner interface, header-only, on top of the C ABI
// (vslam_b200.h).  Same class and method names, argument meaning and error behaviour as the reference for this
// path, without Eigen / Sophus / OpenCV (paths relative to /root/reference/src/vslam/src/):
//
//   pd::vslam::Camera                     core/src/Camera.h                  -> vslam_b200::Camera
//   pd::vslam::PoseWithCovariance         core/src/PoseWithCovariance.h:31-43 -> vslam_b200::PoseWithCovariance
//   pd::vslam::Frame                      core/src/Frame.h:35-37,43,94        -> vslam_b200::Frame
//   least_squares::Scaler / Loss family   least_squares/src/Scaler.h:30-74, Loss.h:27-116 (compute, computeWeight, computeScale)
//   least_squares::NormalEquations        least_squares/src/NormalEquations.h:26-59
//   least_squares::Problem / Solver::Results / GaussNewton::solve   least_squares/src/Solver.h:22-63, GaussNewton.h:35
//   least_squares::Prior, MotionPrior     least_squares/src/Prior.h, SE3Alignment.cpp:27-52
//   lukas_kanade::InverseCompositional    lukas_kanade/src/InverseCompositional.h:30-88 (normal equations on the device)
//   pd::vslam::AlignmentSE3               odometry/src/AlignmentSE3.h:22-26
//   pd::vslam::SE3Alignment               odometry/src/direct_image_alignment/SE3Alignment.h:25-46
//
// Differences that follow from the device: frames are uploaded on first use and stay resident (keyed by
// Frame::id()) until the Frame object dies; computePyramid / computeDerivatives / computePcl only record the
// request (the kernels run at upload); SE3Alignment::align(batch) is an addition for throughput.
// Errors are std::runtime_error (the reference's pd::Exception derives from it, core/src/Exceptions.h:27-31), so
// OdometryRgbd's `catch (const std::runtime_error &)` (odometry/src/Odometry.cpp:52-56) keeps working.
#ifndef VSLAM_B200_HPP_
#define VSLAM_B200_HPP_

#include <algorithm>
#include <array>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <memory>
#include <limits>
#include <mutex>
#include <sstream>
#include <stdexcept>
#include <string>
#include <unordered_map>
#include <utility>
#include <vector>

#include "vslam_b200.h"

namespace vslam_b200
{
using Timestamp = std::uint64_t;

struct Exception : std::runtime_error
{
  explicit Exception(const std::string & m) : std::runtime_error(m) {}
};
inline void check(int code)
{
  if (code != 0) throw Exception(std::string("vslam_b200: ") + vslamb200_last_error());
}

// Dense image; colMajor == true is the memory order of the reference's Eigen types (core/src/types.h:33-35).
template <typename T>
struct Matrix
{
  int rows = 0, cols = 0;
  bool colMajor = false;
  std::vector<T> data;
  Matrix() = default;
  Matrix(int r, int c, bool colMajor_ = false) : rows(r), cols(c), colMajor(colMajor_), data((size_t)r * c) {}
  T & operator()(int r, int c) { return data[colMajor ? (size_t)c * rows + r : (size_t)r * cols + c]; }
  const T & operator()(int r, int c) const { return data[colMajor ? (size_t)c * rows + r : (size_t)r * cols + c]; }
};
using Image = Matrix<std::uint8_t>;
using DepthMap = Matrix<double>;   // the reference's DepthMap is double
using DepthMapF = Matrix<float>;   // sensor-native alternative (halves the upload)

// Sophus::SE3d::data() order: unit quaternion x, y, z, w, then translation.
struct SE3d
{
  std::array<double, 7> q{{0, 0, 0, 1, 0, 0, 0}};
  const double * data() const { return q.data(); }
  double * data() { return q.data(); }
};
using Matd6 = std::array<double, 36>;  // row-major 6x6
using Vec6d = std::array<double, 6>;

// ------------------------------------------------------------------------------------------------
// SE3 algebra with Sophus' conventions: tangent [upsilon; omega], storage unit quaternion (x y z w) + translation
// ------------------------------------------------------------------------------------------------
namespace se3
{
inline void rotate(const double * q, const double * p, double * o)
{
  // o = R(q) p through t = 2 q_v x p:  o = p + w t + q_v x t
  const double x = q[0], y = q[1], z = q[2], w = q[3];
  const double tx = 2 * (y * p[2] - z * p[1]), ty = 2 * (z * p[0] - x * p[2]), tz = 2 * (x * p[1] - y * p[0]);
  o[0] = p[0] + w * tx + (y * tz - z * ty);
  o[1] = p[1] + w * ty + (z * tx - x * tz);
  o[2] = p[2] + w * tz + (x * ty - y * tx);
}
inline SE3d mul(const SE3d & a, const SE3d & b)
{
  SE3d o;
  const double * p = a.data();
  const double * r = b.data();
  double q[4] = {p[3] * r[0] + p[0] * r[3] + p[1] * r[2] - p[2] * r[1], p[3] * r[1] + p[1] * r[3] + p[2] * r[0] - p[0] * r[2],
                 p[3] * r[2] + p[2] * r[3] + p[0] * r[1] - p[1] * r[0], p[3] * r[3] - p[0] * r[0] - p[1] * r[1] - p[2] * r[2]};
  const double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  for (int i = 0; i < 4; i++) o.data()[i] = q[i] / n;
  double rt[3];
  rotate(p, r + 4, rt);
  for (int i = 0; i < 3; i++) o.data()[4 + i] = rt[i] + p[4 + i];
  return o;
}
inline SE3d inverse(const SE3d & a)
{
  SE3d o;
  o.data()[0] = -a.data()[0];
  o.data()[1] = -a.data()[1];
  o.data()[2] = -a.data()[2];
  o.data()[3] = a.data()[3];
  double rt[3];
  rotate(o.data(), a.data() + 4, rt);
  for (int i = 0; i < 3; i++) o.data()[4 + i] = -rt[i];
  return o;
}
inline void cross(const double * a, const double * b, double * o)
{
  o[0] = a[1] * b[2] - a[2] * b[1];
  o[1] = a[2] * b[0] - a[0] * b[2];
  o[2] = a[0] * b[1] - a[1] * b[0];
}
inline SE3d exp(const Vec6d & xi)
{
  const double * u = xi.data();
  const double * w = xi.data() + 3;
  const double th2 = w[0] * w[0] + w[1] * w[1] + w[2] * w[2], th = std::sqrt(th2);
  double imag, real, B, C;
  if (th2 < 1e-20) {
    imag = 0.5 - th2 / 48.0 + th2 * th2 / 3840.0;
    real = 1.0 - th2 / 8.0 + th2 * th2 / 384.0;
    B = 0.5;
    C = 1.0 / 6.0;
  } else {
    imag = std::sin(0.5 * th) / th;
    real = std::cos(0.5 * th);
    B = (1.0 - std::cos(th)) / th2;
    C = (th - std::sin(th)) / (th2 * th);
  }
  SE3d o;
  for (int i = 0; i < 3; i++) o.data()[i] = imag * w[i];
  o.data()[3] = real;
  double c1[3], c2[3];
  cross(w, u, c1);
  cross(w, c1, c2);
  for (int i = 0; i < 3; i++) o.data()[4 + i] = u[i] + B * c1[i] + C * c2[i];  // V(omega) upsilon
  return o;
}
inline Vec6d log(const SE3d & T)
{
  const double * q = T.data();
  const double n2 = q[0] * q[0] + q[1] * q[1] + q[2] * q[2], n = std::sqrt(n2), w = q[3];
  double f;
  if (n2 < 1e-20)
    f = 2.0 / w - 2.0 / 3.0 * n2 / (w * w * w);
  else if (std::fabs(w) < 1e-10)
    f = (w > 0 ? M_PI : -M_PI) / n;
  else
    f = 2.0 * std::atan(n / w) / n;
  const double om[3] = {f * q[0], f * q[1], f * q[2]};
  const double th2 = om[0] * om[0] + om[1] * om[1] + om[2] * om[2], th = std::sqrt(th2);
  double c;
  if (th2 < 1e-12)
    c = 1.0 / 12.0;
  else
    c = (1.0 - th * std::cos(0.5 * th) / (2.0 * std::sin(0.5 * th))) / th2;
  double c1[3], c2[3];
  cross(om, q + 4, c1);
  cross(om, c1, c2);
  Vec6d xi;
  for (int i = 0; i < 3; i++) {
    xi[i] = q[4 + i] - 0.5 * c1[i] + c * c2[i];  // V^-1(omega) t
    xi[3 + i] = om[i];
  }
  return xi;
}
}  // namespace se3


class PoseWithCovariance
{
public:
  typedef std::shared_ptr<PoseWithCovariance> ShPtr;
  typedef std::shared_ptr<const PoseWithCovariance> ConstShPtr;
  typedef std::unique_ptr<PoseWithCovariance> UnPtr;
  PoseWithCovariance() { _cov.fill(0.0); for (int i = 0; i < 6; i++) _cov[7 * i] = 1.0; }
  PoseWithCovariance(const SE3d & pose, const Matd6 & cov) : _pose(pose), _cov(cov) {}
  const SE3d & pose() const { return _pose; }
  SE3d & pose() { return _pose; }
  const Matd6 & cov() const { return _cov; }

private:
  SE3d _pose;
  Matd6 _cov;
};

class Camera
{
public:
  typedef std::shared_ptr<const Camera> ConstShPtr;
  Camera(double f, double cx, double cy) : _fx(f), _fy(f), _cx(cx), _cy(cy) {}
  Camera(double fx, double fy, double cx, double cy) : _fx(fx), _fy(fy), _cx(cx), _cy(cy) {}
  double fx() const { return _fx; }
  double fy() const { return _fy; }
  double cx() const { return _cx; }
  double cy() const { return _cy; }

private:
  double _fx, _fy, _cx, _cy;
};

// Host-side frame: owns the level-0 intensity and depth, the camera and the pose.  Device residency is managed by the
// aligner that uses it.
class Frame
{
public:
  typedef std::shared_ptr<Frame> ShPtr;
  typedef std::shared_ptr<const Frame> ConstShPtr;

  Frame(Image intensity, DepthMap depth, Camera::ConstShPtr cam, Timestamp t = 0,
        const PoseWithCovariance & pose = PoseWithCovariance())
  : _id(nextId()), _intensity(std::move(intensity)), _depth(std::move(depth)), _cam(std::move(cam)), _t(t), _pose(pose)
  {
    validate(_depth.rows, _depth.cols, _depth.colMajor);
  }
  Frame(Image intensity, DepthMapF depth, Camera::ConstShPtr cam, Timestamp t = 0,
        const PoseWithCovariance & pose = PoseWithCovariance())
  : _id(nextId()), _intensity(std::move(intensity)), _depthF(std::move(depth)), _cam(std::move(cam)), _t(t), _pose(pose)
  {
    validate(_depthF.rows, _depthF.cols, _depthF.colMajor);
  }

  std::uint64_t id() const { return _id; }                       // Frame.h:43
  Timestamp t() const { return _t; }
  const PoseWithCovariance & pose() const { return _pose; }      // P_frame = pose * P_world (Frame.h:94)
  void set(const PoseWithCovariance & pose) { _pose = pose; }
  size_t width(size_t level = 0) const { return (size_t)((_intensity.cols + (1 << level) - 1) >> level); }
  size_t height(size_t level = 0) const { return (size_t)((_intensity.rows + (1 << level) - 1) >> level); }
  size_t nLevels() const { return _nLevels; }
  Camera::ConstShPtr camera() const { return _cam; }
  // Camera::resize(cam, 0.5^level) (Camera.cpp:53-57, Frame.cpp:267): fx, fy, cx, cy each halved per level
  Camera::ConstShPtr camera(size_t level) const
  {
    const double s = std::pow(0.5, (double)level);
    return std::make_shared<const Camera>(_cam->fx() * s, _cam->fy() * s, _cam->cx() * s, _cam->cy() * s);
  }
  const Image & intensity() const { return _intensity; }
  // Per-level accessors (Frame.h:45-47,80,83).  The pyramid lives on the device: the level is read back once
  // (vslamb200_frame_download) and kept on the host.  They need a frame that is resident - one that has been aligned, or
  // handed to SE3Alignment::upload - and throw otherwise.  dIx / dIy / depth are widened to double like the reference's.
  const Image & intensity(size_t level) const { return level == 0 ? _intensity : fetched(level).I; }
  const Matrix<double> & dIx(size_t level = 0) const { return fetched(level).gx; }
  const Matrix<double> & dIy(size_t level = 0) const { return fetched(level).gy; }
  const DepthMap & depth(size_t level) const { return fetched(level).Z; }
  // Frame::computePcl (Frame.cpp:233-253): p = K_l^-1 (u, v, 1) d for finite d > 0, else 0; index v * width + u
  std::vector<std::array<double, 3>> pcl(size_t level = 0) const
  {
    const Level & L = fetched(level);
    const auto cam = camera(level);
    std::vector<std::array<double, 3>> out((size_t)L.Z.rows * L.Z.cols);
    for (int v = 0; v < L.Z.rows; v++)
      for (int u = 0; u < L.Z.cols; u++) {
        const double d = L.Z(v, u);
        std::array<double, 3> p{{0.0, 0.0, 0.0}};
        if (std::isfinite(d) && d > 0.0) p = {{(u - cam->cx()) / cam->fx() * d, (v - cam->cy()) / cam->fy() * d, d}};
        out[(size_t)v * L.Z.cols + u] = p;
      }
    return out;
  }
  // called by the aligner that made the frame resident (and again whenever it is re-uploaded)
  void attach(vslamb200_ctx * ctx) const
  {
    _ctx = ctx;
    _levels.clear();
  }
  bool depthIsFloat() const { return !_depthF.data.empty(); }
  const DepthMap & depth() const { return _depth; }
  const DepthMapF & depthF() const { return _depthF; }

  // The reference's caller runs these three, in this order, before aligning (NodeRgbdAlignment.cpp:194-197).  Here
  // they record the request; pyr_down / depth_down / blur_sobel run on the device when the frame is uploaded.
  void computePyramid(size_t nLevels)
  {
    if (nLevels < 1 || nLevels > VSLAMB200_MAX_LEVELS) throw Exception("computePyramid: nLevels out of range");
    _nLevels = nLevels;
    _version++;
  }
  void computeDerivatives() {}
  void computePcl() {}
  std::uint64_t version() const { return _version; }

private:
  static std::uint64_t nextId()
  {
    static std::atomic<std::uint64_t> ctr{0};
    return ctr++;
  }
  void validate(int drows, int dcols, bool dColMajor) const
  {
    // Frame.cpp:150-155
    if (!_cam || _intensity.cols != dcols || _intensity.rows != drows || _intensity.colMajor != dColMajor ||
        std::abs(_intensity.cols / 2 - _cam->cx()) > 10)
      throw Exception("Inconsistent camera parameters / image / depth dimensions detected.");
  }
  struct Level
  {
    Image I;
    Matrix<double> gx, gy;
    DepthMap Z;
  };
  const Level & fetched(size_t level) const
  {
    if (!_ctx) throw Exception("Frame: pyramid levels live on the device; align the frame or call SE3Alignment::upload first");
    auto it = _levels.find(level);
    if (it != _levels.end()) return it->second;
    int d[4];
    check(vslamb200_frame_dims(_ctx, _id, (int)level, d));
    std::vector<std::uint8_t> I((size_t)d[0] * d[1]);
    std::vector<float> Z((size_t)d[2] * d[3]);
    std::vector<std::int16_t> gx(I.size()), gy(I.size());
    check(vslamb200_frame_download(_ctx, _id, (int)level, I.data(), Z.data(), nullptr, gx.data(), gy.data()));
    Level L;
    L.I = Image(d[1], d[0]);
    L.gx = Matrix<double>(d[1], d[0]);
    L.gy = Matrix<double>(d[1], d[0]);
    L.Z = DepthMap(d[3], d[2]);
    L.I.data = I;
    for (size_t i = 0; i < I.size(); i++) {
      L.gx.data[i] = gx[i];
      L.gy.data[i] = gy[i];
    }
    for (size_t i = 0; i < Z.size(); i++) L.Z.data[i] = Z[i];
    return _levels.emplace(level, std::move(L)).first->second;
  }
  std::uint64_t _id;
  mutable vslamb200_ctx * _ctx = nullptr;
  mutable std::unordered_map<size_t, Level> _levels;
  Image _intensity;
  DepthMap _depth;
  DepthMapF _depthF;
  Camera::ConstShPtr _cam;
  Timestamp _t;
  PoseWithCovariance _pose;
  size_t _nLevels = 1;
  std::uint64_t _version = 0;
};

namespace least_squares
{
// Stand-ins for Eigen::VectorXd / Eigen::MatrixXd on this path (the interfaces below take and return them where the
// reference's take Eigen types; core/src/types.h:56-63)
using VecXd = std::vector<double>;
class MatXd
{
public:
  MatXd() = default;
  MatXd(size_t rows, size_t cols) : _rows(rows), _cols(cols), _d(rows * cols, 0.0) {}
  static MatXd Zero(size_t rows, size_t cols) { return MatXd(rows, cols); }
  static MatXd Identity(size_t n)
  {
    MatXd m(n, n);
    for (size_t i = 0; i < n; i++) m(i, i) = 1.0;
    return m;
  }
  size_t rows() const { return _rows; }
  size_t cols() const { return _cols; }
  double & operator()(size_t r, size_t c) { return _d[r * _cols + c]; }
  const double & operator()(size_t r, size_t c) const { return _d[r * _cols + c]; }
  VecXd row(size_t r) const { return VecXd(_d.begin() + r * _cols, _d.begin() + (r + 1) * _cols); }
  void setRow(size_t r, const VecXd & v) { std::copy(v.begin(), v.end(), _d.begin() + r * _cols); }
  const std::vector<double> & data() const { return _d; }

private:
  size_t _rows = 0, _cols = 0;
  std::vector<double> _d;
};

namespace detail
{
// algorithm::median (core/src/algorithm.cpp:109-120) as the reference has it: sort, odd count -> the middle element,
// even count -> (v[n/2-1] + v[n/2+1]) / 2 (sic: not the two middle elements); n == 2 reads past the end in the reference
// and is defined here as the smaller sample; empty -> 0
inline double median(VecXd v)
{
  const size_t n = v.size();
  if (n == 0) return 0.0;
  std::sort(v.begin(), v.end());
  if (n % 2 == 1) return v[n / 2];
  if (n == 2) return v[0];
  return (v[n / 2 - 1] + v[n / 2 + 1]) / 2.0;
}
inline double meanAbsDeviationScale(const VecXd & r, double centre)  // sqrt(sum |r - centre| / (n - 1)), Scaler.cpp:33,44
{
  double s = 0.0;
  for (double x : r) s += std::fabs(x - centre);
  return std::sqrt(s / ((double)r.size() - 1.0));
}
}  // namespace detail

// Scaler.h:30-74, Scaler.cpp:21-67.  compute() is the host evaluation (arbitrary residual vectors); on the device the
// same scalers are evaluated on the residual histogram (csrc/gn_level.cu compute_scale).
class Scaler
{
public:
  typedef std::shared_ptr<Scaler> ShPtr;
  typedef std::shared_ptr<const Scaler> ConstShPtr;
  struct Scale
  {
    double offset;
    double scale;
  };
  virtual ~Scaler() = default;
  virtual Scale compute(const VecXd &) const { return {0.0, 1.0}; }
  virtual int kind() const { return VSLAMB200_SCALER_IDENTITY; }
  virtual double v() const { return 5.0; }
  // the device scalers run the reference's defaults (30 iterations, 1e-5): anything else cannot be honoured there
  virtual bool deviceCompatible() const { return true; }
};
class MedianScaler : public Scaler
{
public:
  Scale compute(const VecXd & r) const override
  {
    if (r.empty()) return {0.0, 1.0};
    const double m = detail::median(r);
    return {m, detail::meanAbsDeviationScale(r, m)};
  }
  int kind() const override { return VSLAMB200_SCALER_MEDIAN; }
};
class MeanScaler : public Scaler
{
public:
  Scale compute(const VecXd & r) const override
  {
    if (r.empty()) return {0.0, 1.0};  // Scaler.cpp:39-42
    double s = 0.0;
    for (double x : r) s += x;
    const double m = s / (double)r.size();
    return {m, detail::meanAbsDeviationScale(r, m)};
  }
  int kind() const override { return VSLAMB200_SCALER_MEAN; }
};
class ScalerTDistribution : public Scaler
{
public:
  explicit ScalerTDistribution(double v = 5.0, std::uint64_t maxIterations = 30, double minStepSize = 1e-5)
  : _v(v), _maxIterations(maxIterations), _minStepSize(minStepSize)
  {
  }
  // fixed point sigma^2 <- mean(r^2 (v + 1) / (v + (r / sigma)^2)) from sigma = 1 (Scaler.cpp:49-67)
  Scale compute(const VecXd & r) const override
  {
    double sigma = 1.0, step = std::numeric_limits<double>::max();
    for (std::uint64_t it = 0; it < _maxIterations && step > _minStepSize; it++) {
      double acc = 0.0;
      for (double x : r) acc += x * x * (_v + 1) / (_v + std::pow(x / sigma, 2));
      const double next = std::sqrt(acc / (double)r.size());
      step = std::fabs(sigma - next);
      sigma = next;
    }
    return {0.0, sigma};
  }
  int kind() const override { return VSLAMB200_SCALER_TDIST; }
  double v() const override { return _v; }
  bool deviceCompatible() const override { return _maxIterations == 30 && _minStepSize == 1e-5; }

private:
  double _v;
  std::uint64_t _maxIterations;
  double _minStepSize;
};

// Loss.h:27-116, Loss.cpp:23-100: compute / computeDerivative / computeWeight of a scaled residual, computeScale of a
// residual vector.  Constructor signatures and defaults are the reference's.
class Loss
{
public:
  typedef std::shared_ptr<Loss> ShPtr;
  typedef std::shared_ptr<const Loss> ConstShPtr;
  Loss(Scaler::ShPtr scaler = std::make_shared<Scaler>()) : _scaler(std::move(scaler)) {}
  virtual ~Loss() = default;
  virtual double compute(double r) const = 0;
  virtual double computeDerivative(double r) const = 0;
  virtual double computeWeight(double r) const { return computeDerivative(r) / r; }
  virtual Scaler::Scale computeScale(const VecXd & residuals) const { return _scaler->compute(residuals); }
  virtual int kind() const = 0;
  virtual double c() const { return 1.345; }
  virtual double v() const { return 5.0; }
  const Scaler & scaler() const { return *_scaler; }

private:
  Scaler::ConstShPtr _scaler;
};
class QuadraticLoss : public Loss
{
public:
  QuadraticLoss(Scaler::ShPtr scaler = std::make_shared<Scaler>()) : Loss(std::move(scaler)) {}
  double compute(double r) const override { return 0.5 * r * r; }
  double computeDerivative(double r) const override { return r; }
  double computeWeight(double) const override { return 1.0; }
  int kind() const override { return VSLAMB200_LOSS_QUADRATIC; }
};
class TukeyLoss : public Loss
{
public:
  static constexpr double C = 4.6851;  // 95 % asymptotic efficiency on the standard normal distribution
  static constexpr double C2_6 = C / 6.0;
  TukeyLoss(Scaler::ShPtr scaler = std::make_shared<MedianScaler>()) : Loss(std::move(scaler)) {}
  double compute(double r) const override { return inside(r) ? C2_6 * (1.0 - std::pow(bell(r), 3)) : C2_6; }
  double computeDerivative(double r) const override { return inside(r) ? r * std::pow(bell(r), 2) : 0.0; }
  double computeWeight(double r) const override { return inside(r) ? std::pow(bell(r), 2) : 0.0; }
  int kind() const override { return VSLAMB200_LOSS_TUKEY; }

private:
  static bool inside(double r) { return std::fabs(r) < C; }
  static double bell(double r) { return 1.0 - (r / C) * (r / C); }
};
class HuberLoss : public Loss
{
public:
  HuberLoss(Scaler::ShPtr scaler = std::make_shared<MedianScaler>(), double c = 1.345f) : Loss(std::move(scaler)), _c(c) {}
  const double _c;
  // outside the quadratic zone the reference's weight is sign(c r) / r, i.e. 1 / |r| for c > 0 (Loss.cpp:65-72)
  double computeWeight(double r) const override { return std::fabs(r) < _c ? 1.0 : sgn(r) / r; }
  double computeDerivative(double r) const override { return std::fabs(r) < _c ? r : sgn(r); }
  double compute(double r) const override { return std::fabs(r) < _c ? 0.5 * r * r : _c * std::fabs(r) - 0.5 * r * r; }
  int kind() const override { return VSLAMB200_LOSS_HUBER; }
  double c() const override { return _c; }

private:
  double sgn(double r) const { return _c * r > 0.0 ? 1.0 : -1.0; }
};
class LossTDistribution : public Loss
{
public:
  LossTDistribution(Scaler::ShPtr scaler = std::make_shared<MedianScaler>(), double v = 5.0) : Loss(std::move(scaler)), _v(v) {}
  double computeWeight(double r) const override { return (_v + 1.0) / (_v + r * r); }
  double computeDerivative(double) const override { return 0.0; }  // left undefined by the reference (Loss.cpp:93-100)
  double compute(double) const override { return 0.0; }
  int kind() const override { return VSLAMB200_LOSS_TDIST; }
  double v() const override { return _v; }

private:
  const double _v;
};

// NormalEquations.h:26-59, NormalEquations.cpp:19-87: A = J^T W J, b = J^T W r, nConstraints = rows.  chi2: the header
// documents r^T W r, and addConstraint accumulates that; the (J, r, w) constructor's `(r * w).transpose() * r`
// (NormalEquations.cpp:58) is a product of two column vectors, which the reference's Release build evaluates as
// w[0] * sum r_k^2 (see VSLAMB200_CHI2_REFERENCE in vslam_b200.h) - reproduced here, as on the device.
class NormalEquations
{
public:
  typedef std::shared_ptr<NormalEquations> ShPtr;
  typedef std::unique_ptr<NormalEquations> UnPtr;
  typedef std::shared_ptr<const NormalEquations> ConstShPtr;
  typedef std::unique_ptr<const NormalEquations> ConstUnPtr;

  explicit NormalEquations(size_t nParameters) : _A(nParameters, nParameters), _b(nParameters, 0.0) {}
  explicit NormalEquations(const std::vector<NormalEquations> & parts) : NormalEquations(parts.at(0))
  {
    for (size_t i = 1; i < parts.size(); i++) combine(parts[i]);
  }
  explicit NormalEquations(const std::vector<NormalEquations::ConstShPtr> & parts) : NormalEquations(*parts.at(0))
  {
    for (size_t i = 1; i < parts.size(); i++) combine(*parts[i]);
  }
  // J: N x P, r and w: N
  NormalEquations(const MatXd & J, const VecXd & r, const VecXd & w) : NormalEquations(J.cols())
  {
    for (size_t k = 0; k < J.rows(); k++) accumulate(J.row(k), r[k], w[k]);
    _chi2 = 0.0;
    for (size_t k = 0; k < r.size(); k++) _chi2 += (r[k] * w[0]) * r[k];
    _nConstraints = r.size();
  }
  void addConstraint(const VecXd & J, double r, double w)
  {
    accumulate(J, r, w);
    _nConstraints++;
  }
  void combine(const NormalEquations & that)
  {
    for (size_t i = 0; i < _b.size(); i++) {
      for (size_t j = 0; j < _b.size(); j++) _A(i, j) += that._A(i, j);
      _b[i] += that._b[i];
    }
    _chi2 += that._chi2;
    _nConstraints += that._nConstraints;
  }
  std::string toString() const
  {
    std::stringstream ss;
    ss << "A=\n";
    for (size_t i = 0; i < _A.rows(); i++) {
      for (size_t j = 0; j < _A.cols(); j++) ss << (j ? " " : "") << _A(i, j);
      ss << "\n";
    }
    ss << "b=\n";
    for (size_t i = 0; i < _b.size(); i++) ss << (i ? " " : "") << _b[i];
    ss << "\nnConstraints=" << _nConstraints << " chi2=" << _chi2;
    return ss.str();
  }
  MatXd & A() { return _A; }
  VecXd & b() { return _b; }
  double & chi2() { return _chi2; }
  size_t & nConstraints() { return _nConstraints; }
  const MatXd & A() const { return _A; }
  const VecXd & b() const { return _b; }
  const double & chi2() const { return _chi2; }
  const size_t & nConstraints() const { return _nConstraints; }

private:
  void accumulate(const VecXd & J, double r, double w)
  {
    for (size_t i = 0; i < _b.size(); i++) {
      for (size_t j = 0; j < _b.size(); j++) _A(i, j) += J[i] * J[j] * w;
      _b[i] += J[i] * r * w;
    }
    _chi2 += r * r * w;
  }
  MatXd _A;
  VecXd _b;
  double _chi2 = 0.0;
  size_t _nConstraints = 0;
};

// Prior.h:25-33
class Prior
{
public:
  typedef std::shared_ptr<Prior> ShPtr;
  typedef std::shared_ptr<const Prior> ConstShPtr;
  virtual ~Prior() = default;
  virtual void apply(NormalEquations::ShPtr ne, const VecXd & x) const = 0;
};

// Solver.h:22-41
class Problem
{
public:
  typedef std::shared_ptr<Problem> ShPtr;
  typedef std::shared_ptr<const Problem> ConstShPtr;
  explicit Problem(size_t nParameters) : _nParameters(nParameters) {}
  virtual ~Problem() = default;
  size_t nParameters() const { return _nParameters; }
  virtual void updateX(const VecXd & dx) = 0;
  virtual void setX(const VecXd & x) = 0;
  virtual VecXd x() const = 0;
  virtual NormalEquations::ConstShPtr computeNormalEquations() = 0;

private:
  size_t _nParameters;
};

namespace detail
{
// determinant by LU with partial pivoting (the role of Eigen's MatrixXd::determinant, GaussNewton.cpp:52)
inline double determinant(MatXd m)
{
  const size_t n = m.rows();
  double det = 1.0;
  for (size_t c = 0; c < n; c++) {
    size_t p = c;
    for (size_t r = c + 1; r < n; r++)
      if (std::fabs(m(r, c)) > std::fabs(m(p, c))) p = r;
    if (m(p, c) == 0.0) return 0.0;
    if (p != c) {
      for (size_t k = 0; k < n; k++) std::swap(m(p, k), m(c, k));
      det = -det;
    }
    det *= m(c, c);
    for (size_t r = c + 1; r < n; r++) {
      const double f = m(r, c) / m(c, c);
      for (size_t k = c + 1; k < n; k++) m(r, k) -= f * m(c, k);
    }
  }
  return det;
}
// A = L D L^T, then forward / diagonal / backward substitution (the role of A.ldlt().solve(b), GaussNewton.cpp:76)
inline VecXd ldltSolve(const MatXd & A, const VecXd & b)
{
  const size_t n = b.size();
  MatXd L = MatXd::Identity(n);
  VecXd D(n), y(n), x(n);
  for (size_t j = 0; j < n; j++) {
    double d = A(j, j);
    for (size_t k = 0; k < j; k++) d -= L(j, k) * L(j, k) * D[k];
    D[j] = d;
    for (size_t i = j + 1; i < n; i++) {
      double s = A(i, j);
      for (size_t k = 0; k < j; k++) s -= L(i, k) * L(j, k) * D[k];
      L(i, j) = s / d;
    }
  }
  for (size_t i = 0; i < n; i++) {
    double s = b[i];
    for (size_t k = 0; k < i; k++) s -= L(i, k) * y[k];
    y[i] = s;
  }
  for (size_t i = n; i-- > 0;) {
    double s = y[i] / D[i];
    for (size_t k = i + 1; k < n; k++) s -= L(k, i) * x[k];
    x[i] = s;
  }
  return x;
}
}  // namespace detail

// Solver.h:43-63 / GaussNewton.h:35.  solve() runs the Gauss-Newton iteration on the host for ANY Problem (the role of
// GaussNewton.cpp:33-102); SE3Alignment::align does not go through it - it hands the solver's parameters to the device,
// where the same loop runs inside the level kernels (csrc/solve.cuh) for every job of a batch.
class Solver
{
public:
  typedef std::shared_ptr<Solver> ShPtr;
  typedef std::shared_ptr<const Solver> ConstShPtr;
  struct Results
  {
    typedef std::shared_ptr<Results> ShPtr;
    typedef std::unique_ptr<Results> UnPtr;
    typedef std::shared_ptr<const Results> ConstShPtr;
    typedef std::unique_ptr<const Results> ConstUnPtr;
    VecXd chi2, stepSize;
    MatXd x;
    std::vector<NormalEquations::ConstShPtr> normalEquations;
    size_t iteration = 0;  // (left uninitialised by the reference, Solver.h:59)
    // additions: why the loop ended (VSLAMB200_STOP_*) and how many steps were completed (normalEquations.size())
    int stop = VSLAMB200_STOP_MAXITER;
    size_t nSteps() const { return normalEquations.size(); }
  };
  virtual ~Solver() = default;
  virtual Results::ConstUnPtr solve(std::shared_ptr<Problem> problem) = 0;
  virtual double minStepSize() const = 0;
  virtual size_t maxIterations() const = 0;
  virtual int kind() const { return VSLAMB200_SOLVER_GN; }
  virtual double lambda0() const { return 1e-3; }
  virtual double up() const { return 10.0; }
  virtual double down() const { return 0.1; }
};
class GaussNewton : public Solver
{
public:
  GaussNewton(double minStepSize, size_t maxIterations)
  : _minStepSize(minStepSize), _minGradient(minStepSize), _minReduction(minStepSize), _maxIterations(maxIterations)
  {
  }
  double minStepSize() const override { return _minStepSize; }
  size_t maxIterations() const override { return _maxIterations; }

  Results::ConstUnPtr solve(std::shared_ptr<Problem> problem) override
  {
    auto res = std::make_unique<Results>();
    const size_t nP = problem->nParameters();
    res->chi2.assign(_maxIterations, 0.0);
    res->stepSize.assign(_maxIterations, 0.0);
    res->x = MatXd::Zero(_maxIterations, nP);
    res->normalEquations.reserve(_maxIterations);
    for (size_t it = 0; it < _maxIterations; it++) {
      const NormalEquations::ConstShPtr ne = problem->computeNormalEquations();
      const double det = detail::determinant(ne->A());
      if (ne->nConstraints() < nP) return stopped(res, VSLAMB200_STOP_CONSTRAINTS);
      if (!std::isfinite(det) || std::fabs(det) < 1e-6) return stopped(res, VSLAMB200_STOP_HESSIAN);
      res->chi2[it] = ne->chi2();
      const double dChi2 = it > 0 ? res->chi2[it] - res->chi2[it - 1] : 0.0;
      if (it > 0 && dChi2 > 0) {  // no improvement: x of the previous step is set again, the step is not taken
        problem->setX(res->x.row(it - 1));
        return stopped(res, VSLAMB200_STOP_CHI2);
      }
      const VecXd dx = detail::ldltSolve(ne->A(), ne->b());
      problem->updateX(dx);
      res->x.setRow(it, problem->x());
      double n2 = 0.0;
      for (double v : dx) n2 += v * v;
      res->stepSize[it] = std::sqrt(n2);
      res->normalEquations.push_back(ne);
      res->iteration = it;
      const double bMax = *std::max_element(ne->b().begin(), ne->b().end());  // signed maximum, then its magnitude
      if (it > 0 && (res->stepSize[it] < _minStepSize || std::fabs(bMax) < _minGradient || std::fabs(dChi2) < _minReduction))
        return stopped(res, VSLAMB200_STOP_CONVERGED);
      if (!std::isfinite(res->stepSize[it])) {
        if (it > 0) problem->setX(res->x.row(it - 1));  // (the reference reads row -1 when this happens at once)
        return stopped(res, VSLAMB200_STOP_NAN);
      }
    }
    return stopped(res, VSLAMB200_STOP_MAXITER);
  }

private:
  static Results::ConstUnPtr stopped(std::unique_ptr<Results> & r, int why)
  {
    r->stop = why;
    return Results::ConstUnPtr(r.release());
  }
  double _minStepSize, _minGradient, _minReduction;
  size_t _maxIterations;
};
// EXTENSION (no counterpart in the reference): Levenberg-Marquardt damping on the same normal equations,
// (A + lambda diag A) dx = b; lambda *= up after a rejected step, *= down after an accepted one.
class LevenbergMarquardt : public GaussNewton
{
public:
  LevenbergMarquardt(double minStepSize, size_t maxIterations, double lambda0 = 1e-3, double up = 10.0, double down = 0.1)
  : GaussNewton(minStepSize, maxIterations), _lambda0(lambda0), _up(up), _down(down)
  {
  }
  int kind() const override { return VSLAMB200_SOLVER_LM; }
  Results::ConstUnPtr solve(std::shared_ptr<Problem>) override
  {
    throw Exception("LevenbergMarquardt::solve: the damped loop runs on the device only (SE3Alignment::align)");
  }
  double lambda0() const override { return _lambda0; }
  double up() const override { return _up; }
  double down() const override { return _down; }

private:
  double _lambda0, _up, _down;
};
}  // namespace least_squares

// AlignmentSE3.h:22-26
class AlignmentSE3
{
public:
  virtual ~AlignmentSE3() = default;
  virtual PoseWithCovariance::UnPtr align(Frame::ConstShPtr from, Frame::ConstShPtr to) const = 0;
};

// SE3Alignment.h:25-46.  One device context per aligner; thread-safe through one mutex (the reference's align is
// const and called from a single callback thread).
class SE3Alignment : public AlignmentSE3
{
public:
  typedef std::shared_ptr<SE3Alignment> ShPtr;
  typedef std::unique_ptr<SE3Alignment> UnPtr;
  struct Pair
  {
    std::vector<Frame::ConstShPtr> from;
    Frame::ConstShPtr to;
    bool multi = false;
  };

  SE3Alignment(double minGradient, least_squares::Solver::ShPtr solver, least_squares::Loss::ShPtr loss,
               bool includePrior = false, int device = 0, int sampling = VSLAMB200_SAMPLING_NEAREST,
               double depthWeight = 0.0,  // these two: extension (bilinear sampling, depth residual), see vslam_b200.h
               int chi2Mode = VSLAMB200_CHI2_REFERENCE)  // NormalEquations.cpp:58 as built, or r^T W r (vslam_b200.h)
  : _solver(std::move(solver)), _loss(std::move(loss))
  {
    if (!_solver) throw Exception("SE3Alignment: solver is null");
    if (_loss && !_loss->scaler().deviceCompatible())
      throw Exception("SE3Alignment: the device scalers run ScalerTDistribution's defaults (30 iterations, 1e-5)");
    vslamb200_config cfg{};
    cfg.min_gradient = minGradient;
    cfg.loss_kind = _loss ? _loss->kind() : VSLAMB200_LOSS_NONE;  // loss == nullptr: InverseCompositional.cpp:129
    cfg.huber_c = _loss ? _loss->c() : 1.345;
    cfg.loss_v = _loss ? _loss->v() : 5.0;
    cfg.scaler_kind = _loss ? _loss->scaler().kind() : VSLAMB200_SCALER_IDENTITY;
    cfg.scaler_v = _loss ? _loss->scaler().v() : 5.0;
    cfg.min_step_size = _solver->minStepSize();
    cfg.max_iterations = (int)_solver->maxIterations();
    cfg.include_prior = includePrior ? 1 : 0;
    cfg.solver_kind = _solver->kind();
    cfg.lm_lambda0 = _solver->lambda0();
    cfg.lm_up = _solver->up();
    cfg.lm_down = _solver->down();
    cfg.sampling = sampling;
    cfg.depth_weight = depthWeight;
    cfg.chi2_mode = chi2Mode;
    check(vslamb200_create(&cfg, device, &_ctx));
  }
  ~SE3Alignment() override { vslamb200_destroy(_ctx); }
  SE3Alignment(const SE3Alignment &) = delete;
  SE3Alignment & operator=(const SE3Alignment &) = delete;

  PoseWithCovariance::UnPtr align(Frame::ConstShPtr from, Frame::ConstShPtr to) const override
  {
    Pair p;
    p.from = {std::move(from)};
    p.to = std::move(to);
    return std::move(align(std::vector<Pair>{p})[0]);
  }
  PoseWithCovariance::UnPtr align(const std::vector<Frame::ConstShPtr> & from, Frame::ConstShPtr to) const
  {
    Pair p;
    p.from = from;
    p.to = std::move(to);
    p.multi = true;
    return std::move(align(std::vector<Pair>{p})[0]);
  }
  // Addition: many independent alignments in one batched device call (one cluster per alignment).
  std::vector<PoseWithCovariance::UnPtr> align(const std::vector<Pair> & pairs) const
  {
    std::lock_guard<std::mutex> lock(_mutex);
    purge();
    std::vector<vslamb200_job> jobs(pairs.size());
    std::vector<std::vector<std::uint64_t>> ids(pairs.size());
    std::vector<std::vector<double>> poses(pairs.size());
    for (size_t i = 0; i < pairs.size(); i++) {
      const Pair & p = pairs[i];
      if (!p.to || p.from.empty()) throw Exception("SE3Alignment::align: null frame");
      ensureResident(p.to);
      for (const auto & f : p.from) {
        if (!f) throw Exception("SE3Alignment::align: null frame");
        ensureResident(f);
        ids[i].push_back(f->id());
        poses[i].insert(poses[i].end(), f->pose().pose().data(), f->pose().pose().data() + 7);
      }
      vslamb200_job & j = jobs[i];
      j.to = p.to->id();
      for (int k = 0; k < 7; k++) j.to_pose[k] = p.to->pose().pose().data()[k];
      j.n_from = (int)p.from.size();
      j.multi = p.multi ? 1 : 0;
      j.from = ids[i].data();
      j.from_poses = poses[i].data();
    }
    std::vector<vslamb200_result> res(pairs.size());
    check(vslamb200_align(_ctx, jobs.data(), (int)jobs.size(), res.data()));
    std::vector<PoseWithCovariance::UnPtr> out;
    for (const auto & r : res) {
      SE3d pose;
      Matd6 cov;
      for (int k = 0; k < 7; k++) pose.data()[k] = r.pose[k];
      for (int k = 0; k < 36; k++) cov[k] = r.cov[k];
      out.push_back(std::make_unique<PoseWithCovariance>(pose, cov));
    }
    return out;
  }
  vslamb200_ctx * context() const { return _ctx; }
  // Makes a frame resident (pyramid, gradients) without aligning it; Frame's per-level accessors work afterwards.
  void upload(const Frame::ConstShPtr & f) const
  {
    if (!f) throw Exception("SE3Alignment::upload: null frame");
    std::lock_guard<std::mutex> lock(_mutex);
    ensureResident(f);
  }
  // One InverseCompositional::computeNormalEquations (InverseCompositional.cpp:105-154) of keyframe `from` against `to` on
  // `level` at the relative transform se3 = poseCur * poseRef^-1, on the device, WITHOUT the motion prior: A, b, chi2 are
  // divided by nConstraints as the reference does (:139-143).  multiMask: the interest mask of the vector overload.
  least_squares::NormalEquations::ShPtr evaluate(const Frame::ConstShPtr & from, const Frame::ConstShPtr & to, size_t level,
                                                 const SE3d & se3, bool multiMask = false) const
  {
    if (!from || !to) throw Exception("SE3Alignment::evaluate: null frame");
    std::lock_guard<std::mutex> lock(_mutex);
    ensureResident(from);
    ensureResident(to);
    double ne64[64];
    check(vslamb200_evaluate(_ctx, from->id(), to->id(), (int)level, se3.data(), multiMask ? 1 : 0, nullptr, ne64));
    auto ne = std::make_shared<least_squares::NormalEquations>(6);
    ne->nConstraints() = (size_t)ne64[0];
    ne->chi2() = ne64[1];
    for (int i = 0; i < 6; i++) {
      ne->b()[i] = ne64[4 + i];
      for (int j = 0; j < 6; j++) ne->A()(i, j) = ne64[10 + 6 * i + j];
    }
    return ne;
  }

private:
  struct Resident
  {
    std::weak_ptr<const Frame> frame;
    std::uint64_t version;
  };
  void ensureResident(const Frame::ConstShPtr & f) const
  {
    auto it = _resident.find(f->id());
    if (it != _resident.end() && it->second.version == f->version() && !it->second.frame.expired()) return;
    const std::uint64_t id = f->id();
    const double cam[4] = {f->camera()->fx(), f->camera()->fy(), f->camera()->cx(), f->camera()->cy()};
    const Image & I = f->intensity();
    const void * depth = f->depthIsFloat() ? (const void *)f->depthF().data.data() : (const void *)f->depth().data.data();
    check(vslamb200_frames_upload(
      _ctx, 1, &id, I.cols, I.rows, I.data.data(), 0, depth, f->depthIsFloat() ? VSLAMB200_DEPTH_F32 : VSLAMB200_DEPTH_F64,
      0, I.colMajor ? VSLAMB200_COL_MAJOR : VSLAMB200_ROW_MAJOR, cam, (int)f->nLevels()));
    _resident[id] = Resident{f, f->version()};
    f->attach(_ctx);
  }
  void purge() const  // frames that died since the last call leave the device
  {
    for (auto it = _resident.begin(); it != _resident.end();) {
      if (it->second.frame.expired()) {
        vslamb200_frame_release(_ctx, it->first);
        it = _resident.erase(it);
      } else {
        ++it;
      }
    }
  }

  least_squares::Solver::ShPtr _solver;
  least_squares::Loss::ShPtr _loss;
  vslamb200_ctx * _ctx = nullptr;
  mutable std::mutex _mutex;
  mutable std::unordered_map<std::uint64_t, Resident> _resident;
};

// MotionPrior (SE3Alignment.cpp:27-52): information I, prediction x_pred = log(predicted * reference^-1) fixed at
// construction; apply: A <- A / 255^2 + I, b <- b / 255^2 + (x_pred - x).
class MotionPrior : public least_squares::Prior
{
public:
  MotionPrior(const PoseWithCovariance & predictedPose, const PoseWithCovariance & referencePose)
  : _xPred(se3::log(se3::mul(predictedPose.pose(), se3::inverse(referencePose.pose()))))
  {
  }
  void apply(least_squares::NormalEquations::ShPtr ne, const least_squares::VecXd & x) const override
  {
    const double norm = 255.0 * 255.0;
    for (size_t i = 0; i < 6; i++) {
      for (size_t j = 0; j < 6; j++) ne->A()(i, j) = ne->A()(i, j) / norm + (i == j ? 1.0 : 0.0);
      ne->b()[i] = ne->b()[i] / norm + (_xPred[i] - x[i]);
    }
  }

private:
  Vec6d _xPred;
};

namespace lukas_kanade
{
// The least-squares problem SE3Alignment solves on one pyramid level (InverseCompositional.h:30-88 with WarpSE3,
// Warp.cpp:135-157,233-239), with the normal equations evaluated on the device: a least_squares::Solver can drive the
// device path iteration by iteration - solver->solve(problem) reproduces the iterates of SE3Alignment::align on that
// level.  x is log(se3), se3 = poseCur * poseRef^-1; updateX(dx): se3 <- se3 * exp(-dx) (InverseCompositional.cpp:104).
class InverseCompositional : public least_squares::Problem
{
public:
  InverseCompositional(const SE3Alignment & aligner, Frame::ConstShPtr from, Frame::ConstShPtr to, size_t level,
                       const SE3d & se3 = SE3d(), least_squares::Prior::ConstShPtr prior = nullptr, bool multiMask = false)
  : least_squares::Problem(6), _aligner(aligner), _from(std::move(from)), _to(std::move(to)), _level(level), _se3(se3),
    _prior(std::move(prior)), _multiMask(multiMask)
  {
  }
  void updateX(const least_squares::VecXd & dx) override
  {
    Vec6d m;
    for (int i = 0; i < 6; i++) m[i] = -dx[i];
    _se3 = se3::mul(_se3, se3::exp(m));
  }
  void setX(const least_squares::VecXd & x) override
  {
    Vec6d v;
    for (int i = 0; i < 6; i++) v[i] = x[i];
    _se3 = se3::exp(v);
  }
  least_squares::VecXd x() const override
  {
    const Vec6d v = se3::log(_se3);
    return least_squares::VecXd(v.begin(), v.end());
  }
  const SE3d & se3() const { return _se3; }
  least_squares::NormalEquations::ConstShPtr computeNormalEquations() override
  {
    auto ne = _aligner.evaluate(_from, _to, _level, _se3, _multiMask);
    if (_prior) _prior->apply(ne, x());
    return ne;
  }

private:
  const SE3Alignment & _aligner;
  Frame::ConstShPtr _from, _to;
  size_t _level;
  SE3d _se3;
  least_squares::Prior::ConstShPtr _prior;
  bool _multiMask;
};
}  // namespace lukas_kanade
}  // namespace vslam_b200

#endif  // VSLAM_B200_HPP_
