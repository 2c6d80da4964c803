// vslam_b200.hpp — C++17 host mirror of the reference's aligner interface, header-only, on top of the C ABI
// (vslam_b200.h).  Same class and method names, argument meaning and error behaviour as the reference for this
// path, without Eigen / Sophus / OpenCV (paths relative to /root/reference/src/vslam/src/):
//
//   pd::vslam::Camera                     core/src/Camera.h                  -> vslam_b200::Camera
//   pd::vslam::PoseWithCovariance         core/src/PoseWithCovariance.h:31-43 -> vslam_b200::PoseWithCovariance
//   pd::vslam::Frame                      core/src/Frame.h:35-37,43,94        -> vslam_b200::Frame
//   least_squares::Scaler / Loss family   least_squares/src/Scaler.h:30-74, Loss.h:37-116
//   least_squares::Solver / GaussNewton   least_squares/src/Solver.h:22-63, GaussNewton.h:35
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

#include <array>
#include <atomic>
#include <cmath>
#include <cstdint>
#include <memory>
#include <mutex>
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
  const Image & intensity() const { return _intensity; }
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
  std::uint64_t _id;
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
// Scaler.h:30-74
class Scaler
{
public:
  typedef std::shared_ptr<Scaler> ShPtr;
  virtual ~Scaler() = default;
  virtual int kind() const { return VSLAMB200_SCALER_IDENTITY; }
  virtual double v() const { return 5.0; }
};
class MedianScaler : public Scaler
{
public:
  int kind() const override { return VSLAMB200_SCALER_MEDIAN; }
};
class MeanScaler : public Scaler
{
public:
  int kind() const override { return VSLAMB200_SCALER_MEAN; }
};
class ScalerTDistribution : public Scaler
{
public:
  explicit ScalerTDistribution(double v = 5.0) : _v(v) {}
  int kind() const override { return VSLAMB200_SCALER_TDIST; }
  double v() const override { return _v; }

private:
  double _v;
};

// Loss.h:37-116
class Loss
{
public:
  typedef std::shared_ptr<Loss> ShPtr;
  explicit Loss(Scaler::ShPtr scaler = std::make_shared<Scaler>()) : _scaler(std::move(scaler)) {}
  virtual ~Loss() = default;
  virtual int kind() const = 0;
  virtual double c() const { return 1.345; }
  virtual double v() const { return 5.0; }
  const Scaler & scaler() const { return *_scaler; }

protected:
  Scaler::ShPtr _scaler;
};
class QuadraticLoss : public Loss
{
public:
  using Loss::Loss;
  int kind() const override { return VSLAMB200_LOSS_QUADRATIC; }
};
class TukeyLoss : public Loss
{
public:
  explicit TukeyLoss(Scaler::ShPtr scaler = std::make_shared<MedianScaler>()) : Loss(std::move(scaler)) {}
  int kind() const override { return VSLAMB200_LOSS_TUKEY; }
};
class HuberLoss : public Loss
{
public:
  explicit HuberLoss(double c, Scaler::ShPtr scaler = std::make_shared<MedianScaler>()) : Loss(std::move(scaler)), _c(c) {}
  int kind() const override { return VSLAMB200_LOSS_HUBER; }
  double c() const override { return _c; }

private:
  double _c;
};
class LossTDistribution : public Loss
{
public:
  explicit LossTDistribution(Scaler::ShPtr scaler = std::make_shared<ScalerTDistribution>(), double v = 5.0)
  : Loss(std::move(scaler)), _v(v) {}
  int kind() const override { return VSLAMB200_LOSS_TDIST; }
  double v() const override { return _v; }

private:
  double _v;
};

// Solver.h:22-63 / GaussNewton.h:35.  The iteration loop itself runs on the device (gn_step_block); this object carries
// its parameters.
class Solver
{
public:
  typedef std::shared_ptr<Solver> ShPtr;
  virtual ~Solver() = default;
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
  GaussNewton(double minStepSize, size_t maxIterations) : _minStepSize(minStepSize), _maxIterations(maxIterations) {}
  double minStepSize() const override { return _minStepSize; }
  size_t maxIterations() const override { return _maxIterations; }

private:
  double _minStepSize;
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
               double depthWeight = 0.0)  // the last two: extension (bilinear sampling, depth residual), see vslam_b200.h
  : _solver(std::move(solver)), _loss(std::move(loss))
  {
    if (!_solver) throw Exception("SE3Alignment: solver is null");
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
}  // namespace vslam_b200

#endif  // VSLAM_B200_HPP_
