// C entry points around the REFERENCE's own aligner, compiled from /root/reference by oracle/build_ref.py into
// oracle/_ref/libvslam_ref.so.  Test infrastructure: it exists so that tests/test_oracle_ref_cpu.py can pin the oracle's
// restatement (oracle/vslam_oracle.cpp) against the reference's control flow and arithmetic.  Nothing here restates the
// algorithm: frames go through the reference's Frame (computePyramid / computeDerivatives / computePcl, the order of
// NodeRgbdAlignment.cpp), the alignment through SE3Alignment::align (SE3Alignment.cpp:63 and :107) with the
// reference's GaussNewton, losses and scalers.  The solver handed to SE3Alignment is the reference's GaussNewton inside
// a wrapper that only copies each level's Solver::Results out.
#include <cmath>
#include <cstdint>
#include <cstring>
#include <memory>
#include <vector>

#include "KeyFrameSelection.h"
#include "MotionPrediction.h"
#include "Odometry.h"
#include "SE3Alignment.h"
#include "core/core.h"
#include "least_squares/least_squares.h"

using namespace pd::vslam;
namespace ls = pd::vslam::least_squares;

// The two algorithm.cpp functions named by InverseCompositional's image-only constructor (not on the alignment path;
// the rest of algorithm.cpp needs Eigen's geometry and eigenvalue modules and is not built).
namespace pd::vslam::algorithm
{
Eigen::MatrixXi gradX(const Image &) { throw pd::Exception("ref_harness: algorithm::gradX is not built"); }
Eigen::MatrixXi gradY(const Image &) { throw pd::Exception("ref_harness: algorithm::gradY is not built"); }
}  // namespace pd::vslam::algorithm

// Declared by headers the odometry sources include, never run here: the Kalman motion model (kalman/src needs more of
// Eigen than the stand-in has) and the PCL-based ICP aligner.  Odometry.cpp / MotionPrediction.cpp reference them.
namespace pd::vslam::kalman
{
EKFConstantVelocitySE3::State::UnPtr EKFConstantVelocitySE3::predict(Timestamp) const
{
  throw pd::Exception("ref_harness: the Kalman motion model is not built");
}
void EKFConstantVelocitySE3::update(const Vec6d &, const Matd<6, 6> &, Timestamp)
{
  throw pd::Exception("ref_harness: the Kalman motion model is not built");
}
}  // namespace pd::vslam::kalman
namespace pd::vslam
{
PoseWithCovariance::UnPtr IterativeClosestPoint::align(Frame::ConstShPtr, Frame::ConstShPtr) const
{
  throw pd::Exception("ref_harness: the ICP aligner is not built");
}
}  // namespace pd::vslam

namespace
{
constexpr int TRACE_STRIDE = 64;  // the row layout of oracle/vslam_oracle.cpp's trace

struct LevelRecord
{
  std::vector<double> rows;  // one TRACE_STRIDE row per pushed normal equation
  int iteration = -1;
  int nPushed = 0;
};

class RecordingSolver : public ls::Solver
{
public:
  RecordingSolver(double minStepSize, size_t maxIterations) : _inner(std::make_shared<ls::GaussNewton>(minStepSize, maxIterations)) {}
  Results::ConstUnPtr solve(std::shared_ptr<ls::Problem> problem) override
  {
    auto r = _inner->solve(problem);
    LevelRecord rec;
    rec.nPushed = (int)r->normalEquations.size();
    rec.iteration = rec.nPushed ? (int)r->iteration : -1;  // Results::iteration is not initialised before the first push
    rec.rows.assign((size_t)rec.nPushed * TRACE_STRIDE, std::nan(""));
    for (int i = 0; i < rec.nPushed; i++) {
      double * t = rec.rows.data() + (size_t)i * TRACE_STRIDE;
      const auto & ne = r->normalEquations[(size_t)i];
      t[0] = 1;
      t[1] = (double)ne->nConstraints();
      t[2] = r->chi2(i);
      t[3] = r->stepSize(i);
      for (int k = 0; k < 6; k++) t[6 + k] = r->x(i, k);
      for (int k = 0; k < 6; k++) t[12 + k] = ne->b()(k);
      for (int a = 0; a < 6; a++)
        for (int b = 0; b < 6; b++) t[18 + a * 6 + b] = ne->A()(a, b);
    }
    levels.push_back(rec);
    return r;
  }
  std::vector<LevelRecord> levels;  // in the order solve() was called: coarsest level first

private:
  std::shared_ptr<ls::GaussNewton> _inner;
};

ls::Scaler::ShPtr make_scaler(int kind, double v)
{
  switch (kind) {
    case 1: return std::make_shared<ls::MedianScaler>();
    case 2: return std::make_shared<ls::MeanScaler>();
    case 3: return std::make_shared<ls::ScalerTDistribution>(v);
    default: return std::make_shared<ls::Scaler>();
  }
}
ls::Loss::ShPtr make_loss(int kind, double huberC, double v, ls::Scaler::ShPtr scaler)
{
  switch (kind) {
    case 1: return std::make_shared<ls::QuadraticLoss>(scaler);
    case 2: return std::make_shared<ls::TukeyLoss>(scaler);
    case 3: return std::make_shared<ls::HuberLoss>(scaler, huberC);
    case 4: return std::make_shared<ls::LossTDistribution>(scaler, v);
    default: return nullptr;
  }
}

Frame::ShPtr make_frame(int W, int H, const uint8_t * gray, const double * depth, const double * cam4, const double * twist, int nLevels)
{
  Image I(H, W);
  DepthMap Z(H, W);
  for (int v = 0; v < H; v++)
    for (int u = 0; u < W; u++) {
      I(v, u) = gray[(size_t)v * W + u];
      Z(v, u) = depth[(size_t)v * W + u];
    }
  Vec6d x;
  for (int k = 0; k < 6; k++) x(k) = twist[k];
  auto cam = std::make_shared<Camera>(cam4[0], cam4[1], cam4[2], cam4[3]);
  auto f = std::make_shared<Frame>(I, Z, cam, 0U, PoseWithCovariance(x));
  f->computePyramid((size_t)nLevels);
  f->computeDerivatives();
  f->computePcl();
  return f;
}
}  // namespace

extern "C" {

struct ref_config  // the fields of oracle/vslam_oracle.cpp's vo_config that the reference has a counterpart for
{
  double min_gradient;
  int loss_kind;  // 0 none (null loss), 1 quadratic, 2 Tukey, 3 Huber, 4 t-distribution
  double huber_c, loss_v;
  int scaler_kind;  // 0 identity, 1 median, 2 mean, 3 t-distribution
  double scaler_v;
  double min_step_size;
  int max_iterations;
  int include_prior;
};

// Frames: nFrom keyframes then the frame to align, each W x H (gray u8, depth f64 in metres, 0 / NaN invalid), poses as
// twists (PoseWithCovariance(Vec6d)).  multi != 0 calls the vector overload (also for nFrom == 1).
// Out: twist6 = log of the aligned pose, Rt12 = rotation matrix row-major + translation, cov36, per level (index =
// pyramid level) pushed normal equations, last completed iteration, and trace rows [nLevels][max_iterations][64].
// Returns 0, or -1 with the exception text in err.
int ref_align(
  const ref_config * c, int nFrom, int W, int H, const uint8_t * gray, const double * depth, const double * cam4,
  const double * twists, int nLevels, int multi, double * twist6, double * Rt12, double * cov36, int * pushed, int * lastIteration,
  double * trace, char * err, int errLen)
{
  try {
    std::vector<Frame::ConstShPtr> from;
    const size_t npx = (size_t)W * H;
    for (int i = 0; i < nFrom; i++) from.push_back(make_frame(W, H, gray + i * npx, depth + i * npx, cam4, twists + 6 * i, nLevels));
    Frame::ConstShPtr to = make_frame(W, H, gray + nFrom * npx, depth + nFrom * npx, cam4, twists + 6 * nFrom, nLevels);
    auto solver = std::make_shared<RecordingSolver>(c->min_step_size, (size_t)c->max_iterations);
    auto loss = make_loss(c->loss_kind, c->huber_c, c->loss_v, make_scaler(c->scaler_kind, c->scaler_v));
    SE3Alignment aligner(c->min_gradient, solver, loss, c->include_prior != 0);
    PoseWithCovariance::UnPtr pose = multi ? aligner.align(from, to) : aligner.align(from[0], to);
    const Vec6d x = pose->pose().log();
    for (int k = 0; k < 6; k++) twist6[k] = x(k);
    const auto R = pose->pose().rotationMatrix();
    for (int a = 0; a < 3; a++)
      for (int b = 0; b < 3; b++) Rt12[a * 3 + b] = R(a, b);
    for (int a = 0; a < 3; a++) Rt12[9 + a] = pose->pose().translation()(a);
    const auto cov = pose->cov();
    for (int a = 0; a < 6; a++)
      for (int b = 0; b < 6; b++) cov36[a * 6 + b] = cov(a, b);
    for (size_t k = 0; k < solver->levels.size(); k++) {
      const int level = nLevels - 1 - (int)k;
      const LevelRecord & rec = solver->levels[k];
      pushed[level] = rec.nPushed;
      lastIteration[level] = rec.iteration;
      if (trace)
        std::memcpy(trace + (size_t)level * c->max_iterations * TRACE_STRIDE, rec.rows.data(), rec.rows.size() * sizeof(double));
    }
    return 0;
  } catch (const std::exception & e) {
    if (err && errLen > 0) {
      std::strncpy(err, e.what(), (size_t)errLen - 1);
      err[errLen - 1] = 0;
    }
    return -1;
  }
}

// One pyramid level of a frame as the reference's Frame builds it (for the pyramid / derivative / depth checks).
int ref_frame_level(
  int W, int H, const uint8_t * gray, const double * depth, const double * cam4, int nLevels, int level, int * dims4, uint8_t * I,
  double * Z, double * dIx, double * dIy, double * pcl, double * camOut4)
{
  try {
    const double zero[6] = {0, 0, 0, 0, 0, 0};
    auto f = make_frame(W, H, gray, depth, cam4, zero, nLevels);
    const int w = (int)f->width(level), h = (int)f->height(level);
    const auto & z = f->depth(level);
    dims4[0] = w;
    dims4[1] = h;
    dims4[2] = (int)z.cols();
    dims4[3] = (int)z.rows();
    if (!I) return 0;
    if (z.rows() != h || z.cols() != w) return -2;
    const auto pc = f->pcl(level, false);
    for (int v = 0; v < h; v++)
      for (int u = 0; u < w; u++) {
        const size_t o = (size_t)v * w + u;
        I[o] = f->intensity(level)(v, u);
        Z[o] = z(v, u);
        dIx[o] = f->dIx(level)(v, u);
        dIy[o] = f->dIy(level)(v, u);
        for (int k = 0; k < 3; k++) pcl[o * 3 + k] = pc[o](k);
      }
    camOut4[0] = f->camera(level)->fx();
    camOut4[1] = f->camera(level)->fy();
    camOut4[2] = f->camera(level)->principalPoint().x();
    camOut4[3] = f->camera(level)->principalPoint().y();
    return 0;
  } catch (const std::exception &) {
    return -1;
  }
}

// The aligner's direct caller, SURVEY.md 8(f)-2: the reference's OdometryRgbd (Odometry.cpp:21-62), MotionPrediction
// (MotionPrediction.cpp:22-55), KeyFrameSelectionIdx (KeyFrameSelection.h:36-51) and Map (Map.cpp:19-33), driven in the
// order of the ROS node's processFrame (src/ros/nodes/NodeRgbdAlignment.cpp:121-150 - the node itself is ROS and cannot
// be built; these six calls are its loop body).  motionModel: 0 NoMotion, 1 ConstantMotion.  stamps in nanoseconds.
// Out: per frame the pose (twist) the frame ends up with and the speed (twist per second).
int ref_odometry(
  const ref_config * c, int nFrames, int W, int H, const uint8_t * gray, const double * depth, const double * cam4,
  const uint64_t * stamps, int nLevels, int motionModel, int kfPeriod, double * poses6, double * speeds6, char * err, int errLen)
{
  try {
    auto solver = std::make_shared<ls::GaussNewton>(c->min_step_size, (size_t)c->max_iterations);
    auto loss = make_loss(c->loss_kind, c->huber_c, c->loss_v, make_scaler(c->scaler_kind, c->scaler_v));
    auto map = std::make_shared<Map>();
    auto odometry = std::make_shared<OdometryRgbd>(c->min_gradient, solver, loss, map);
    MotionPrediction::ShPtr prediction = MotionPrediction::make(motionModel ? "ConstantMotion" : "NoMotion");
    auto keyFrameSelection = std::make_shared<KeyFrameSelectionIdx>((uint64_t)kfPeriod);
    const size_t npx = (size_t)W * H;
    const double zero[6] = {0, 0, 0, 0, 0, 0};
    for (int i = 0; i < nFrames; i++) {
      Image I(H, W);
      DepthMap Z(H, W);
      for (int v = 0; v < H; v++)
        for (int u = 0; u < W; u++) {
          I(v, u) = gray[i * npx + (size_t)v * W + u];
          Z(v, u) = depth[i * npx + (size_t)v * W + u];
        }
      Vec6d x0;
      for (int k = 0; k < 6; k++) x0(k) = zero[k];
      auto cam = std::make_shared<Camera>(cam4[0], cam4[1], cam4[2], cam4[3]);
      auto frame = std::make_shared<Frame>(I, Z, cam, stamps[i], PoseWithCovariance(x0));
      frame->computePyramid((size_t)nLevels);  // createFrame (NodeRgbdAlignment.cpp:194-197)
      frame->computeDerivatives();
      frame->computePcl();
      frame->set(*prediction->predict(frame->t()));
      odometry->update(frame);
      frame->set(*odometry->pose());
      prediction->update(odometry->pose(), frame->t());
      keyFrameSelection->update(frame);
      map->insert(frame, keyFrameSelection->isKeyFrame());
      const Vec6d p = odometry->pose()->pose().log(), s = odometry->speed()->pose().log();
      for (int k = 0; k < 6; k++) {
        poses6[6 * i + k] = p(k);
        speeds6[6 * i + k] = s(k);
      }
    }
    return 0;
  } catch (const std::exception & e) {
    if (err && errLen > 0) {
      std::strncpy(err, e.what(), (size_t)errLen - 1);
      err[errLen - 1] = 0;
    }
    return -1;
  }
}
}
