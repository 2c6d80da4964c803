// vslam_b200_odometry.hpp — C++17 host mirror of the DIRECT CALLER of the alignment path,
// header-only, on top of vslam_b200.hpp.  Same class and method names, argument meaning and error behaviour as the
// reference (paths relative to /root/reference/src/vslam/src/):
//
//   Sophus::SE3d exp / log / * / inverse            (un-vendored dependency)      -> vslam_b200::se3::
//   algorithm::computeRelativeTransform             core/src/algorithm.cpp:82-85  -> vslam_b200::algorithm::
//   pd::vslam::Map (frame / key-frame rings)        odometry/src/mapping/Map.cpp:19-33, Map.h:41-48
//   pd::vslam::KeyFrameSelectionIdx                 odometry/src/KeyFrameSelection.h:36-51
//   pd::vslam::MotionPrediction{NoMotion,Constant}  odometry/src/MotionPrediction.h:41-76, MotionPrediction.cpp:22-55
//   pd::vslam::Odometry / OdometryRgbd              odometry/src/Odometry.h, Odometry.cpp:21-62
//   utils::loadTrajectory / writeTrajectory         utils/src/utils.cpp:76-117,134-163 (TUM trajectory files)
//
// The frames the odometry aligns against live in the Map's rings (7 + 7); SE3Alignment keeps exactly those resident
// on the device (a frame leaves the device when its last shared_ptr dies, vslam_b200.hpp).  Not mirrored: the Kalman
// motion model (kalman/, out of the path), KeyFrameSelectionCustom (needs the feature map), OdometryIcp.
#ifndef VSLAM_B200_ODOMETRY_HPP_
#define VSLAM_B200_ODOMETRY_HPP_

#include <deque>
#include <fstream>
#include <map>
#include <sstream>

#include "vslam_b200.hpp"

namespace vslam_b200
{
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

namespace algorithm
{
// core/src/algorithm.cpp:82-85
inline SE3d computeRelativeTransform(const SE3d & t0, const SE3d & t1) { return se3::mul(t1, se3::inverse(t0)); }
}  // namespace algorithm

// ------------------------------------------------------------------------------------------------
// Map: rings of the last frames and key frames, newest first (Map.cpp:19-33)
// ------------------------------------------------------------------------------------------------
class Map
{
public:
  typedef std::shared_ptr<Map> ShPtr;
  typedef std::shared_ptr<const Map> ConstShPtr;
  Map() : _maxFrames(7), _maxKeyFrames(7) {}
  void insert(Frame::ShPtr frame, bool isKeyFrame)
  {
    if (_frames.size() >= _maxFrames) _frames.pop_back();
    _frames.push_front(frame);
    if (isKeyFrame) {
      if (_keyFrames.size() >= _maxKeyFrames) _keyFrames.pop_back();
      _keyFrames.push_front(frame);
    }
  }
  Frame::ConstShPtr lastKf(size_t idx = 0) const { return _keyFrames.size() <= idx ? nullptr : _keyFrames.at(idx); }
  Frame::ConstShPtr lastFrame(size_t idx = 0) const { return _frames.size() <= idx ? nullptr : _frames.at(idx); }
  std::vector<Frame::ConstShPtr> frames() const { return {_frames.begin(), _frames.end()}; }
  std::vector<Frame::ConstShPtr> keyFrames() const { return {_keyFrames.begin(), _keyFrames.end()}; }
  // Map.cpp:87-102
  void updatePose(std::uint64_t id, const PoseWithCovariance & pose)
  {
    for (auto & f : _keyFrames)
      if (f->id() == id) {
        f->set(pose);
        return;
      }
    for (auto & f : _frames)
      if (f->id() == id) {
        f->set(pose);
        return;
      }
    throw Exception("Frame not part of map: " + std::to_string(id));
  }

private:
  std::deque<Frame::ShPtr> _frames, _keyFrames;
  const size_t _maxFrames, _maxKeyFrames;
};

// ------------------------------------------------------------------------------------------------
// key-frame selection (KeyFrameSelection.h:22-51)
// ------------------------------------------------------------------------------------------------
class KeyFrameSelection
{
public:
  typedef std::shared_ptr<KeyFrameSelection> ShPtr;
  virtual ~KeyFrameSelection() = default;
  virtual void update(Frame::ConstShPtr frame) = 0;
  virtual bool isKeyFrame() const = 0;
};
class KeyFrameSelectionIdx : public KeyFrameSelection
{
public:
  explicit KeyFrameSelectionIdx(std::uint64_t period = 2) : _period(period), _ctr(0U) {}
  void update(Frame::ConstShPtr) override { _ctr++; }
  bool isKeyFrame() const override { return _ctr == 0 || _ctr % _period == 0; }

private:
  const std::uint64_t _period;
  std::uint64_t _ctr;
};

// ------------------------------------------------------------------------------------------------
// motion prediction (MotionPrediction.h:24-76, MotionPrediction.cpp:22-55)
// ------------------------------------------------------------------------------------------------
class MotionPrediction
{
public:
  typedef std::shared_ptr<MotionPrediction> ShPtr;
  virtual ~MotionPrediction() = default;
  virtual void update(PoseWithCovariance::ConstShPtr pose, Timestamp timestamp) = 0;
  virtual PoseWithCovariance::UnPtr predict(Timestamp timestamp) const = 0;
  static ShPtr make(const std::string & model);
};
class MotionPredictionNoMotion : public MotionPrediction
{
public:
  MotionPredictionNoMotion() : _lastPose(std::make_shared<PoseWithCovariance>()) {}
  void update(PoseWithCovariance::ConstShPtr pose, Timestamp) override { _lastPose = pose; }
  PoseWithCovariance::UnPtr predict(Timestamp) const override
  {
    return std::make_unique<PoseWithCovariance>(_lastPose->pose(), _lastPose->cov());
  }

private:
  PoseWithCovariance::ConstShPtr _lastPose;
};
class MotionPredictionConstant : public MotionPrediction
{
public:
  MotionPredictionConstant() : _lastPose(std::make_shared<PoseWithCovariance>()), _lastT(0U) { _speed.fill(0.0); }
  void update(PoseWithCovariance::ConstShPtr pose, Timestamp timestamp) override
  {
    if (timestamp < _lastT) throw Exception("New timestamp is older than last one!");
    const double dT = (static_cast<double>(timestamp) - static_cast<double>(_lastT)) / 1e9;
    _speed = se3::log(algorithm::computeRelativeTransform(_lastPose->pose(), pose->pose()));
    for (double & v : _speed) v /= dT;
    _lastPose = pose;
    _lastT = timestamp;
  }
  PoseWithCovariance::UnPtr predict(Timestamp timestamp) const override
  {
    const double dT = (static_cast<double>(timestamp) - static_cast<double>(_lastT)) / 1e9;
    Vec6d x = _speed;
    for (double & v : x) v *= dT;
    return std::make_unique<PoseWithCovariance>(se3::mul(se3::exp(x), _lastPose->pose()), PoseWithCovariance().cov());
  }

private:
  Vec6d _speed;
  PoseWithCovariance::ConstShPtr _lastPose;
  Timestamp _lastT;
};
inline MotionPrediction::ShPtr MotionPrediction::make(const std::string & model)
{
  if (model == "NoMotion") return std::make_shared<MotionPredictionNoMotion>();
  if (model == "Kalman") throw Exception("MotionPrediction: the Kalman model is not part of this path");
  return std::make_shared<MotionPredictionConstant>();  // "ConstantMotion" and the reference's fall-back (:33-37)
}

// ------------------------------------------------------------------------------------------------
// Odometry (Odometry.h:27-71, Odometry.cpp:21-62)
// ------------------------------------------------------------------------------------------------
class Odometry
{
public:
  typedef std::shared_ptr<Odometry> ShPtr;
  virtual ~Odometry() = default;
  virtual void update(Frame::ConstShPtr frame) = 0;
  virtual PoseWithCovariance::ConstShPtr pose() const = 0;
  virtual PoseWithCovariance::ConstShPtr speed() const = 0;
};

class OdometryRgbd : public Odometry
{
public:
  typedef std::shared_ptr<OdometryRgbd> ShPtr;
  OdometryRgbd(double minGradient, least_squares::Solver::ShPtr solver, least_squares::Loss::ShPtr loss, Map::ConstShPtr map,
               int device = 0)
  : _aligner(std::make_shared<SE3Alignment>(minGradient, std::move(solver), std::move(loss), true, device)),  // prior on (:24)
    _map(std::move(map)),
    _includeKeyFrame(true),
    _trackKeyFrame(false)
  {
  }
  void update(Frame::ConstShPtr frame) override
  {
    if (_map->lastFrame()) {
      try {
        if (_includeKeyFrame && _map->lastKf() && _map->lastKf() != _map->lastFrame()) {
          _pose = _aligner->align(std::vector<Frame::ConstShPtr>{_map->lastKf(), _map->lastFrame()}, frame);
        } else if (_trackKeyFrame && _map->lastKf()) {
          _pose = _aligner->align(_map->lastKf(), frame);
        } else {
          _pose = _aligner->align(_map->lastFrame(), frame);
        }
        const double dT = static_cast<double>(frame->t() - _map->lastFrame()->t()) / 1e9;
        Vec6d x = se3::log(algorithm::computeRelativeTransform(_map->lastFrame()->pose().pose(), _pose->pose()));
        for (double & v : x) v /= dT;
        _speed = std::make_shared<PoseWithCovariance>(se3::exp(x), _pose->cov());
      } catch (const std::runtime_error & e) {  // Odometry.cpp:52-56: log, fall back to the predicted pose
        _lastError = e.what();
        _pose = std::make_shared<PoseWithCovariance>(frame->pose());
        _speed = std::make_shared<PoseWithCovariance>();
      }
    } else {
      _pose = std::make_shared<PoseWithCovariance>(frame->pose());
      _speed = std::make_shared<PoseWithCovariance>();
    }
  }
  PoseWithCovariance::ConstShPtr pose() const override { return _pose; }
  PoseWithCovariance::ConstShPtr speed() const override { return _speed; }
  const SE3Alignment & aligner() const { return *_aligner; }
  // what the reference writes to its "odometry" log (LOG_ODOM(ERROR) << e.what(), Odometry.cpp:53); empty if none
  const std::string & lastError() const { return _lastError; }

protected:
  const std::shared_ptr<const SE3Alignment> _aligner;
  const Map::ConstShPtr _map;
  const bool _includeKeyFrame, _trackKeyFrame;
  PoseWithCovariance::ConstShPtr _speed;
  PoseWithCovariance::ConstShPtr _pose;
  std::string _lastError;
};

// ------------------------------------------------------------------------------------------------
// TUM trajectory files (utils.cpp:76-117,134-163).  The quirks are the reference's: the fractional part of the time
// stamp is read as 1e-4 s units (x 100000 ns: TUM ground truth has four decimals) and written back without zero
// padding.
// ------------------------------------------------------------------------------------------------
namespace utils
{
inline std::map<Timestamp, SE3d> loadTrajectory(const std::string & path)
{
  std::ifstream gtFile(path);
  if (!gtFile.is_open()) throw Exception("Could not open file at: " + path);  // the reference builds the exception but never throws it
  std::map<Timestamp, SE3d> poses;
  std::string line;
  while (std::getline(gtFile, line)) {
    std::vector<std::string> elements;
    std::string s;
    std::istringstream lines(line);
    while (std::getline(lines, s, ' ')) elements.push_back(s);
    if (elements.empty() || elements[0] == "#" || elements[0].empty() || elements[0][0] == '#') continue;
    if (elements.size() < 8) throw Exception("loadTrajectory: malformed line: " + line);
    SE3d T;
    for (int i = 0; i < 3; i++) T.data()[4 + i] = std::stod(elements[1 + i]);
    double q[4] = {std::stod(elements[4]), std::stod(elements[5]), std::stod(elements[6]), std::stod(elements[7])};
    const double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);  // Sophus normalises the quaternion
    for (int i = 0; i < 4; i++) T.data()[i] = q[i] / n;
    std::vector<std::string> tElements;
    std::string st;
    std::istringstream tLine(elements[0]);
    while (std::getline(tLine, st, '.')) tElements.push_back(st);
    const auto sec = std::stoull(tElements[0]);
    const auto nanosec = tElements.size() > 1 ? std::stoull(tElements[1]) * 100000 : 0ull;
    poses.insert({static_cast<Timestamp>(sec * 1e9 + nanosec), T});
  }
  return poses;
}

inline void writeTrajectory(const std::map<Timestamp, PoseWithCovariance::ConstShPtr> & traj, const std::string & path,
                            bool writeCovariance = false)
{
  std::ofstream algoFile(path);
  if (!algoFile.is_open()) throw Exception("Could not open file at: " + path);
  algoFile << "# Algorithm Trajectory\n";
  algoFile << "# file: \"" << path << "\"\n";
  algoFile << "# timestamp tx ty tz qx qy qz qw\n";
  for (const auto & pose : traj) {
    const Timestamp sec = static_cast<Timestamp>(static_cast<double>(pose.first) / 1e9);
    const double * d = pose.second->pose().data();
    algoFile << sec << "." << static_cast<Timestamp>(static_cast<double>(pose.first) - static_cast<double>(sec) * 1e9) << " " << d[4]
             << " " << d[5] << " " << d[6] << " " << d[0] << " " << d[1] << " " << d[2] << " " << d[3];
    if (writeCovariance)
      for (int i = 0; i < 36; i++) algoFile << " " << pose.second->cov()[i];
    algoFile << "\n";
  }
}
}  // namespace utils
}  // namespace vslam_b200

#endif  // VSLAM_B200_ODOMETRY_HPP_
