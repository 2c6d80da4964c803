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
// (Vec6d and the SE3 algebra, namespace se3, live in vslam_b200.hpp: the least-squares problem of the aligner uses them too)

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
  // Behaviour of Odometry.cpp:31-62.  Which frames the new one is aligned against:
  //   key frame AND last frame (stacked) when both exist and differ and key frames are included (the default);
  //   the key frame alone in key-frame tracking mode; otherwise the last frame.
  // Without a previous frame, or when the aligner throws (the reference logs e.what() and carries on), the frame's own
  // pose - the motion model's prediction - is the estimate and the speed is the identity.
  void update(Frame::ConstShPtr frame) override
  {
    _pose = nullptr;
    const Frame::ConstShPtr last = _map->lastFrame(), key = _map->lastKf();
    if (last) {
      try {
        _pose = alignAgainst(last, key, frame);
      } catch (const std::runtime_error & e) {
        _lastError = e.what();
      }
    }
    if (!_pose) {
      _pose = std::make_shared<PoseWithCovariance>(frame->pose());
      _speed = std::make_shared<PoseWithCovariance>();
      return;
    }
    // speed = exp(log(motion since the last frame) / elapsed seconds), with the covariance of the new pose
    const double seconds = static_cast<double>(frame->t() - last->t()) / 1e9;
    Vec6d twist = se3::log(algorithm::computeRelativeTransform(last->pose().pose(), _pose->pose()));
    for (double & v : twist) v /= seconds;
    _speed = std::make_shared<PoseWithCovariance>(se3::exp(twist), _pose->cov());
  }
  PoseWithCovariance::ConstShPtr pose() const override { return _pose; }
  PoseWithCovariance::ConstShPtr speed() const override { return _speed; }
  const SE3Alignment & aligner() const { return *_aligner; }
  // what the reference writes to its "odometry" log (LOG_ODOM(ERROR) << e.what(), Odometry.cpp:53); empty if none
  const std::string & lastError() const { return _lastError; }

protected:
  PoseWithCovariance::ConstShPtr alignAgainst(const Frame::ConstShPtr & last, const Frame::ConstShPtr & key,
                                              const Frame::ConstShPtr & frame) const
  {
    if (_includeKeyFrame && key && key != last) return _aligner->align(std::vector<Frame::ConstShPtr>{key, last}, frame);
    if (_trackKeyFrame && key) return _aligner->align(key, frame);
    return _aligner->align(last, frame);
  }
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
namespace detail
{
inline std::vector<std::string> split(const std::string & text, char sep)
{
  std::vector<std::string> out;
  size_t from = 0;
  for (;;) {
    const size_t to = text.find(sep, from);
    out.push_back(text.substr(from, to == std::string::npos ? std::string::npos : to - from));
    if (to == std::string::npos) return out;
    from = to + 1;
  }
}
// "1305031102.1758" -> nanoseconds.  The digits behind the point count units of 1e-4 s whatever their number (TUM ground
// truth has four decimals), and the sum goes through a double - both as utils.cpp:104-112 has it.
inline Timestamp stampFromTum(const std::string & field)
{
  const std::vector<std::string> parts = split(field, '.');
  const unsigned long long whole = std::stoull(parts[0]);
  const unsigned long long frac = parts.size() > 1 ? std::stoull(parts[1]) * 100000ull : 0ull;
  return static_cast<Timestamp>(whole * 1e9 + frac);
}
// nanoseconds -> "sec.nanosec" without zero padding of the fraction (utils.cpp:150-152)
inline std::string stampToTum(Timestamp ns)
{
  const Timestamp whole = static_cast<Timestamp>(static_cast<double>(ns) / 1e9);
  const Timestamp rest = static_cast<Timestamp>(static_cast<double>(ns) - static_cast<double>(whole) * 1e9);
  return std::to_string(whole) + "." + std::to_string(rest);
}
}  // namespace detail

// One pose per line: "timestamp tx ty tz qx qy qz qw", '#' lines are comments.
inline std::map<Timestamp, SE3d> loadTrajectory(const std::string & path)
{
  std::ifstream in(path);
  if (!in.is_open()) throw Exception("Could not open file at: " + path);  // the reference builds the exception but never throws it
  std::map<Timestamp, SE3d> trajectory;
  for (std::string row; std::getline(in, row);) {
    const std::vector<std::string> f = detail::split(row, ' ');
    if (f[0].empty() || f[0][0] == '#') continue;
    if (f.size() < 8) throw Exception("loadTrajectory: malformed line: " + row);
    double v[7];  // tx ty tz qx qy qz qw
    for (int i = 0; i < 7; i++) v[i] = std::stod(f[1 + i]);
    const double qn = std::sqrt(v[3] * v[3] + v[4] * v[4] + v[5] * v[5] + v[6] * v[6]);  // Sophus normalises the quaternion
    SE3d T;
    for (int i = 0; i < 4; i++) T.data()[i] = v[3 + i] / qn;
    for (int i = 0; i < 3; i++) T.data()[4 + i] = v[i];
    trajectory.emplace(detail::stampFromTum(f[0]), T);
  }
  return trajectory;
}

// Three header lines, then one pose per line in the order above (optionally followed by the 36 covariance entries).
inline void writeTrajectory(const std::map<Timestamp, PoseWithCovariance::ConstShPtr> & traj, const std::string & path,
                            bool writeCovariance = false)
{
  std::ofstream out(path);
  if (!out.is_open()) throw Exception("Could not open file at: " + path);
  out << "# Algorithm Trajectory\n# file: \"" << path << "\"\n# timestamp tx ty tz qx qy qz qw\n";
  static const int order[7] = {4, 5, 6, 0, 1, 2, 3};  // translation first, then the quaternion
  for (const auto & entry : traj) {
    out << detail::stampToTum(entry.first);
    for (int k : order) out << " " << entry.second->pose().data()[k];
    if (writeCovariance)
      for (double c : entry.second->cov()) out << " " << c;
    out << "\n";
  }
}
}  // namespace utils
}  // namespace vslam_b200

#endif  // VSLAM_B200_ODOMETRY_HPP_
