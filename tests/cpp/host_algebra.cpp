// Host-only checks of include/vslam_b200_odometry.hpp that need no GPU: SE3 algebra, Map rings, key-frame
// selection, motion prediction, TUM trajectory files.  Reads twists from argv, prints results for the Python test.
//   usage: host_algebra trajectory_in.txt trajectory_out.txt x0 .. x5 y0 .. y5
#include <cstdio>
#include <cstdlib>

#include "vslam_b200_odometry.hpp"

using namespace vslam_b200;

static void print7(const char * tag, const SE3d & T)
{
  printf("%s", tag);
  for (int i = 0; i < 7; i++) printf(" %.17g", T.data()[i]);
  printf("\n");
}

int main(int argc, char ** argv)
{
  if (argc < 15) return 2;
  Vec6d x, y;
  for (int i = 0; i < 6; i++) {
    x[i] = atof(argv[3 + i]);
    y[i] = atof(argv[9 + i]);
  }
  const SE3d X = se3::exp(x), Y = se3::exp(y);
  print7("expx", X);
  print7("mul", se3::mul(X, Y));
  print7("inv", se3::inverse(X));
  print7("rel", algorithm::computeRelativeTransform(X, Y));
  const Vec6d l = se3::log(se3::mul(X, Y));
  printf("log");
  for (double v : l) printf(" %.17g", v);
  printf("\n");

  // Map rings (Map.cpp:19-33) and KeyFrameSelectionIdx (KeyFrameSelection.h:36-51)
  auto cam = std::make_shared<const Camera>(100.0, 100.0, 15.5, 11.5);
  Map map;
  KeyFrameSelectionIdx sel(3);
  printf("kf");
  for (int k = 0; k < 12; k++) {
    auto f = std::make_shared<Frame>(Image(24, 32), DepthMap(24, 32), cam, (Timestamp)k);
    sel.update(f);
    map.insert(f, sel.isKeyFrame());
    printf(" %d", sel.isKeyFrame() ? 1 : 0);
  }
  printf("\nrings %zu %zu %llu %llu\n", map.frames().size(), map.keyFrames().size(), (unsigned long long)map.lastFrame()->t(),
         (unsigned long long)map.lastKf()->t());

  // constant-motion prediction (MotionPrediction.cpp:38-55)
  MotionPredictionConstant pred;
  pred.update(std::make_shared<PoseWithCovariance>(X, PoseWithCovariance().cov()), 1000000000ull);
  pred.update(std::make_shared<PoseWithCovariance>(se3::mul(Y, X), PoseWithCovariance().cov()), 1500000000ull);
  print7("predict", pred.predict(2500000000ull)->pose());
  bool threw = false;
  try {
    pred.update(std::make_shared<PoseWithCovariance>(), 1u);
  } catch (const std::runtime_error &) {
    threw = true;
  }
  printf("older %d\n", threw ? 1 : 0);

  // TUM trajectory files (utils.cpp:76-117,134-163)
  const auto traj = utils::loadTrajectory(argv[1]);
  for (const auto & p : traj) {
    printf("stamp %llu", (unsigned long long)p.first);
    for (int i = 0; i < 7; i++) printf(" %.17g", p.second.data()[i]);
    printf("\n");
  }
  std::map<Timestamp, PoseWithCovariance::ConstShPtr> out;
  for (const auto & p : traj) out[p.first] = std::make_shared<PoseWithCovariance>(p.second, PoseWithCovariance().cov());
  utils::writeTrajectory(out, argv[2]);
  printf("ok\n");
  return 0;
}
