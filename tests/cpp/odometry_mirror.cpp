// Drives include/vslam_b200_odometry.hpp the way NodeRgbdAlignment::processFrame drives the reference
// (src/ros/nodes/NodeRgbdAlignment.cpp:121-150): predict -> OdometryRgbd::update -> prediction update -> key-frame
// selection -> Map::insert, over a rendered synthetic sequence (the one the Python tests and bench.py use).
//   usage: odometry_mirror W H frames levels kf_period model out_trajectory.txt
// prints "pose k qx qy qz qw tx ty tz", "speed k ..." and "kind k single|stacked|first" per frame.
#include <cstdio>
#include <cstdlib>

#include "vslam_b200_odometry.hpp"

extern "C" {
void vsynth_sequence_pose(uint64_t seed, int k, double R[9], double t[3]);
void vsynth_render(int W, int H, double fx, double fy, double cx, double cy, const double R[9], const double t[3],
                   uint64_t scene_seed, int holes_permille, uint64_t hole_seed, uint8_t * gray, float * depth);
}

using namespace vslam_b200;

int main(int argc, char ** argv)
{
  if (argc < 8) {
    fprintf(stderr, "usage: odometry_mirror W H frames levels kf_period model out_trajectory\n");
    return 2;
  }
  const int W = atoi(argv[1]), H = atoi(argv[2]), F = atoi(argv[3]);
  const size_t levels = (size_t)atoi(argv[4]);
  const uint64_t period = strtoull(argv[5], nullptr, 10), seed = 1234;
  const double f = 525.0 * W / 640.0;
  auto cam = std::make_shared<const Camera>(f, f, W / 2 - 0.5, H / 2 - 0.5);
  try {
    auto map = std::make_shared<Map>();
    auto solver = std::make_shared<least_squares::GaussNewton>(1e-7, 50);
    auto loss = std::make_shared<least_squares::LossTDistribution>(std::make_shared<least_squares::ScalerTDistribution>(5.0), 5.0);
    OdometryRgbd odometry(30.0, solver, loss, map);
    auto prediction = MotionPrediction::make(argv[6]);
    KeyFrameSelectionIdx keyFrameSelection(period);
    std::map<Timestamp, PoseWithCovariance::ConstShPtr> trajectory;
    for (int k = 0; k < F; k++) {
      double R[9], t[3];
      vsynth_sequence_pose(seed, k, R, t);
      std::vector<uint8_t> g((size_t)W * H);
      std::vector<float> d((size_t)W * H);
      vsynth_render(W, H, cam->fx(), cam->fy(), cam->cx(), cam->cy(), R, t, seed, 0, 0, g.data(), d.data());
      Image I(H, W);
      DepthMapF Z(H, W);
      I.data = g;
      Z.data = d;
      const Timestamp stamp = 1000000000ull + (Timestamp)k * 33333333ull;
      auto frame = std::make_shared<Frame>(std::move(I), std::move(Z), cam, stamp);
      frame->computePyramid(levels);
      frame->computeDerivatives();
      frame->computePcl();
      // NodeRgbdAlignment::processFrame
      const bool first = !map->lastFrame();
      const bool stacked = !first && map->lastKf() && map->lastKf() != map->lastFrame();
      frame->set(*prediction->predict(frame->t()));
      odometry.update(frame);
      if (!odometry.lastError().empty()) {
        fprintf(stderr, "odometry fell back to the prediction at frame %d: %s\n", k, odometry.lastError().c_str());
        return 1;
      }
      frame->set(*odometry.pose());
      prediction->update(odometry.pose(), frame->t());
      keyFrameSelection.update(frame);
      map->insert(frame, keyFrameSelection.isKeyFrame());
      trajectory[stamp] = odometry.pose();
      printf("pose %d", k);
      for (int i = 0; i < 7; i++) printf(" %.17g", odometry.pose()->pose().data()[i]);
      printf("\nspeed %d", k);
      for (int i = 0; i < 7; i++) printf(" %.17g", odometry.speed()->pose().data()[i]);
      printf("\nkind %d %s\n", k, first ? "first" : stacked ? "stacked" : "single");
    }
    if ((int)map->frames().size() != std::min(F, 7)) {
      fprintf(stderr, "frame ring holds %zu frames\n", map->frames().size());
      return 1;
    }
    utils::writeTrajectory(trajectory, argv[7]);
    printf("ok\n");
  } catch (const std::runtime_error & e) {
    fprintf(stderr, "runtime_error: %s\n", e.what());
    return 3;
  }
  return 0;
}
