// Exercises include/vslam_b200.hpp the way the reference's own alignment test drives SE3Alignment
// (odometry/test/test_alignment_se3.cpp:66-77,119-170): build frames, computePyramid / Derivatives / Pcl, align.
// The frame pair is rendered by the synthetic-scene library the Python tests use, in the Eigen memory order
// (column-major u8 intensity, column-major double depth) so that the transposing ingest path is covered.
//   usage: host_mirror W H seed levels [multi]
// prints "pose qx qy qz qw tx ty tz" (and "batch ..." lines); exit 3 + message when no CUDA device exists.
#include <cstdio>
#include <cstdlib>

#include "vslam_b200.hpp"

extern "C" {
void vsynth_se3_exp(const double xi[6], double R[9], double t[3]);
void vsynth_pair_twist(uint64_t seed, double xi[6]);
void vsynth_render(int W, int H, double fx, double fy, double cx, double cy, const double R[9], const double t[3],
                   uint64_t scene_seed, int holes_permille, uint64_t hole_seed, uint8_t * gray, float * depth);
}

using namespace vslam_b200;

static Frame::ShPtr make_frame(int W, int H, const Camera::ConstShPtr & cam, const double R[9], const double t[3],
                               uint64_t seed, size_t levels)
{
  std::vector<uint8_t> g((size_t)W * H);
  std::vector<float> d((size_t)W * H);
  vsynth_render(W, H, cam->fx(), cam->fy(), cam->cx(), cam->cy(), R, t, seed, 0, 0, g.data(), d.data());
  Image I(H, W, /*colMajor=*/true);
  DepthMap Z(H, W, /*colMajor=*/true);
  for (int v = 0; v < H; v++)
    for (int u = 0; u < W; u++) {
      I(v, u) = g[(size_t)v * W + u];
      Z(v, u) = d[(size_t)v * W + u];
    }
  auto f = std::make_shared<Frame>(std::move(I), std::move(Z), cam);
  f->computePyramid(levels);
  f->computeDerivatives();
  f->computePcl();
  return f;
}

int main(int argc, char ** argv)
{
  const int W = argc > 1 ? atoi(argv[1]) : 320, H = argc > 2 ? atoi(argv[2]) : 240;
  const uint64_t seed = argc > 3 ? strtoull(argv[3], nullptr, 10) : 1234;
  const size_t levels = argc > 4 ? (size_t)atoi(argv[4]) : 3;
  const double f = 525.0 * W / 640.0;
  auto cam = std::make_shared<const Camera>(f, f, W / 2 - 0.5, H / 2 - 0.5);
  try {
    // an inconsistent camera must be rejected like Frame.cpp:150-155
    bool threw = false;
    try {
      Frame bad(Image(H, W), DepthMap(H, W), std::make_shared<const Camera>(f, f, 5.0, 5.0));
    } catch (const std::runtime_error &) {
      threw = true;
    }
    if (!threw) {
      fprintf(stderr, "inconsistent camera accepted\n");
      return 1;
    }
    const double I3[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, z3[3] = {0, 0, 0};
    double xi[6], R[9], t[3];
    vsynth_pair_twist(seed, xi);
    vsynth_se3_exp(xi, R, t);
    auto f0 = make_frame(W, H, cam, I3, z3, seed, levels);
    auto f1 = make_frame(W, H, cam, R, t, seed, levels);

    auto solver = std::make_shared<least_squares::GaussNewton>(1e-7, 30);
    auto loss = std::make_shared<least_squares::LossTDistribution>(std::make_shared<least_squares::ScalerTDistribution>(5.0), 5.0);
    SE3Alignment aligner(30.0, solver, loss, false);
    const AlignmentSE3 & iface = aligner;  // through the abstract interface, like OdometryRgbd does
    auto pose = iface.align(f0, f1);
    printf("pose");
    for (int k = 0; k < 7; k++) printf(" %.17g", pose->pose().data()[k]);
    printf("\n");
    // the vector overload with one keyframe, and the batched addition (same pair twice: identical results)
    auto poseM = aligner.align(std::vector<Frame::ConstShPtr>{f0}, f1);
    printf("multi");
    for (int k = 0; k < 7; k++) printf(" %.17g", poseM->pose().data()[k]);
    printf("\n");
    SE3Alignment::Pair p;
    p.from = {f0};
    p.to = f1;
    auto both = aligner.align(std::vector<SE3Alignment::Pair>{p, p});
    for (int k = 0; k < 7; k++)
      if (both[0]->pose().data()[k] != pose->pose().data()[k] || both[1]->pose().data()[k] != pose->pose().data()[k]) {
        fprintf(stderr, "batched result differs from the single call\n");
        return 1;
      }
    printf("cov00 %.17g\n", pose->cov()[0]);
    // a frame that dies is released from the device on the next call
    f0.reset();
    auto f2 = make_frame(W, H, cam, I3, z3, seed, levels);
    auto again = aligner.align(f2, f1);
    for (int k = 0; k < 7; k++)
      if (again->pose().data()[k] != pose->pose().data()[k]) {
        fprintf(stderr, "re-uploaded frame gives a different result\n");
        return 1;
      }
    printf("ok\n");
  } catch (const std::runtime_error & e) {
    fprintf(stderr, "runtime_error: %s\n", e.what());
    return 3;
  }
  return 0;
}
