// Exercises include/vslam_b200.hpp the way the reference's own alignment test drives SE3Alignment
// (odometry/test/test_alignment_se3.cpp:66-77,119-170): build frames, computePyramid / Derivatives / Pcl, align.
// The frame pair is rendered by the synthetic-scene library the Python tests use, in the Eigen memory order
// (column-major u8 intensity, column-major double depth) so that the transposing ingest path is covered.
//   usage: host_mirror W H seed levels [multi]
// prints "pose qx qy qz qw tx ty tz" (and "batch ..." lines); exit 3 + message when no CUDA device exists.
#include <algorithm>
#include <cmath>
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
    // ---- the least-squares interfaces (Solver.h:22-63): solver->solve(problem) on ONE level drives the device path
    // iteration by iteration and reproduces the iterates of SE3Alignment::align on frames with a one-level pyramid
    for (int withPrior = 0; withPrior < 2; withPrior++) {
      auto g0 = make_frame(W, H, cam, I3, z3, seed, 1);
      auto g1 = make_frame(W, H, cam, R, t, seed, 1);
      const size_t maxIt = 6;
      auto gn = std::make_shared<least_squares::GaussNewton>(1e-7, maxIt);
      SE3Alignment one(30.0, gn, loss, withPrior != 0);
      check(vslamb200_set_trace(one.context(), 1));
      auto viaAlign = one.align(g0, g1);
      std::vector<double> tr(maxIt * VSLAMB200_TRACE_STRIDE);
      check(vslamb200_get_trace(one.context(), 0, tr.data()));
      size_t nTrace = 0;
      while (nTrace < maxIt && tr[nTrace * VSLAMB200_TRACE_STRIDE] == 1.0) nTrace++;
      least_squares::Prior::ConstShPtr prior;
      if (withPrior) prior = std::make_shared<MotionPrior>(g1->pose(), g0->pose());
      auto problem = std::make_shared<lukas_kanade::InverseCompositional>(one, g0, g1, 0, SE3d(), prior);
      std::shared_ptr<least_squares::Solver> solver = gn;
      auto res = solver->solve(problem);
      double worst = 0.0;
      for (size_t i = 0; i < std::min(nTrace, res->nSteps()); i++) {
        const double * rec = &tr[i * VSLAMB200_TRACE_STRIDE];
        for (int k = 0; k < 6; k++) worst = std::max(worst, std::fabs(res->x(i, k) - rec[6 + k]));
        worst = std::max(worst, std::fabs(res->chi2[i] - rec[2]) / std::fabs(rec[2]));
        worst = std::max(worst, std::fabs(res->stepSize[i] - rec[3]));
        for (int a = 0; a < 6; a++)
          for (int b = 0; b < 6; b++)
            worst = std::max(worst, std::fabs(res->normalEquations[i]->A()(a, b) - rec[18 + 6 * a + b]) / std::fabs(rec[18]));
      }
      double poseDiff = 0.0;
      for (int k = 0; k < 7; k++) poseDiff = std::max(poseDiff, std::fabs(problem->se3().data()[k] - viaAlign->pose().data()[k]));
      printf("solve%d %zu %zu %.3g %.3g %d\n", withPrior, res->nSteps(), nTrace, worst, poseDiff, res->stop);
    }
    // ---- per-level accessors of a resident frame (Frame.h:45-83): checksums for the Python test
    {
      printf("levels");
      for (size_t l = 0; l < levels; l++) {
        const Image & I = f1->intensity(l);
        const auto & gx = f1->dIx(l);
        const auto & gy = f1->dIy(l);
        const DepthMap & Z = f1->depth(l);
        double sI = 0, sx = 0, sy = 0, sZ = 0;
        for (auto v : I.data) sI += v;
        for (auto v : gx.data) sx += std::fabs(v);
        for (auto v : gy.data) sy += std::fabs(v);
        for (auto v : Z.data) sZ += v;
        const auto pc = f1->pcl(l);
        const size_t mid = pc.size() / 2 + (size_t)Z.cols / 3;
        printf(" %d %d %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g", I.cols, I.rows, sI, sx, sy, sZ, f1->camera(l)->fx(),
               pc[mid][0], pc[mid][1], pc[mid][2]);
      }
      printf("\n");
    }
    printf("ok\n");
  } catch (const std::runtime_error & e) {
    fprintf(stderr, "runtime_error: %s\n", e.what());
    return 3;
  }
  return 0;
}
