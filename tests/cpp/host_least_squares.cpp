// Host-only checks of the least-squares half of include/vslam_b200.hpp (no GPU): scalers, losses and their reference
// constructor signatures (Loss.h:54-116, Scaler.h:30-74), NormalEquations (NormalEquations.cpp:53-76), and
// GaussNewton::solve (GaussNewton.cpp:33-102) on a linear problem.  Prints values for the Python test.
#include <cstdio>

#include "vslam_b200.hpp"

using namespace vslam_b200;
namespace ls = vslam_b200::least_squares;

// r(x) = J x - y; x <- x - dx (the sign convention of the aligner's problems: the step is subtracted)
class LinearProblem : public ls::Problem
{
public:
  LinearProblem(ls::MatXd J, ls::VecXd y) : ls::Problem(J.cols()), _J(std::move(J)), _y(std::move(y)), _x(_J.cols(), 0.0) {}
  void updateX(const ls::VecXd & dx) override
  {
    for (size_t i = 0; i < _x.size(); i++) _x[i] -= dx[i];
  }
  void setX(const ls::VecXd & x) override { _x = x; }
  ls::VecXd x() const override { return _x; }
  ls::NormalEquations::ConstShPtr computeNormalEquations() override
  {
    ls::VecXd r(_y.size()), w(_y.size(), 1.0);
    for (size_t k = 0; k < _y.size(); k++) {
      r[k] = -_y[k];
      for (size_t i = 0; i < _x.size(); i++) r[k] += _J(k, i) * _x[i];
    }
    return std::make_shared<ls::NormalEquations>(_J, r, w);
  }

private:
  ls::MatXd _J;
  ls::VecXd _y, _x;
};

int main()
{
  // the residual vector the Python test regenerates
  ls::VecXd r(200);
  for (int k = 0; k < 200; k++) r[k] = (double)((k * 37) % 101 - 50) * 0.37;
  const ls::Scaler::Scale id = ls::Scaler().compute(r), md = ls::MedianScaler().compute(r), mn = ls::MeanScaler().compute(r),
                          td = ls::ScalerTDistribution(5.0).compute(r);
  printf("scale %.17g %.17g %.17g %.17g %.17g %.17g %.17g %.17g\n", id.offset, id.scale, md.offset, md.scale, mn.offset, mn.scale,
         td.offset, td.scale);
  printf("median4 %.17g median2 %.17g median5 %.17g\n", ls::detail::median({4, 1, 3, 2}), ls::detail::median({7, 5}),
         ls::detail::median({5, 1, 4, 2, 3}));

  // constructor signatures and defaults of the reference (Loss.h:70,87,101; NodeRgbdAlignment.cpp:76)
  ls::TukeyLoss tukey;
  ls::HuberLoss huberDefault, huberMean(std::make_shared<ls::MeanScaler>()), huberNode(std::make_shared<ls::MedianScaler>(), 10.0);
  ls::LossTDistribution tDefault, tNode(std::make_shared<ls::ScalerTDistribution>(5.0), 5.0);
  ls::QuadraticLoss quad;
  printf("kinds %d %d %d %d %d %d %d  scalers %d %d %d %d %d %d %d  c %.17g %.17g\n", tukey.kind(), huberDefault.kind(),
         huberMean.kind(), huberNode.kind(), tDefault.kind(), tNode.kind(), quad.kind(), tukey.scaler().kind(),
         huberDefault.scaler().kind(), huberMean.scaler().kind(), huberNode.scaler().kind(), tDefault.scaler().kind(),
         tNode.scaler().kind(), quad.scaler().kind(), huberDefault.c(), huberNode.c());
  const double xs[6] = {-20.0, -3.0, -0.5, 0.25, 2.0, 12.0};
  for (double x : xs)
    printf("w %.17g %.17g %.17g %.17g %.17g  rho %.17g %.17g %.17g\n", x, tukey.computeWeight(x), huberNode.computeWeight(x),
           tNode.computeWeight(x), quad.computeWeight(x), tukey.compute(x), huberNode.compute(x), quad.compute(x));
  const ls::Scaler::Scale viaLoss = tNode.computeScale(r);
  printf("lossscale %.17g %.17g\n", viaLoss.offset, viaLoss.scale);

  // normal equations of a deterministic 40 x 6 system with weights
  const int N = 40, P = 6;
  ls::MatXd J(N, P);
  ls::VecXd y(N), w(N);
  for (int k = 0; k < N; k++) {
    for (int i = 0; i < P; i++) J(k, i) = std::sin(0.7 * k + 1.3 * i) + (i == k % P ? 1.5 : 0.0);
    y[k] = std::cos(0.31 * k) * 2.0;
    w[k] = 0.5 + (k % 5) * 0.25;
  }
  ls::NormalEquations ne(J, y, w);
  printf("ne %zu %.17g", ne.nConstraints(), ne.chi2());
  for (int i = 0; i < P; i++) printf(" %.17g", ne.b()[i]);
  for (int i = 0; i < P; i++)
    for (int j = 0; j < P; j++) printf(" %.17g", ne.A()(i, j));
  printf("\n");
  ls::NormalEquations sum(6);
  for (int k = 0; k < N; k++) sum.addConstraint(J.row(k), y[k], w[k]);
  ls::NormalEquations twice(std::vector<ls::NormalEquations>{ne, sum});
  printf("combine %zu %.17g %.17g %.17g\n", twice.nConstraints(), twice.chi2(), twice.A()(2, 3), 2.0 * ne.A()(2, 3));
  printf("det %.17g\n", ls::detail::determinant(ne.A()));
  const ls::VecXd sol = ls::detail::ldltSolve(ne.A(), ne.b());
  printf("solve");
  for (double v : sol) printf(" %.17g", v);
  printf("\n");

  // Gauss-Newton on the linear problem: one step reaches the minimum, the next evaluation stops the loop
  auto problem = std::make_shared<LinearProblem>(J, y);
  ls::GaussNewton gn(1e-9, 10);
  auto res = gn.solve(problem);
  printf("gn %zu %zu %d %.17g", res->nSteps(), res->iteration, res->stop, res->chi2[0]);
  for (double v : problem->x()) printf(" %.17g", v);
  printf("\n");
  bool threw = false;
  try {
    ls::LevenbergMarquardt(1e-7, 10).solve(problem);
  } catch (const std::runtime_error &) {
    threw = true;
  }
  // per-level accessors need a resident frame
  bool threw2 = false;
  try {
    Frame f(Image(24, 32), DepthMap(24, 32), std::make_shared<const Camera>(100.0, 100.0, 15.5, 11.5));
    f.computePyramid(2);
    (void)f.camera(1)->fx();
    f.dIx(1);
  } catch (const std::runtime_error &) {
    threw2 = true;
  }
  printf("throws %d %d\nok\n", threw ? 1 : 0, threw2 ? 1 : 0);
  return 0;
}
