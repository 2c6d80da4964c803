// Robust-loss helpers shared by the Gauss-Newton level kernels (Loss.cpp:51-92, Scaler.cpp:49-67).
#pragma once
#include "common.cuh"

namespace vb
{
__device__ __forceinline__ double loss_weight(const SolverParams & p, double r)
{
  switch (p.lossKind) {
    case VSLAMB200_LOSS_TUKEY: {
      const double C = 4.6851;
      if (fabs(r) < C) {
        const double rc = r / C;
        const double q = 1.0 - rc * rc;
        return q * q;
      }
      return 0.0;
    }
    case VSLAMB200_LOSS_HUBER:
      if (fabs(r) < p.huberC) return 1.0;
      return (p.huberC * r > 0.0 ? 1.0 : -1.0) / r;
    case VSLAMB200_LOSS_TDIST:
      return (p.lossV + 1.0) / (p.lossV + r * r);
    default:
      return 1.0;
  }
}

// 1 / d for a positive, normal d: 20-bit seed refined by two Newton steps (relative error ~1e-16).  The scaler's
// recurrence is a chain of dependent fp64 operations on one warp; the correctly rounded division's ~25
// instruction sequence per term is what its latency is made of.
__device__ __forceinline__ double rcp_newton(double d)
{
  double y;
  asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(d));
  double e = fma(-d, y, 1.0);
  y = fma(y, e, y);
  e = fma(-d, y, 1.0);
  y = fma(y, e, y);
  e = fma(-d, y, 1.0);
  return fma(y, e, y);
}

// ScalerTDistribution::compute (Scaler.cpp:49-67) on a residual histogram, executed by ONE full warp:
//   sigma^2 <- mean(r^2 (nu+1) / (nu + (r/sigma)^2)), from sigma = 1, <= 30 steps, until |step| <= 1e-5.
// h[b] = number of points with residual b - 255; zerosExtra more points have residual 0 (they only count in n).
// The steps are strictly sequential, so one warp runs them without block barriers: lane l owns the residual values
// whose offset from 0 is l + 32 m (both signs folded: the summand is even in r), slot m = 0..7; slots nobody
// populates are skipped warp-wide.  r^2 (nu+1) / (nu + r^2/sigma^2) == r^2 (nu+1) sigma^2 / (nu sigma^2 + r^2): one
// division per slot.
__device__ __forceinline__ double tdist_sigma_warp(const int * h, int zerosExtra, long n, double nu)
{
  const int lane = threadIdx.x & 31;
  double av[8], r2[8];
  unsigned populated = 0;
#pragma unroll
  for (int m = 0; m < 8; m++) {
    const int r = lane + 32 * m;  // 0..255
    const int cnt = r == 0 ? h[255] + zerosExtra : h[255 + r] + h[255 - r];
    r2[m] = (double)(r * r);
    av[m] = (double)cnt * (r2[m] * (nu + 1));
    if (__any_sync(0xffffffffu, cnt > 0 && r > 0)) populated |= 1u << m;
  }
  // The recurrence is carried in x = sigma^2 (x' = mean(a x / (nu x + r^2))); sigma' = sqrt(x') only feeds the
  // stopping test, so it is evaluated next to the following step's terms and the loop runs one step ahead
  // (the look-ahead step is discarded on exit).
  const double invn = 1.0 / (double)n;
  auto next_x = [&](double x) {
    const double vs2 = nu * x;
    double t0 = 0.0, t1 = 0.0;
#pragma unroll
    for (int m = 0; m < 8; m += 2) {
      if (populated & (1u << m)) t0 = fma(av[m] * x, rcp_newton(vs2 + r2[m]), t0);
      if (populated & (2u << m)) t1 = fma(av[m + 1] * x, rcp_newton(vs2 + r2[m + 1]), t1);
    }
    double term = t0 + t1;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) term += __shfl_xor_sync(0xffffffffu, term, o);
    return term * invn;
  };
  double sigma = 1.0;
  double x1 = next_x(1.0);  // sigma_1^2
  for (int it = 0; it < 30; it++) {
    const double x2 = next_x(x1);  // look-ahead: sigma_{it+2}^2
    const double sg = sqrt(x1);    // sigma_{it+1}
    const double step = fabs(sigma - sg);
    sigma = sg;
    x1 = x2;
    if (!(step > 1e-5)) break;
  }
  return sigma;
}
}  // namespace vb
