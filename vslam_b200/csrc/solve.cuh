// Device-side 6x6 dense algebra and the Gauss-Newton step (one thread per job).
#pragma once
#include "common.cuh"

namespace vb
{
// The 6x6 routines keep their matrices in registers: every loop is fully unrolled and every index is a
// compile-time constant.  Partial pivoting is done with predicated row exchanges (a sequence of compare-and-
// swap steps that brings the largest remaining |entry| of the column to the pivot row); in exact arithmetic
// this is the same elimination as picking the arg-max row, only the order of the not-yet-eliminated rows differs.

// determinant by LU with partial pivoting (GaussNewton.cpp:52)
__device__ __forceinline__ double det6(const double * A)
{
  double m[6][6];
#pragma unroll
  for (int i = 0; i < 6; i++)
#pragma unroll
    for (int j = 0; j < 6; j++) m[i][j] = A[i * 6 + j];
  double det = 1.0;
#pragma unroll
  for (int c = 0; c < 6; c++) {
#pragma unroll
    for (int r = c + 1; r < 6; r++) {
      const bool sw = fabs(m[r][c]) > fabs(m[c][c]);
#pragma unroll
      for (int k = c; k < 6; k++) {
        const double x = m[c][k], y = m[r][k];
        m[c][k] = sw ? y : x;
        m[r][k] = sw ? x : y;
      }
      det = sw ? -det : det;
    }
    if (m[c][c] == 0.0) return 0.0;
    det *= m[c][c];
    const double inv = 1.0 / m[c][c];
#pragma unroll
    for (int r = c + 1; r < 6; r++) {
      const double f = m[r][c] * inv;
#pragma unroll
      for (int k = c + 1; k < 6; k++) m[r][k] = fma(-f, m[c][k], m[r][k]);
    }
  }
  return det;
}

// LDL^T solve (GaussNewton.cpp:76)
__device__ __forceinline__ void ldlt_solve(const double * A, const double * b, double * x)
{
  double L[6][6], D[6], y[6];
#pragma unroll
  for (int j = 0; j < 6; j++) {
    double d = A[j * 6 + j];
#pragma unroll
    for (int k = 0; k < j; k++) d = fma(-L[j][k] * L[j][k], D[k], d);
    D[j] = d;
    const double inv = 1.0 / d;
#pragma unroll
    for (int i = j + 1; i < 6; i++) {
      double s = A[i * 6 + j];
#pragma unroll
      for (int k = 0; k < j; k++) s = fma(-L[i][k] * L[j][k], D[k], s);
      L[i][j] = s * inv;
    }
  }
#pragma unroll
  for (int i = 0; i < 6; i++) {
    double s = b[i];
#pragma unroll
    for (int k = 0; k < i; k++) s = fma(-L[i][k], y[k], s);
    y[i] = s;
  }
#pragma unroll
  for (int i = 5; i >= 0; i--) {
    double s = y[i] / D[i];
#pragma unroll
    for (int k = i + 1; k < 6; k++) s = fma(-L[k][i], x[k], s);
    x[i] = s;
  }
}

// inverse by Gauss-Jordan with partial pivoting (SE3Alignment.cpp:101)
static __device__ __noinline__ void inverse6(const double * A, double * out)
{
  double m[6][12];
#pragma unroll
  for (int i = 0; i < 6; i++)
#pragma unroll
    for (int j = 0; j < 6; j++) {
      m[i][j] = A[i * 6 + j];
      m[i][6 + j] = i == j ? 1.0 : 0.0;
    }
#pragma unroll
  for (int c = 0; c < 6; c++) {
#pragma unroll
    for (int r = c + 1; r < 6; r++) {
      const bool sw = fabs(m[r][c]) > fabs(m[c][c]);
#pragma unroll
      for (int k = 0; k < 12; k++) {
        const double x = m[c][k], y = m[r][k];
        m[c][k] = sw ? y : x;
        m[r][k] = sw ? x : y;
      }
    }
    const double inv = 1.0 / m[c][c];
#pragma unroll
    for (int k = 0; k < 12; k++) m[c][k] *= inv;
#pragma unroll
    for (int r = 0; r < 6; r++) {
      if (r == c) continue;
      const double f = m[r][c];
#pragma unroll
      for (int k = 0; k < 12; k++) m[r][k] = fma(-f, m[c][k], m[r][k]);
    }
  }
#pragma unroll
  for (int i = 0; i < 6; i++)
#pragma unroll
    for (int j = 0; j < 6; j++) out[i * 6 + j] = m[i][6 + j];
}

__host__ __device__ inline void view_set_warp(ViewState & V)
{
  quat_to_R(V.se3.q, V.R);
  V.t[0] = V.se3.t[0];
  V.t[1] = V.se3.t[1];
  V.t[2] = V.se3.t[2];
  double rImax = 0.0, tmax = 0.0;
  for (int i = 0; i < 9; i++) {
    const double d = V.R[i] - ((i % 4 == 0) ? 1.0 : 0.0);
    V.rI[i] = (float)d;
    rImax = fmax(rImax, fabs(d));
  }
  for (int i = 0; i < 3; i++) {
    V.t32[i] = (float)V.t[i];
    tmax = fmax(tmax, fabs(V.t[i]));
  }
  // fast-path error bound: eps = epsA*d/qz + epsB/qz + ... (derivation in DESIGN.md, "filtered warp")
  const double u128 = 128.0 / 16777216.0, fmx = fmax(V.fx, V.fy);
  V.epsA = (float)(u128 * fmx * rImax * 4.25);
  V.epsB = (float)(u128 * fmx * tmax * 1.7);
}

// End of a level's solve: covariance and pose hand-off (SE3Alignment.cpp:100-102)
// jv: the job's views (jv[0] is the job's first keyframe)
static __device__ __noinline__ void finish_level(JobState & J, ViewState * jv, int level, int stop)
{
  J.stop[level] = stop;
  J.iters[level] = J.nPushed;
  J.active = 0;
  if (J.nPushed == 0) {
    // reference: results->iteration never set -> out_of_range / UB.  Defined here: give the input pose back.
    J.status = 1;
    J.dead = 1;
    J.pose = J.toPose;
    for (int i = 0; i < 36; i++) J.cov[i] = (i % 7 == 0) ? 1.0 : 0.0;
    return;
  }
  inverse6(J.lastA, J.cov);
  const ViewState & V0 = jv[0];
  J.pose = se3_mul(V0.se3, V0.poseRef);  // WarpSE3::poseCur (Warp.cpp:239)
}

// One GaussNewton::solve iteration (GaussNewton.cpp:44-99) on the stacked, normalised, prior-augmented
// normal equations.  sums: per view 28 raw terms (21 upper A, 6 b, chi2), view-major.
static __device__ __noinline__ void gn_step(
  JobState & J, ViewState * jv, const double * sums, const SolverParams & p, int level, double * trace)
{
  double A[36], b[6], chi2 = 0.0;
  for (int i = 0; i < 36; i++) A[i] = 0.0;
  for (int i = 0; i < 6; i++) b[i] = 0.0;
  long nTotal = 0;
  for (int k = 0; k < J.nViews; k++) {
    ViewState & V = jv[k];
    const double * s = sums + k * NE_TERMS;
    double Av[36], bv[6], c2 = s[27];
    int idx = 0;
    for (int i = 0; i < 6; i++)
      for (int j = i; j < 6; j++) {
        Av[i * 6 + j] = s[idx];
        Av[j * 6 + i] = s[idx++];
      }
    for (int i = 0; i < 6; i++) bv[i] = s[21 + i];
    const long n = V.nIp;
    if (n > 1) {  // InverseCompositional.cpp:139-143
      const double dn = (double)n;
      for (int i = 0; i < 36; i++) Av[i] /= dn;
      for (int i = 0; i < 6; i++) bv[i] /= dn;
      c2 /= dn;
    }
    if (p.includePrior) {  // MotionPrior::apply (SE3Alignment.cpp:37-47)
      const double normalizer = 1.0 / (255.0 * 255.0);
      double x[6];
      se3_log(V.se3, x);
      for (int i = 0; i < 36; i++) Av[i] *= normalizer;
      for (int i = 0; i < 6; i++) bv[i] *= normalizer;
      for (int i = 0; i < 6; i++) {
        Av[i * 6 + i] += 1.0;
        bv[i] += V.xPred[i] - x[i];
      }
    }
    for (int i = 0; i < 36; i++) A[i] += Av[i];
    for (int i = 0; i < 6; i++) b[i] += bv[i];
    chi2 += c2;
    nTotal += n;
  }
  const int i = J.iter;
  const double det = det6(A);
  if (nTotal < 6) return finish_level(J, jv, level, VSLAMB200_STOP_CONSTRAINTS);
  if (!isfinite(det) || fabs(det) < 1e-6) return finish_level(J, jv, level, VSLAMB200_STOP_HESSIAN);
  const double dChi2 = i > 0 ? chi2 - J.chi2Prev : 0.0;
  if (i > 0 && dChi2 > 0) {
    // problem->setX(x.row(i-1)) (GaussNewton.cpp:73): se3 = exp(x) for every stacked frame
    for (int k = 0; k < J.nViews; k++) jv[k].se3 = se3_exp(J.xPrev);
    return finish_level(J, jv, level, VSLAMB200_STOP_CHI2);
  }
  double dx[6], mdx[6];
  ldlt_solve(A, b, dx);
  for (int k = 0; k < 6; k++) mdx[k] = -dx[k];
  const Pose inc = se3_exp(mdx);  // updateCompositional(-dx) (InverseCompositional.cpp:104, Warp.cpp:153-157)
  for (int k = 0; k < J.nViews; k++) {
    ViewState & V = jv[k];
    V.se3 = se3_mul(V.se3, inc);
  }
  double x[6];
  se3_log(jv[0].se3, x);
  double nrm = 0;
  for (int k = 0; k < 6; k++) nrm += dx[k] * dx[k];
  const double step = sqrt(nrm);
  for (int k = 0; k < 36; k++) J.lastA[k] = A[k];
  J.nPushed++;
  if (trace) {
    double * t = trace + ((size_t)level * p.maxIterations + i) * TRACE_STRIDE;
    t[0] = 1;
    t[1] = (double)nTotal;
    t[2] = chi2;
    t[3] = step;
    t[4] = jv[0].scale;
    t[5] = jv[0].offset;
    for (int k = 0; k < 6; k++) t[6 + k] = x[k];
    for (int k = 0; k < 6; k++) t[12 + k] = b[k];
    for (int k = 0; k < 36; k++) t[18 + k] = A[k];
    for (int k = 0; k < 6; k++) t[54 + k] = dx[k];
  }
  double bmax = b[0];
  for (int k = 1; k < 6; k++) bmax = fmax(bmax, b[k]);
  if (i > 0 && (step < p.minStepSize || fabs(bmax) < p.minStepSize || fabs(dChi2) < p.minStepSize))
    return finish_level(J, jv, level, VSLAMB200_STOP_CONVERGED);
  if (!isfinite(step)) {
    if (i > 0)
      for (int k = 0; k < J.nViews; k++) jv[k].se3 = se3_exp(J.xPrev);
    return finish_level(J, jv, level, VSLAMB200_STOP_NAN);
  }
  for (int k = 0; k < 6; k++) J.xPrev[k] = x[k];
  J.chi2Prev = chi2;
  J.iter = i + 1;
  if (J.iter >= p.maxIterations) return finish_level(J, jv, level, VSLAMB200_STOP_MAXITER);
  for (int k = 0; k < J.nViews; k++) view_set_warp(jv[k]);
}
}  // namespace vb
