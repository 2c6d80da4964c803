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

// x() of the view's problem: log of the current relative transform (Warp.cpp:157)
__host__ __device__ inline void view_set_x(ViewState & V) { se3_log(V.se3, V.xCur); }

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

// Scratch of the block-cooperative Gauss-Newton step (shared memory).
struct SolveScratch
{
  double ne[NE_TERMS];  // stacked, normalised, prior-augmented terms: 21 upper A, 6 b, chi2
  double A[36];
  double dx[6];
  double det;
  double dChi2;
  Pose inc;             // exp(-dx)
  long nTotal;
  int decision;         // -1: accept the step, else VSLAMB200_STOP_*
};

// One GaussNewton::solve iteration (GaussNewton.cpp:44-99) on the stacked, normalised, prior-augmented
// normal equations, executed by the whole thread block (>= 64 threads; contains block barriers):
//   1. threads 0..27 assemble one term each over the job's views (InverseCompositional.cpp:139-147,
//      MotionPrior::apply SE3Alignment.cpp:37-47, InverseCompositionalStacked.cpp:45-46)
//   2. thread 0 takes the determinant while thread 32 solves and exponentiates the step
//   3. thread 0 decides (GaussNewton.cpp:53-75); on acceptance the views are updated in parallel
//      (one thread per view composes, then logarithm and warp constants side by side)
//   4. thread 0 records the step and applies the convergence rules (GaussNewton.cpp:77-98)
// sums: per view 28 raw terms (21 upper A, 6 b, chi2), view-major.  forceStop: end the level after this step.
__device__ __forceinline__ void gn_step_block(
  JobState & J, ViewState * jv, const double * sums, SolveScratch & X, const SolverParams & p, int level,
  double * trace, bool forceStop)
{
  const int tid = threadIdx.x;
  const int K = J.nViews;
  // ---- 1. assemble
  if (tid < NE_TERMS) {
    // which entry of A / b this term is
    int di = -1, bi = -1;  // diagonal index / b index
    {
      int idx = 0;
      for (int i = 0; i < 6; i++)
        for (int j = i; j < 6; j++) {
          if (idx == tid && i == j) di = i;
          idx++;
        }
      if (tid >= 21 && tid < 27) bi = tid - 21;
    }
    double acc = 0.0;
    long nTotal = 0;
    for (int k = 0; k < K; k++) {
      const ViewState & V = jv[k];
      double v = sums[k * NE_TERMS + tid];
      const long n = V.nIp;
      if (n > 1) v /= (double)n;                 // InverseCompositional.cpp:139-143
      if (p.includePrior && tid < 27) {          // MotionPrior::apply (chi2 is left alone)
        v *= 1.0 / (255.0 * 255.0);
        if (di >= 0) v += 1.0;
        if (bi >= 0) v += V.xPred[bi] - V.xCur[bi];
      }
      acc += v;
      nTotal += n;
    }
    X.ne[tid] = acc;
    if (tid == 0) X.nTotal = nTotal;
  }
  __syncthreads();
  // ---- 2. determinant || solve
  if (tid == 0 || tid == 32) {
    double A[36], b[6];
    int idx = 0;
#pragma unroll
    for (int i = 0; i < 6; i++)
#pragma unroll
      for (int j = i; j < 6; j++) {
        A[i * 6 + j] = X.ne[idx];
        A[j * 6 + i] = X.ne[idx];
        idx++;
      }
    if (tid == 0) {
#pragma unroll
      for (int i = 0; i < 36; i++) X.A[i] = A[i];
      X.det = det6(A);
    } else {
#pragma unroll
      for (int i = 0; i < 6; i++) b[i] = X.ne[21 + i];
      double dx[6], mdx[6];
      ldlt_solve(A, b, dx);
#pragma unroll
      for (int i = 0; i < 6; i++) {
        X.dx[i] = dx[i];
        mdx[i] = -dx[i];
      }
      X.inc = se3_exp(mdx);  // updateCompositional(-dx) (InverseCompositional.cpp:104, Warp.cpp:153-157)
    }
  }
  __syncthreads();
  // ---- 3. decision
  const int it = J.iter;
  if (tid == 0) {
    const double chi2 = X.ne[27];
    int dec = -1;
    X.dChi2 = it > 0 ? chi2 - J.chi2Prev : 0.0;
    if (X.nTotal < 6)
      dec = VSLAMB200_STOP_CONSTRAINTS;
    else if (!isfinite(X.det) || fabs(X.det) < 1e-6)
      dec = VSLAMB200_STOP_HESSIAN;
    else if (it > 0 && X.dChi2 > 0)
      dec = VSLAMB200_STOP_CHI2;
    X.decision = dec;
  }
  __syncthreads();
  if (X.decision < 0) {
    if (tid < K) jv[tid].se3 = se3_mul(jv[tid].se3, X.inc);
    __syncthreads();
    if (tid < K) se3_log(jv[tid].se3, jv[tid].xCur);
    if (tid >= 32 && tid < 32 + K) view_set_warp(jv[tid - 32]);
    __syncthreads();
  }
  // ---- 4. bookkeeping and stop rules
  if (tid == 0) {
    int stop = X.decision;
    bool rollback = stop == VSLAMB200_STOP_CHI2;  // problem->setX(x.row(i-1)) (GaussNewton.cpp:73)
    if (stop < 0) {
      const double * x = jv[0].xCur;
      double nrm = 0;
#pragma unroll
      for (int k = 0; k < 6; k++) nrm += X.dx[k] * X.dx[k];
      const double step = sqrt(nrm);
#pragma unroll
      for (int k = 0; k < 36; k++) J.lastA[k] = X.A[k];
      J.nPushed++;
      if (trace) {
        double * t = trace + ((size_t)level * p.maxIterations + it) * TRACE_STRIDE;
        t[0] = 1;
        t[1] = (double)X.nTotal;
        t[2] = X.ne[27];
        t[3] = step;
        t[4] = jv[0].scale;
        t[5] = jv[0].offset;
        for (int k = 0; k < 6; k++) t[6 + k] = x[k];
        for (int k = 0; k < 6; k++) t[12 + k] = X.ne[21 + k];
        for (int k = 0; k < 36; k++) t[18 + k] = X.A[k];
        for (int k = 0; k < 6; k++) t[54 + k] = X.dx[k];
      }
      double bmax = X.ne[21];
#pragma unroll
      for (int k = 1; k < 6; k++) bmax = fmax(bmax, X.ne[21 + k]);
      if (it > 0 && (step < p.minStepSize || fabs(bmax) < p.minStepSize || fabs(X.dChi2) < p.minStepSize)) {
        stop = VSLAMB200_STOP_CONVERGED;
      } else if (!isfinite(step)) {
        stop = VSLAMB200_STOP_NAN;
        rollback = it > 0;
      } else {
#pragma unroll
        for (int k = 0; k < 6; k++) J.xPrev[k] = x[k];
        J.chi2Prev = X.ne[27];
        J.iter = it + 1;
        if (J.iter >= p.maxIterations || forceStop) stop = VSLAMB200_STOP_MAXITER;
      }
    }
    if (stop >= 0) {
      if (rollback)
        for (int k = 0; k < K; k++) jv[k].se3 = se3_exp(J.xPrev);  // se3 = exp(x) for every stacked frame
      finish_level(J, jv, level, stop);
    }
  }
  __syncthreads();
}
// ------------------------------------------------------------------------------------------------
// EXTENSION: one Levenberg-Marquardt evaluation on the same stacked, normalised, prior-augmented normal equations
// (oracle/vslam_oracle.cpp: levenberg_marquardt is the definition; the reference has no LM solver).  Same calling
// convention as gn_step_block.  J.lastA / J.bAcc / J.chi2Acc / jv[k].se3Acc hold the last accepted evaluation.
//   worse than the accepted state: reject (pose <- accepted pose, lambda *= up); else accept (lambda *= down from the
//   second accept on); then dx = ldlt(A_acc + lambda diag A_acc)^-1 b_acc, se3 <- se3_acc * exp(-dx).
// ------------------------------------------------------------------------------------------------
static __device__ __noinline__ void lm_step_block(
  JobState & J, ViewState * jv, const double * sums, SolveScratch & X, const SolverParams & p, int level,
  double * trace, bool forceStop)
{
  const int tid = threadIdx.x;
  const int K = J.nViews;
  // ---- 1. assemble (as gn_step_block)
  if (tid < NE_TERMS) {
    int di = -1, bi = -1;
    {
      int idx = 0;
      for (int i = 0; i < 6; i++)
        for (int j = i; j < 6; j++) {
          if (idx == tid && i == j) di = i;
          idx++;
        }
      if (tid >= 21 && tid < 27) bi = tid - 21;
    }
    double acc = 0.0;
    long nTotal = 0;
    for (int k = 0; k < K; k++) {
      const ViewState & V = jv[k];
      double v = sums[k * NE_TERMS + tid];
      const long n = V.nIp;
      if (n > 1) v /= (double)n;
      if (p.includePrior && tid < 27) {
        v *= 1.0 / (255.0 * 255.0);
        if (di >= 0) v += 1.0;
        if (bi >= 0) v += V.xPred[bi] - V.xCur[bi];
      }
      acc += v;
      nTotal += n;
    }
    X.ne[tid] = acc;
    if (tid == 0) X.nTotal = nTotal;
  }
  __syncthreads();
  // ---- 2. accept / reject
  const int it = J.iter;
  if (tid == 0) {
    if (it == 0) {
      J.lambda = p.lmLambda0;
      J.accepted = 0;
    }
    const double chi2 = X.ne[27];
    int dec = -1, rej = 0;
    if (X.nTotal < 6) {
      dec = VSLAMB200_STOP_CONSTRAINTS;
    } else if (J.accepted && chi2 > J.chi2Acc) {
      rej = 1;
      J.lambda *= p.lmUp;
      if (J.lambda > 1e12) dec = VSLAMB200_STOP_CONVERGED;
    } else {
      const bool converged = J.accepted && fabs(chi2 - J.chi2Acc) < p.minStepSize;
      if (J.accepted) J.lambda = fmax(J.lambda * p.lmDown, 1e-12);
      int idx = 0;
      for (int i = 0; i < 6; i++)
        for (int j = i; j < 6; j++) {
          J.lastA[i * 6 + j] = X.ne[idx];
          J.lastA[j * 6 + i] = X.ne[idx];
          idx++;
        }
      for (int i = 0; i < 6; i++) J.bAcc[i] = X.ne[21 + i];
      J.chi2Acc = chi2;
      J.accepted = 1;
      if (converged) dec = VSLAMB200_STOP_CONVERGED;  // (a previous step exists: nPushed >= 1)
    }
    X.decision = dec;
    X.dChi2 = (double)rej;
  }
  __syncthreads();
  const bool rejected = X.dChi2 != 0.0;
  if (tid < K) {
    if (rejected) {
      jv[tid].se3 = jv[tid].se3Acc;
    } else {
      jv[tid].se3Acc = jv[tid].se3;
      jv[tid].scaleAcc = jv[tid].scale;
      jv[tid].offsetAcc = jv[tid].offset;
    }
  }
  __syncthreads();
  // ---- 3. damped solve (thread 0: determinant, thread 32: step)
  if (X.decision < 0 && (tid == 0 || tid == 32)) {
    double A[36];
#pragma unroll
    for (int i = 0; i < 36; i++) A[i] = J.lastA[i];
#pragma unroll
    for (int i = 0; i < 6; i++) A[i * 7] += J.lambda * J.lastA[i * 7];
    if (tid == 0) {
#pragma unroll
      for (int i = 0; i < 36; i++) X.A[i] = J.lastA[i];
      X.det = det6(A);
    } else {
      double b[6], dx[6], mdx[6];
#pragma unroll
      for (int i = 0; i < 6; i++) b[i] = J.bAcc[i];
      ldlt_solve(A, b, dx);
#pragma unroll
      for (int i = 0; i < 6; i++) {
        X.dx[i] = dx[i];
        mdx[i] = -dx[i];
      }
      X.inc = se3_exp(mdx);
    }
  }
  __syncthreads();
  if (tid == 0 && X.decision < 0 && (!isfinite(X.det) || fabs(X.det) < 1e-6)) X.decision = VSLAMB200_STOP_HESSIAN;
  __syncthreads();
  if (X.decision < 0) {
    if (tid < K) jv[tid].se3 = se3_mul(jv[tid].se3Acc, X.inc);
    __syncthreads();
  }
  if (X.decision < 0 || rejected) {  // the pose changed (a step, or the roll-back of a rejected one that ends the level)
    if (tid < K) se3_log(jv[tid].se3, jv[tid].xCur);
    if (tid >= 32 && tid < 32 + K) view_set_warp(jv[tid - 32]);
    __syncthreads();
  }
  // ---- 4. bookkeeping and stop rules
  if (tid == 0) {
    int stop = X.decision;
    if (stop < 0) {
      double nrm = 0;
#pragma unroll
      for (int k = 0; k < 6; k++) nrm += X.dx[k] * X.dx[k];
      const double step = sqrt(nrm);
      J.nPushed++;
      if (trace) {
        double * t = trace + ((size_t)level * p.maxIterations + it) * TRACE_STRIDE;
        t[0] = 1;
        t[1] = (double)X.nTotal;
        t[2] = J.chi2Acc;
        t[3] = step;
        t[4] = jv[0].scaleAcc;
        t[5] = jv[0].offsetAcc;
        for (int k = 0; k < 6; k++) t[6 + k] = jv[0].xCur[k];
        for (int k = 0; k < 6; k++) t[12 + k] = J.bAcc[k];
        for (int k = 0; k < 36; k++) t[18 + k] = J.lastA[k];
        for (int k = 0; k < 6; k++) t[54 + k] = X.dx[k];
        t[60] = J.lambda;
      }
      if (!isfinite(step)) {
        for (int k = 0; k < K; k++) jv[k].se3 = jv[k].se3Acc;
        stop = VSLAMB200_STOP_NAN;
      } else if (step < p.minStepSize) {
        stop = VSLAMB200_STOP_CONVERGED;
      } else {
        J.iter = it + 1;
        if (J.iter >= p.maxIterations || forceStop) stop = VSLAMB200_STOP_MAXITER;
      }
    }
    if (stop >= 0) finish_level(J, jv, level, stop);
  }
  __syncthreads();
}

// the level kernels' single entry point
__device__ __forceinline__ void solver_step_block(
  JobState & J, ViewState * jv, const double * sums, SolveScratch & X, const SolverParams & p, int level,
  double * trace, bool forceStop)
{
  if (p.solverKind == VSLAMB200_SOLVER_LM)
    lm_step_block(J, jv, sums, X, p, level, trace, forceStop);
  else
    gn_step_block(J, jv, sums, X, p, level, trace, forceStop);
}
}  // namespace vb
