// Shared device-side declarations of the B200 direct-alignment path.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/vslam_b200.h"

namespace vb
{
constexpr int MAX_LEVELS = VSLAMB200_MAX_LEVELS;
constexpr int MAX_VIEWS = 16;          // keyframes per job
constexpr int HIST_BINS = 512;         // residual r in [-255,255] -> bin r+255 (bin 511 unused)
constexpr int NE_TERMS = 28;           // 21 (upper A) + 6 (b) + 1 (chi2)
constexpr int TRACE_STRIDE = VSLAMB200_TRACE_STRIDE;
constexpr int16_t R_NOT_INTEREST = -32768;
constexpr int16_t R_INVISIBLE = -32767;
constexpr int R_CODE_INVISIBLE = HIST_BINS - 1;  // device-side residual code: r + 255, or 511 when not visible

// One frame's pyramid in HBM.  All planes are row-major, tightly packed (pitch == width).
struct FrameDev
{
  int nLevels;
  int W[MAX_LEVELS], H[MAX_LEVELS];      // intensity / gradient dims  ((W+1)/2 rule, Frame.cpp:264)
  int Wd[MAX_LEVELS], Hd[MAX_LEVELS];    // depth dims                 (int(W*0.5) rule, algorithm.h:93-94)
  uint8_t * I[MAX_LEVELS];               // intensity(l)
  float * Z[MAX_LEVELS];                 // depth(l)
  float * Zsf[MAX_LEVELS];               // depth(l) zeroed where depth(0)(v*2^l, u*2^l) <= 0 (SE3Alignment.cpp:90-91); [0] aliases Z[0]
  int16_t * gx[MAX_LEVELS];              // dIx(l)
  int16_t * gy[MAX_LEVELS];              // dIy(l)
  double cam[MAX_LEVELS][4];             // fx fy cx cy per level (Camera.cpp:53-57)
};

struct Pose
{
  double q[4];  // x y z w
  double t[3];
};

// Per (job, keyframe) state of the current pyramid level.
constexpr int LIST_TILE = 128;
constexpr int TILE_OFF_D = 0, TILE_OFF_UVT = 512, LIST_TILE_BYTES = 1024;
// Per interest point, in list order: what the Jacobian needs (Warp.cpp:166-201, InverseCompositional.cpp:79-80)
struct alignas(16) ListRec
{
  uint32_t uvt;  // u | v << 12 | T << 24
  uint32_t g;    // dIx (low 16) | dIy (high 16)
  double zi;     // 1 / depth in fp64 (Warp.cpp:170)
};
constexpr int JSTAT = 12;  // per view: max |dIx P_i| (6), max |dIy Q_i| (6) over the list, J_i = fxJ dIx P_i + fyJ dIy Q_i

struct ViewState
{
  // static per level
  int job;
  int fromSlot, toSlot;
  int W, H;                 // template dims at this level
  int Wt, Ht;               // target image dims
  const uint8_t * I;        // target intensity(level)
  double fx, fy, cx, cy;    // projection intrinsics (_camCur = from->camera(level), Warp.cpp:136-137)
  double ifx, ify, kx, ky;  // analytic K^-1 of the keyframe camera (Camera.cpp:32-35)
  double fxJd, fyJd;        // Jacobian scale (_camRef = to->camera(level), Warp.cpp:198-199)
  // fp32 hi/lo splits of the fp64 constants: keeps the fp32 Jacobian free of coherent rounding bias
  float ifxH, ifxL, ifyH, ifyL, kxH, kxL, kyH, kyL, fxJH, fxJL, fyJH, fyJL;
  float fx32, fy32;
  // interest list, two streams in the same (row-major) order:
  //   tiles  LIST_TILE-point tiles  d f32[128] | uvt u32[128]  - what the warp phase streams (one bulk copy per tile)
  //   rec    one ListRec per point  {uvt, g, zi}               - what the Jacobian needs
  //   uvt = u | v<<12 | T<<24, d = depth, g = dIx (low 16) | dIy (high 16), zi = 1 / depth in fp64 (Warp.cpp:170)
  unsigned char * tiles;
  ListRec * rec;
  int16_t * r;              // residual code of the current iteration: r + 255, R_CODE_INVISIBLE if not visible
  uint32_t * perm;          // moments kernel: list indices of a CTA's slice ordered by residual code (first evaluation)
  unsigned jstat[JSTAT];    // float bits (non-negative): bounds of the Jacobian components over the list (lists.cu)
  int nIp;                  // retained interest points (nConstraints of this view)
  int listTicket;           // lists.cu: next chunk of this view to hand to a block (zero at the start of a level)
  // warp
  Pose poseRef;             // keyframe pose
  Pose se3;                 // poseCur * poseRef^-1
  double R[9], t[3];        // rotation matrix / translation of se3 for the residual kernel
  // fp32 fast path of the residual kernel (displacement form, see align_kernels.cu)
  float rI[9], t32[3];      // R - I and t rounded to fp32
  float epsA, epsB;         // per-view terms of the fast-path error bound
  double xPred[6];          // MotionPrior::_xPred (SE3Alignment.cpp:32)
  double xCur[6];           // log(se3), kept in step with se3
  Pose se3Acc;              // Levenberg-Marquardt: se3 at the last accepted evaluation
  // bilinear extension (ext.cu): per list entry {r_I, r_Z} of the current iteration (NaN = not visible / no depth),
  // per 2048-entry block the 28 partial sums, scale of the depth residuals
  double2 * extR;
  double * extPart;
  double scaleZ;
  const float * Zt;         // target depth(level)
  const int16_t * gxt;      // target dIx(level), dIy(level)
  const int16_t * gyt;
  double scale, offset;
  double scaleAcc, offsetAcc;  // Levenberg-Marquardt: the scale of the last accepted evaluation (what the trace reports)
};

struct JobState
{
  int nViews;
  int view0;                // index of the first view (views are contiguous)
  int multi;
  int nLevels;
  Pose toPose;              // initial guess (to->pose())
  Pose pose;                // current estimate
  double cov[36];
  double lastA[36];
  double xPrev[6];
  double chi2Prev;
  // Levenberg-Marquardt (extension): damping, b and chi2 of the last accepted evaluation (its A is lastA)
  double lambda, bAcc[6], chi2Acc;
  int accepted;
  int extCount;            // ext.cu: blocks of the job that have delivered their partial sums in this evaluation
  int iter;                 // Gauss-Newton iteration index within the level
  int nPushed;              // completed steps within the level
  int active;               // 1 while the level's solve is running
  int dead;                 // 1 after a level ended without a completed step
  int status;
  int iters[MAX_LEVELS];
  int stop[MAX_LEVELS];
  int totalPoints;          // sum of nIp over the job's views at the current level
  int evals[MAX_LEVELS];    // computeNormalEquations calls per level (completed steps + a final rejected one)
  unsigned long long pointIters;  // sum over evaluations of the interest points visited
  // SM clocks spent by the cluster's rank-0 CTA per level: phase A, histogram barrier, scale + weights, phase C,
  // block reduction, sums barrier, solve
  long long clk[MAX_LEVELS][10];
};

struct SolverParams
{
  double minGradient2;
  int lossKind;
  double huberC, lossV;
  int scalerKind;
  double scalerV;
  double minStepSize;
  int maxIterations;
  int includePrior;
  int solverKind;            // VSLAMB200_SOLVER_*
  double lmLambda0, lmUp, lmDown;
  int sampling;              // VSLAMB200_SAMPLING_*
  double depthWeight;
  int chi2Mode;              // VSLAMB200_CHI2_*
};

// ------------------------------------------------------------------------------------------------
// SE3 (unit quaternion + translation, Sophus conventions; tangent order [upsilon; omega])
// ------------------------------------------------------------------------------------------------
__host__ __device__ inline void quat_to_R(const double q[4], double R[9])
{
  const double x = q[0], y = q[1], z = q[2], w = q[3];
  const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
  const double twx = tx * w, twy = ty * w, twz = tz * w;
  const double txx = tx * x, txy = ty * x, txz = tz * x;
  const double tyy = ty * y, tyz = tz * y, tzz = tz * z;
  R[0] = 1 - (tyy + tzz);
  R[1] = txy - twz;
  R[2] = txz + twy;
  R[3] = txy + twz;
  R[4] = 1 - (txx + tzz);
  R[5] = tyz - twx;
  R[6] = txz - twy;
  R[7] = tyz + twx;
  R[8] = 1 - (txx + tyy);
}
__host__ __device__ inline void rotate(const double q[4], const double p[3], double o[3])
{
  double R[9];
  quat_to_R(q, R);
  for (int i = 0; i < 3; i++) o[i] = R[i * 3] * p[0] + R[i * 3 + 1] * p[1] + R[i * 3 + 2] * p[2];
}
__host__ __device__ inline Pose se3_mul(const Pose & a, const Pose & b)
{
  Pose o;
  const double ax = a.q[0], ay = a.q[1], az = a.q[2], aw = a.q[3];
  const double bx = b.q[0], by = b.q[1], bz = b.q[2], bw = b.q[3];
  o.q[0] = aw * bx + ax * bw + ay * bz - az * by;
  o.q[1] = aw * by + ay * bw + az * bx - ax * bz;
  o.q[2] = aw * bz + az * bw + ax * by - ay * bx;
  o.q[3] = aw * bw - ax * bx - ay * by - az * bz;
  const double n = sqrt(o.q[0] * o.q[0] + o.q[1] * o.q[1] + o.q[2] * o.q[2] + o.q[3] * o.q[3]);
  for (int i = 0; i < 4; i++) o.q[i] /= n;
  double rt[3];
  rotate(a.q, b.t, rt);
  for (int i = 0; i < 3; i++) o.t[i] = rt[i] + a.t[i];
  return o;
}
__host__ __device__ inline Pose se3_inv(const Pose & a)
{
  Pose o;
  o.q[0] = -a.q[0];
  o.q[1] = -a.q[1];
  o.q[2] = -a.q[2];
  o.q[3] = a.q[3];
  double rt[3];
  rotate(o.q, a.t, rt);
  for (int i = 0; i < 3; i++) o.t[i] = -rt[i];
  return o;
}
__host__ __device__ inline Pose se3_exp(const double xi[6])
{
  const double wx = xi[3], wy = xi[4], wz = xi[5];
  const double th2 = wx * wx + wy * wy + wz * wz;
  const double th = sqrt(th2);
  double imag, real, B, C;
  if (th2 < 1e-20) {
    const double th4 = th2 * th2;
    imag = 0.5 - th2 / 48.0 + th4 / 3840.0;
    real = 1.0 - th2 / 8.0 + th4 / 384.0;
    B = 0.5;
    C = 1.0 / 6.0;
  } else {
    imag = sin(0.5 * th) / th;
    real = cos(0.5 * th);
    B = (1.0 - cos(th)) / th2;
    C = (th - sin(th)) / (th2 * th);
  }
  Pose o;
  o.q[0] = imag * wx;
  o.q[1] = imag * wy;
  o.q[2] = imag * wz;
  o.q[3] = real;
  const double ux = xi[0], uy = xi[1], uz = xi[2];
  const double c1[3] = {wy * uz - wz * uy, wz * ux - wx * uz, wx * uy - wy * ux};
  const double c2[3] = {wy * c1[2] - wz * c1[1], wz * c1[0] - wx * c1[2], wx * c1[1] - wy * c1[0]};
  o.t[0] = ux + B * c1[0] + C * c2[0];
  o.t[1] = uy + B * c1[1] + C * c2[1];
  o.t[2] = uz + B * c1[2] + C * c2[2];
  return o;
}
__host__ __device__ inline void se3_log(const Pose & T, double xi[6])
{
  const double n2 = T.q[0] * T.q[0] + T.q[1] * T.q[1] + T.q[2] * T.q[2];
  const double n = sqrt(n2), w = T.q[3];
  double f;
  if (n2 < 1e-20) {
    f = 2.0 / w - 2.0 / 3.0 * n2 / (w * w * w);
  } else if (fabs(w) < 1e-10) {
    f = (w > 0 ? 3.14159265358979323846 : -3.14159265358979323846) / n;
  } else {
    f = 2.0 * atan(n / w) / n;
  }
  const double wx = f * T.q[0], wy = f * T.q[1], wz = f * T.q[2];
  const double th2 = wx * wx + wy * wy + wz * wz, th = sqrt(th2);
  double c;
  if (th2 < 1e-12) {
    c = 1.0 / 12.0;
  } else {
    const double hs = sin(0.5 * th), hc = cos(0.5 * th);
    c = (1.0 - th * hc / (2.0 * hs)) / th2;
  }
  const double * t = T.t;
  const double c1[3] = {wy * t[2] - wz * t[1], wz * t[0] - wx * t[2], wx * t[1] - wy * t[0]};
  const double c2[3] = {wy * c1[2] - wz * c1[1], wz * c1[0] - wx * c1[2], wx * c1[1] - wy * c1[0]};
  xi[0] = t[0] - 0.5 * c1[0] + c * c2[0];
  xi[1] = t[1] - 0.5 * c1[1] + c * c2[1];
  xi[2] = t[2] - 0.5 * c1[2] + c * c2[2];
  xi[3] = wx;
  xi[4] = wy;
  xi[5] = wz;
}

__device__ __forceinline__ int reflect101(int i, int n)
{
  if (n == 1) return 0;
  while (i < 0 || i >= n) {
    if (i < 0) i = -i;
    if (i >= n) i = 2 * n - 2 - i;
  }
  return i;
}

// ------------------------------------------------------------------------------------------------
// host-side launchers (each defined next to its kernels)
// ------------------------------------------------------------------------------------------------
// api.cu: sets the message vslamb200_last_error() returns (for entry points defined in other files); returns code
int set_last_error(int code, const char * msg);
// preprocess.cu
void launch_ingest(const uint8_t * stagedGray, const void * stagedDepth, int depthType, int layout, int W, int H,
                   const FrameDev * frames, const int * slots, int n, cudaStream_t s);
void launch_ingest_rgbd(const uint8_t * stagedRgb, const void * stagedDepth, int depthType, int bgr, double divisor, int W,
                        int H, const FrameDev * frames, const int * slots, int n, cudaStream_t s);
void launch_depth_u16(const uint16_t * staged, double divisor, int W, int H, const FrameDev * frames, const int * slots, int n,
                      cudaStream_t s);
void launch_pyramid_level(const FrameDev * frames, const int * slots, int n, int level, int W, int H, int Wd, int Hd,
                          int Wp, int Hp, int Wdp, int Hdp, cudaStream_t s);
void launch_gradients(const FrameDev * frames, const int * slots, int n, int level, int W, int H, cudaStream_t s);
// lists.cu
void launch_build_lists(const FrameDev * frames, ViewState * views, const JobState * jobs, int view0, int nViews, int level,
                        int maxPixels, double minGradient2, int * chunkCounts, int maxChunksPerView, int wantJstat,
                        cudaStream_t s);
int list_chunks_for(int pixels);
// level_misc.cu
void launch_level_setup(const FrameDev * frames, JobState * jobs, ViewState * views, int view0, int nViews, int level,
                        cudaStream_t s);
void launch_level_begin(JobState * jobs, ViewState * views, int job0, int nJobs, int level, cudaStream_t s);
void launch_stage_copy(void * dst, const void * srcPinned, size_t bytes, cudaStream_t s);
void launch_scatter_rcode(const ViewState * views, int view, int16_t * plane, cudaStream_t s);
// gn_level.cu
size_t gn_level_static_smem();
cudaError_t launch_gn_level(JobState * jobs, ViewState * views, int job0, int nJobs, int maxViewsPerJob, int clusterSize,
                            SolverParams p, int level, int maxEvals, double * trace, size_t traceJobStride, double * viewSums, cudaStream_t s);
// gn_mom.cu (moment-based level kernel: K keyframes per job, cpv CTAs per keyframe, scalers without offset)
cudaError_t launch_gn_mom(JobState * jobs, ViewState * views, int job0, int nJobs, int K, int cpv, SolverParams p, int level,
                          int maxEvals, double * trace, size_t traceJobStride, double * viewSums, cudaStream_t s);
size_t gn_mom_smem(int K);
// ext.cu (bilinear / depth-residual extension): one Gauss-Newton (or LM) evaluation + step for every active job
constexpr int EXT_CHUNK = 2048;  // list entries per block
void launch_ext_iteration(JobState * jobs, ViewState * views, int job0, int nJobs, int view0, int nViews, int maxPoints,
                          SolverParams p, int level, bool forceStop, double * trace, size_t traceJobStride, int * activeCount,
                          cudaStream_t s);
}  // namespace vb
