/*
 * vslam_b200 — C ABI of the B200-native dense RGB-D direct-alignment path.
 *
 * Drop-in boundary for phildue/vslam's aligner (paths relative to
 * /root/reference/src/vslam/src/):
 *   - odometry/src/AlignmentSE3.h:22-26            AlignmentSE3::align(from, to)
 *   - odometry/src/direct_image_alignment/SE3Alignment.h:33-39   SE3Alignment(minGradient, solver, loss,
 *     includePrior); align(from, to); align(vector<from>, to)
 *   - core/src/Frame.h:35-37, Frame.cpp:215-275     Frame ctor, computePyramid, computeDerivatives, computePcl
 *   - least_squares/src/{GaussNewton.h:35, Loss.h:54-116, Scaler.h:30-74}  solver / loss / scaler parameters
 *
 * Plain pointers and sizes only; every function returns 0 on success or a negative
 * VSLAMB200_E* code, with a message available from vslamb200_last_error().  There is no CPU
 * fallback: creation fails if no CUDA device is usable.
 */
#ifndef VSLAM_B200_H_
#define VSLAM_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VSLAMB200_MAX_LEVELS 8
#define VSLAMB200_TRACE_STRIDE 64

/* error codes */
#define VSLAMB200_OK 0
#define VSLAMB200_EINVAL (-1)      /* bad argument */
#define VSLAMB200_ECUDA (-2)       /* CUDA runtime failure */
#define VSLAMB200_ENOFRAME (-3)    /* unknown frame id */
#define VSLAMB200_EDIMS (-4)       /* image / depth / camera inconsistent (Frame.cpp:150-155) or pyramid sizes disagree */

/* least_squares::Loss family (Loss.h:54-116); NONE == loss nullptr (InverseCompositional.cpp:129) */
enum { VSLAMB200_LOSS_NONE = 0, VSLAMB200_LOSS_QUADRATIC = 1, VSLAMB200_LOSS_TUKEY = 2, VSLAMB200_LOSS_HUBER = 3, VSLAMB200_LOSS_TDIST = 4 };
/* least_squares::Scaler family (Scaler.h:30-74) */
enum { VSLAMB200_SCALER_IDENTITY = 0, VSLAMB200_SCALER_MEDIAN = 1, VSLAMB200_SCALER_MEAN = 2, VSLAMB200_SCALER_TDIST = 3 };
/* GaussNewton::solve stop reasons per level (GaussNewton.cpp:53-98) */
enum { VSLAMB200_STOP_MAXITER = 0, VSLAMB200_STOP_CONSTRAINTS = 1, VSLAMB200_STOP_HESSIAN = 2, VSLAMB200_STOP_CHI2 = 3, VSLAMB200_STOP_CONVERGED = 4, VSLAMB200_STOP_NAN = 5 };
/* host image layouts / depth element types accepted by frame upload */
enum { VSLAMB200_ROW_MAJOR = 0, VSLAMB200_COL_MAJOR = 1 };   /* COL_MAJOR == Eigen default (core/src/types.h:33-35) */
enum { VSLAMB200_DEPTH_F32 = 0, VSLAMB200_DEPTH_F64 = 1, VSLAMB200_DEPTH_U16 = 2 /* frames_ingest_rgbd only */ };
enum { VSLAMB200_RGB = 0, VSLAMB200_BGR = 1 };
/* solver_kind: Gauss-Newton is the reference's solver (least_squares/src/GaussNewton.cpp); Levenberg-Marquardt is an
 * EXTENSION named by the benchmark configurations (no counterpart in the reference; defined by oracle/vslam_oracle.cpp) */
enum { VSLAMB200_SOLVER_GN = 0, VSLAMB200_SOLVER_LM = 1 };
/* sampling: NEAREST is the reference's aligner (inverse compositional, InverseCompositional.cpp:119-120).  BILINEAR is an
 * EXTENSION (forward formulation: bilinear sampling, Jacobians at the warped point every iteration, optional depth
 * residual; semantics of lukas_kanade/src/ForwardAdditive.cpp:52-137, defined by oracle/vslam_oracle.cpp) */
enum { VSLAMB200_SAMPLING_NEAREST = 0, VSLAMB200_SAMPLING_BILINEAR = 1 };
/* chi2 of one view's normal equations (GaussNewton.cpp:55-61 stops a level when it rises, :86 when it stalls).
 * REFERENCE (0, the default of a zeroed config): w_0 * sum_k r_k^2 - what NormalEquations.cpp:58,
 *   `_chi2 = (r * w).transpose() * r` with r and w both column vectors, evaluates to in the reference's Release build:
 *   Eigen only asserts a product's inner dimension (NDEBUG removes the assertion) and takes it from the left factor,
 *   1 here, so (r * w)(k) = r(k) * w(0), the weight of the FIRST interest point (row-major order; 0 while that point
 *   projects outside the image) times the unweighted sum of squared residuals.  This is what oracle/_ref - the
 *   reference's own sources compiled against the stand-in headers of oracle/ref_shim - computes; whenever w_0 = 0 every
 *   chi2 is 0 and each level ends "converged" after its second step.
 * WEIGHTED (1): sum_k w_k r_k^2, the r^T W r that NormalEquations.h documents.
 * The bilinear extension always uses the weighted sum. */
enum { VSLAMB200_CHI2_REFERENCE = 0, VSLAMB200_CHI2_WEIGHTED = 1 };
/* level_kernel: which Gauss-Newton level kernel runs the nearest-sampling path.  AUTO: the dense kernel
 * (csrc/gn_level.cu) for jobs with one keyframe, the moment-based kernel (csrc/gn_mom.cu: keyframes side by side) for
 * stacked jobs where it applies - scalers without offset (identity, t-distribution), at most 8 keyframes per job, the
 * same number in every job of the call; DENSE: always the dense kernel; MOMENTS: the moment kernel wherever it applies
 * (also with one keyframe).  Results agree to rounding (the sums are formed in another order). */
enum { VSLAMB200_KERNEL_AUTO = 0, VSLAMB200_KERNEL_DENSE = 1, VSLAMB200_KERNEL_MOMENTS = 2 };

typedef struct vslamb200_ctx vslamb200_ctx;

/* SE3Alignment ctor arguments + GaussNewton(minStepSize, maxIterations) + loss / scaler parameters */
typedef struct vslamb200_config
{
  double min_gradient;   /* SE3Alignment.h:33; compared squared against dIx^2+dIy^2 (SE3Alignment.cpp:54,89) */
  int loss_kind;         /* VSLAMB200_LOSS_* */
  double huber_c;        /* HuberLoss::_c (Loss.h:85) */
  double loss_v;         /* LossTDistribution nu (Loss.h:103) */
  int scaler_kind;       /* VSLAMB200_SCALER_* */
  double scaler_v;       /* ScalerTDistribution nu (Scaler.h:61) */
  double min_step_size;  /* GaussNewton.cpp:24-29: also min gradient and min chi2 reduction */
  int max_iterations;    /* GaussNewton.h:35 */
  int include_prior;     /* MotionPrior (SE3Alignment.cpp:27-52) */
  int solver_kind;       /* VSLAMB200_SOLVER_* */
  double lm_lambda0;     /* LM only: initial damping of (A + lambda diag A) dx = b */
  double lm_up;          /* LM only: lambda *= lm_up after a rejected step (> 1) */
  double lm_down;        /* LM only: lambda *= lm_down after an accepted step (0 < lm_down <= 1) */
  int sampling;          /* VSLAMB200_SAMPLING_* */
  double depth_weight;   /* BILINEAR only: weight of the depth residual Z_to(u',v') - q_z against the photometric one; 0 = off */
  int chi2_mode;         /* VSLAMB200_CHI2_*: the chi2 of a view's normal equations that drives the solver's stops (NEAREST only) */
  /* execution choices (results do not depend on them beyond rounding); 0 = automatic */
  int level_kernel;      /* VSLAMB200_KERNEL_*: which Gauss-Newton level kernel runs the nearest-sampling path */
  int cluster_size;      /* CTAs per alignment job of the level kernels (1, 2, 4, 8 or 16) */
  int stream_parts;      /* a batch is cut into this many parts that walk the pyramid on their own streams (1..8) */
} vslamb200_config;

/* One alignment: SE3Alignment::align(from, to) (n_from == 1, multi == 0) or the vector overload (multi != 0).
 * Poses are Sophus SE3d::data() order: quaternion x,y,z,w then translation (P_frame = pose * P_world, Frame.h:94). */
typedef struct vslamb200_job
{
  uint64_t to;                 /* frame id of the target ("to") frame */
  double to_pose[7];           /* to->pose(): the initial guess */
  int n_from;                  /* number of keyframes (1 for the single-frame overload) */
  int multi;                   /* 0: SE3Alignment.cpp:64-105; 1: SE3Alignment.cpp:106-146 */
  const uint64_t * from;       /* n_from frame ids */
  const double * from_poses;   /* n_from x 7 */
} vslamb200_job;

typedef struct vslamb200_result
{
  double pose[7];                      /* aligned pose of the "to" frame */
  double cov[36];                      /* inverse of the last normal-equation matrix (SE3Alignment.cpp:101) */
  int status;                          /* 0 ok; 1: a level ended before one Gauss-Newton step completed
                                          (reference: out_of_range / UB) -> input pose, identity covariance */
  int iters[VSLAMB200_MAX_LEVELS];     /* completed Gauss-Newton steps per level */
  int stop[VSLAMB200_MAX_LEVELS];      /* VSLAMB200_STOP_* per level */
} vslamb200_result;

/* counters of the work done since creation / the last reset (bench roofline accounting) */
typedef struct vslamb200_stats
{
  uint64_t kernel_launches;      /* kernels of this library launched */
  uint64_t pixel_iterations;     /* sum over views, levels, Gauss-Newton iterations of W_l*H_l */
  uint64_t point_iterations;     /* same, counting only retained interest points */
  uint64_t gn_iterations;        /* Gauss-Newton iterations over all jobs and levels */
  uint64_t frames_uploaded;
  uint64_t h2d_bytes, d2h_bytes;
  /* filled only when timing is enabled: interest-list construction and the Gauss-Newton level kernel per level */
  double ms_lists, ms_preprocess, ms_gn_level[VSLAMB200_MAX_LEVELS];
  uint64_t pixel_iterations_level[VSLAMB200_MAX_LEVELS];  /* pixel_iterations split by pyramid level */
  /* SM clocks of the jobs' leading CTA per level, summed over jobs.  Dense kernel: phase A, histogram barrier, scale +
     weights, phase C, block reduction, sums barrier, solve, then the scale section split: histogram gather + scaler,
     publish + second cluster barrier, weight table.  Moment kernel: warp + sample + moment moves (evaluations after the
     first), histogram barrier, scale + weights, A / b from the moments, sums barrier, cluster sums, solve, first
     evaluation's warp + sample, moment build, and (a count, not clocks) points whose residual changed */
  uint64_t phase_clocks[VSLAMB200_MAX_LEVELS][10];
  uint64_t moment_launches;      /* level launches that ran the moment-based kernel */
  uint64_t dense_launches;       /* level launches that ran the dense kernel */
} vslamb200_stats;

const char * vslamb200_last_error(void);
int vslamb200_version(void);

int vslamb200_create(const vslamb200_config * cfg, int device, vslamb200_ctx ** out);
void vslamb200_destroy(vslamb200_ctx * ctx);
int vslamb200_set_config(vslamb200_ctx * ctx, const vslamb200_config * cfg);
/* run on a caller-owned cudaStream_t (passed as void*); NULL selects the context's own stream */
int vslamb200_set_stream(vslamb200_ctx * ctx, void * cuda_stream);

/* Frame ctor + computePyramid(n_levels) + computeDerivatives() + computePcl() (node order,
 * src/ros/nodes/NodeRgbdAlignment.cpp:194-197) for n frames of identical size.  gray / depth are HOST
 * pointers to n consecutive images (gray_stride / depth_stride = bytes between consecutive frames; 0 =
 * tightly packed).  cams: n x 4 (fx, fy, cx, cy).  A frame id that already exists is replaced. */
int vslamb200_frames_upload(
  vslamb200_ctx * ctx, int n, const uint64_t * ids, int W, int H, const uint8_t * gray, size_t gray_stride,
  const void * depth, int depth_type, size_t depth_stride, int layout, const double * cams, int n_levels);
/* Same as vslamb200_frames_upload, but returns as soon as the work is queued (row-major u8 + f32 depth only; other
 * layouts take the synchronous path): the host->device copies run on an internal copy stream and overlap whatever the
 * context computes next; pyramids and gradients are queued behind them.  The host buffers should be pinned and must stay
 * untouched until the next vslamb200_align / vslamb200_synchronize on this context.  Lets a caller pipeline
 * "upload sub-batch k+1" under "align sub-batch k". */
int vslamb200_frames_upload_async(
  vslamb200_ctx * ctx, int n, const uint64_t * ids, int width, int height, const uint8_t * gray, size_t gray_stride,
  const void * depth, int depth_type, size_t depth_stride, int layout, const double * cams, int n_levels);
/* vslamb200_frames_upload_async for what an RGB-D sensor delivers: row-major 8-bit grey and 16-bit depth, tightly packed,
 * depth = value / depth_divisor in double (loadDepth(...) / 5000.0, odometry/test/test_alignment_se3.cpp:99; 0 = no
 * measurement).  3 bytes per pixel cross the bus instead of 5; the conversion runs on the device behind the copy. */
int vslamb200_frames_upload_u16_async(
  vslamb200_ctx * ctx, int n, const uint64_t * ids, int width, int height, const uint8_t * gray, const uint16_t * depth,
  double depth_divisor, const double * cams, int n_levels);
/* same as vslamb200_frames_upload, gray / depth already resident on the context's device (row-major u8 / f32) */
int vslamb200_frames_upload_device(
  vslamb200_ctx * ctx, int n, const uint64_t * ids, int W, int H, const uint8_t * d_gray, const float * d_depth,
  const double * cams, int n_levels);
/* Zero-copy variant of vslamb200_frames_upload_device: the level-0 planes of the n frames ARE the caller's device buffers
 * (tightly packed, 16-byte aligned: frame i at d_gray + i*W*H and d_depth + i*W*H), read in place by the pyramid, gradient,
 * list and level kernels; coarser levels and gradients live in the library's arena.  The buffers must stay valid and
 * unchanged until the frames are released or replaced.  No counterpart in the reference, whose Frame copies its inputs
 * (core/src/Frame.cpp:131-142); it saves the device-to-device copy of 5 bytes per pixel when a producer (decoder,
 * simulator, another kernel) already left the images in HBM. */
int vslamb200_frames_attach_device(
  vslamb200_ctx * ctx, int n, const uint64_t * ids, int W, int H, const uint8_t * d_gray, const float * d_depth,
  const double * cams, int n_levels);
/* Frame ingestion as the node does it, fused into the upload (src/ros/nodes/NodeRgbdAlignment.cpp:177-199, replaces
 * cv::cvtColor(RGB2GRAY) :182, the NaN -> 0 select :190 and, for 16-bit sensor depth, loadDepth(...) / 5000.0 of
 * odometry/test/test_alignment_se3.cpp:99), followed by what vslamb200_frames_upload does.  rgb: n row-major H x W x 3
 * interleaved 8-bit images in channel_order VSLAMB200_RGB / VSLAMB200_BGR; depth: n row-major planes of depth_type
 * (F32, F64 or U16); every depth value is divided by depth_divisor in double (1.0: already metres) and NaN becomes 0. */
int vslamb200_frames_ingest_rgbd(
  vslamb200_ctx * ctx, int n, const uint64_t * ids, int W, int H, const uint8_t * rgb, int channel_order,
  const void * depth, int depth_type, double depth_divisor, const double * cams, int n_levels);
int vslamb200_frame_release(vslamb200_ctx * ctx, uint64_t id);
int vslamb200_frame_release_all(vslamb200_ctx * ctx);
/* dims4 = intensity W, H, depth W, H at that level (they differ at odd sizes: Frame.cpp:264 vs algorithm.h:93-94) */
int vslamb200_frame_dims(vslamb200_ctx * ctx, uint64_t id, int level, int * dims4);
/* copy one pyramid level back (any pointer may be NULL): intensity(l), depth(l), single-frame masked depth,
 * dIx(l), dIy(l) as row-major planes */
int vslamb200_frame_download(
  vslamb200_ctx * ctx, uint64_t id, int level, uint8_t * intensity, float * depth, float * depth_sf, int16_t * dIx,
  int16_t * dIy);

/* SE3Alignment::align for n_jobs independent alignments in one batch */
int vslamb200_align(vslamb200_ctx * ctx, const vslamb200_job * jobs, int n_jobs, vslamb200_result * out);

/* Solver::Results history (Solver.h:49-60): when enabled, align() records per job / level / iteration a
 * 64-double record: 0 valid, 1 nConstraints, 2 chi2, 3 stepSize, 4 scale, 5 offset, 6..11 x, 12..17 b,
 * 18..53 A (row major), 54..59 dx.  get_trace copies job's records: n_levels * max_iterations * 64 doubles. */
int vslamb200_set_trace(vslamb200_ctx * ctx, int enable);
int vslamb200_get_trace(vslamb200_ctx * ctx, int job, double * out);

/* One InverseCompositional::computeNormalEquations (InverseCompositional.cpp:105-154) of keyframe `from`
 * against `to` on `level` at the relative transform se3 (pose7), without prior.  rcode: W*H int16 plane
 * (-32768 not an interest point, -32767 interest point not visible, else the residual); ne64: n, chi2,
 * scale, offset, b[6], A[36] (normalised by n). */
int vslamb200_evaluate(
  vslamb200_ctx * ctx, uint64_t from, uint64_t to, int level, const double * se3_pose7, int multi_mask,
  int16_t * rcode, double * ne64);

int vslamb200_get_stats(vslamb200_ctx * ctx, vslamb200_stats * out);
int vslamb200_reset_stats(vslamb200_ctx * ctx);
/* 1: bracket kernel groups with CUDA events so get_stats reports per-kernel milliseconds (adds syncs) */
int vslamb200_set_timing(vslamb200_ctx * ctx, int enable);
/* block until everything queued on the context's stream has finished */
int vslamb200_synchronize(vslamb200_ctx * ctx);

/* NO COUNTERPART IN THE REFERENCE - added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).  The interest-point list of ONE
 * pyramid level - the candidate mask of SE3Alignment.cpp:83-94 / :123-130, dIx^2 + dIy^2 >= min_gradient^2 and depth > 0
 * and finite, compacted in row-major order exactly like the lists vslamb200_align builds internally - computed by a
 * single kernel that allocates NOTHING: its scratch (one record per 2048 pixels) is a static device array inside the
 * library sized for VSLAMB200_NOALLOC_MAX_PIXELS, about 33 KB per GPU held until exit; larger images are rejected with
 * VSLAMB200_EINVAL.  All pointers are DEVICE pointers on the current device: d_dIx, d_dIy (int16) and d_depth (float)
 * are row-major width x height planes (e.g. what vslamb200_frame_download returns, copied back to the device; pass
 * depth_sf for the single-frame overload, depth for the multi-frame one); d_points receives up to width * height pixel
 * indices v * width + u, d_count their number.  The work is queued on cuda_stream (a cudaStream_t passed as void*; NULL =
 * the default stream) and is complete when that stream reaches this point; the function does not synchronise.  Every call
 * shares the one scratch array and clears the part it uses first: calls that overlap on the GPU (two streams, two host
 * threads) race - serialise them.  The first call on a device resolves the array's address (cudaGetSymbolAddress); make
 * that call outside a graph capture.  Errors: vslamb200_last_error(). */
#define VSLAMB200_NOALLOC_MAX_PIXELS (4096 * 4096)
int vslamb200_interest_points_noalloc(
  const int16_t * d_dIx, const int16_t * d_dIy, const float * d_depth, int width, int height, double min_gradient,
  uint32_t * d_points, int * d_count, void * cuda_stream);

#ifdef __cplusplus
}
#endif
#endif /* VSLAM_B200_H_ */
