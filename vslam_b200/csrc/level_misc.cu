// Small per-level bookkeeping kernels around the Gauss-Newton level kernel (gn_level.cu).
#include "solve.cuh"

namespace vb
{
// ------------------------------------------------------------------------------------------------
// per-level bookkeeping
// ------------------------------------------------------------------------------------------------
__device__ inline void split_hi_lo(double v, float & hi, float & lo)
{
  hi = (float)v;
  lo = (float)(v - (double)hi);
}

// one thread per view: static fields of the level, warp = pose * poseRef^-1 (Warp.cpp:135-147), prior
__global__ void level_setup_kernel(
  const FrameDev * __restrict__ frames, JobState * __restrict__ jobs, ViewState * __restrict__ views, int view0,
  int nViews, int level)
{
  const int vi = blockIdx.x * blockDim.x + threadIdx.x;
  if (vi >= nViews) return;
  ViewState & V = views[view0 + vi];
  JobState & J = jobs[V.job];
  const FrameDev & ff = frames[V.fromSlot];
  const FrameDev & ft = frames[V.toSlot];
  V.W = ff.W[level];
  V.H = ff.H[level];
  V.Wt = ft.W[level];
  V.Ht = ft.H[level];
  V.I = ft.I[level];
  V.Zt = ft.Z[level];
  V.gxt = ft.gx[level];
  V.gyt = ft.gy[level];
  V.scaleZ = 1.0;
  V.fx = ff.cam[level][0];
  V.fy = ff.cam[level][1];
  V.cx = ff.cam[level][2];
  V.cy = ff.cam[level][3];
  V.ifx = 1.0 / V.fx;
  V.ify = 1.0 / V.fy;
  V.kx = -V.cx / V.fx;
  V.ky = -V.cy / V.fy;
  V.fxJd = ft.cam[level][0];
  V.fyJd = ft.cam[level][1];
  split_hi_lo(V.ifx, V.ifxH, V.ifxL);
  split_hi_lo(V.ify, V.ifyH, V.ifyL);
  split_hi_lo(V.kx, V.kxH, V.kxL);
  split_hi_lo(V.ky, V.kyH, V.kyL);
  split_hi_lo(V.fxJd, V.fxJH, V.fxJL);
  split_hi_lo(V.fyJd, V.fyJH, V.fyJL);
  V.fx32 = (float)V.fx;
  V.fy32 = (float)V.fy;
  V.nIp = 0;
  V.listTicket = 0;
  for (int i = 0; i < JSTAT; i++) V.jstat[i] = 0u;
  V.scale = 1.0;
  V.offset = 0.0;
  V.se3 = se3_mul(J.pose, se3_inv(V.poseRef));
  view_set_warp(V);
  view_set_x(V);
  se3_log(se3_mul(J.toPose, se3_inv(V.poseRef)), V.xPred);  // MotionPrior ctor (SE3Alignment.cpp:32)
  if (view0 + vi == J.view0) {
    J.iter = 0;
    J.nPushed = 0;
    J.chi2Prev = 0.0;
    J.accepted = 0;
    J.chi2Acc = 0.0;
    J.active = J.dead ? 0 : 1;
    for (int k = 0; k < 6; k++) J.xPrev[k] = 0.0;
    J.evals[level] = 0;
    for (int k = 0; k < 10; k++) J.clk[level][k] = 0;
  }
}

// one thread per job, after the lists are built: total chunk count; a job without a single interest
// point ends the level at once (NormalEquations all zero, n = 0 < 6: GaussNewton.cpp:53-58)
__global__ void level_begin_kernel(
  JobState * __restrict__ jobs, ViewState * __restrict__ views, int job0, int nJobs, int level)
{
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nJobs) return;
  JobState & J = jobs[job0 + j];
  int tot = 0;
  for (int k = 0; k < J.nViews; k++) tot += views[J.view0 + k].nIp;
  J.totalPoints = tot;
  if (J.active && tot == 0) {
    J.evals[level]++;
    finish_level(J, views + J.view0, level, VSLAMB200_STOP_CONSTRAINTS);
  }
}

void launch_level_setup(const FrameDev * frames, JobState * jobs, ViewState * views, int view0, int nViews, int level,
                        cudaStream_t s)
{
  level_setup_kernel<<<(nViews + 127) / 128, 128, 0, s>>>(frames, jobs, views, view0, nViews, level);
}
void launch_level_begin(JobState * jobs, ViewState * views, int job0, int nJobs, int level, cudaStream_t s)
{
  level_begin_kernel<<<(nJobs + 127) / 128, 128, 0, s>>>(jobs, views, job0, nJobs, level);
}

// Host -> device hand-over of small control structures without the DMA engine: the source is pinned, mapped host
// memory and the SMs read it over PCIe.  (A cudaMemcpyAsync would queue behind a large frame upload that is in flight
// on the host-to-device copy engine and serialise the alignment with it.)
__global__ void stage_copy_kernel(uint32_t * __restrict__ dst, const uint32_t * __restrict__ src, size_t words)
{
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < words; i += (size_t)gridDim.x * blockDim.x)
    dst[i] = src[i];
}
void launch_stage_copy(void * dst, const void * srcPinned, size_t bytes, cudaStream_t s)
{
  const size_t words = (bytes + 3) / 4;
  const int blocks = (int)std::min<size_t>(296, (words + 255) / 256);
  if (words) stage_copy_kernel<<<blocks, 256, 0, s>>>((uint32_t *)dst, (const uint32_t *)srcPinned, words);
}

// parity helper: residual codes of one view scattered back to a W*H plane
__global__ void scatter_rcode_kernel(const ViewState * __restrict__ views, int view, int16_t * __restrict__ plane)
{
  const ViewState & V = views[view];
  const int npx = V.W * V.H;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < npx; i += gridDim.x * blockDim.x) plane[i] = R_NOT_INTEREST;
}
__global__ void scatter_rcode_kernel2(const ViewState * __restrict__ views, int view, int16_t * __restrict__ plane)
{
  const ViewState & V = views[view];
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < V.nIp; i += gridDim.x * blockDim.x) {
    const uint32_t pk = V.rec[i].uvt;
    const int code = V.r[i];
    plane[(size_t)((pk >> 12) & 0xfff) * V.W + (pk & 0xfff)] = code == R_CODE_INVISIBLE ? R_INVISIBLE : (int16_t)(code - 255);
  }
}
void launch_scatter_rcode(const ViewState * views, int view, int16_t * plane, cudaStream_t s)
{
  scatter_rcode_kernel<<<296, 256, 0, s>>>(views, view, plane);
  scatter_rcode_kernel2<<<296, 256, 0, s>>>(views, view, plane);
}
}  // namespace vb
