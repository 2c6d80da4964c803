// Accumulate kernel (Jacobian, weights, 21+6+1 products, deterministic reduction, Gauss-Newton step)
// plus the small per-level bookkeeping kernels.
#include "solve.cuh"

namespace vb
{
constexpr int THREADS = 256;

template <int PPT>
__global__ void __launch_bounds__(THREADS, 3) accumulate_kernel(
  JobState * __restrict__ jobs, ViewState * __restrict__ views, const int * __restrict__ activeViews,
  double * __restrict__ partials, int partialStride, SolverParams p, int level, double * __restrict__ trace,
  size_t traceJobStride)
{
  const int viewIdx = activeViews[blockIdx.y];
  ViewState & V = views[viewIdx];
  if ((int)blockIdx.x >= V.nChunks) return;
  __shared__ float lut[HIST_BINS];
  __shared__ float sacc[NE_TERMS][THREADS];
  __shared__ bool isLast;
  const int tid = threadIdx.x;
  const bool weighted = p.lossKind != VSLAMB200_LOSS_NONE;
  if (weighted) {
    lut[tid] = __ldcg(V.wlut + tid);
    lut[tid + 256] = __ldcg(V.wlut + tid + 256);
  }
  __syncthreads();

  const double ifx = V.ifx, ify = V.ify, kx = V.kx, ky = V.ky, fxJ = V.fxJd, fyJ = V.fyJd;
  const uint32_t * __restrict__ uvt = V.uvt;
  const float * __restrict__ dep = V.d;
  const uint32_t * __restrict__ grad = V.g;
  const int16_t * __restrict__ res = V.r;
  const int nIp = V.nIp;
  const int base = blockIdx.x * (THREADS * PPT);

  float acc[NE_TERMS];
#pragma unroll
  for (int s = 0; s < NE_TERMS; s++) acc[s] = 0.f;

#pragma unroll 4
  for (int k = 0; k < PPT; k++) {
    const int idx = base + k * THREADS + tid;
    if (idx < nIp) {
      const int rr = res[idx];
      if (rr != R_INVISIBLE) {
        const uint32_t pk = __ldg(uvt + idx);
        const float d = __ldg(dep + idx);
        const uint32_t gg = __ldg(grad + idx);
        // WarpSE3::J rows (Warp.cpp:184-199) contracted with the template gradient
        // (InverseCompositional.cpp:79-80), in normalised coordinates xn = x/z, yn = y/z.  The six
        // entries are formed in fp64 and only then rounded to fp32: a coherent fp32 bias in xn / yn
        // would be amplified by the cancellation in b = sum(w J r) near convergence.
        const double u = (double)(pk & 0xfff), v = (double)((pk >> 12) & 0xfff);
        const double a = (double)(int16_t)(gg & 0xffff) * fxJ;   // dIx * fx
        const double b = (double)(int16_t)(gg >> 16) * fyJ;      // dIy * fy
        const float w = weighted ? lut[rr + 255] : 1.f;
        const float r = (float)rr;
        const double zi = 1.0 / (double)d;
        const double xn = fma(ifx, u, kx), yn = fma(ify, v, ky);
        float J[6];
        J[0] = (float)(a * zi);
        J[1] = (float)(b * zi);
        J[2] = (float)(-(a * xn + b * yn) * zi);
        J[3] = (float)(-a * xn * yn - b * (1.0 + yn * yn));
        J[4] = (float)(a * (1.0 + xn * xn) + b * xn * yn);
        J[5] = (float)(b * xn - a * yn);
        int s = 0;
#pragma unroll
        for (int i = 0; i < 6; i++) {
          const float wj = w * J[i];
#pragma unroll
          for (int j = i; j < 6; j++) {
            acc[s] = fmaf(wj, J[j], acc[s]);
            s++;
          }
          acc[21 + i] = fmaf(wj, r, acc[21 + i]);
        }
        acc[27] = fmaf(w * r, r, acc[27]);
      }
    }
  }
  // block reduction in a fixed order: smem transpose, then warp w owns terms w, w+8, w+16, w+24
#pragma unroll
  for (int s = 0; s < NE_TERMS; s++) sacc[s][tid] = acc[s];
  __syncthreads();
  const int lane = tid & 31, warp = tid >> 5;
  double * out = partials + ((size_t)viewIdx * partialStride + blockIdx.x) * NE_TERMS;
  for (int s = warp; s < NE_TERMS; s += THREADS / 32) {
    double v = 0.0;
#pragma unroll
    for (int k = 0; k < THREADS / 32; k++) v += (double)sacc[s][lane + 32 * k];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) out[s] = v;
  }
  // last block of the job: fixed-order grid reduction + solve
  JobState & J = jobs[V.job];
  __threadfence();
  __syncthreads();
  if (tid == 0) isLast = atomicAdd(&J.ticketB, 1u) == (unsigned)(J.totalChunks - 1);
  __syncthreads();
  if (!isLast) return;
  __threadfence();
  double * sums = reinterpret_cast<double *>(&sacc[0][0]);           // MAX_VIEWS*28 doubles
  double * part = sums + MAX_VIEWS * NE_TERMS;                        // 8 x 32 doubles
  for (int k = 0; k < J.nViews; k++) {
    const int vi = J.view0 + k;
    const int nC = views[vi].nChunks;
    const int per = (nC + 7) / 8;
    const int c0 = warp * per, c1 = min(nC, c0 + per);
    double v = 0.0;
    if (lane < NE_TERMS) {
      const double * src = partials + (size_t)vi * partialStride * NE_TERMS + lane;
      for (int c = c0; c < c1; c++) v += __ldcg(src + (size_t)c * NE_TERMS);
    }
    __syncthreads();
    part[warp * 32 + lane] = v;
    __syncthreads();
    if (tid < NE_TERMS) {
      double tsum = 0.0;
#pragma unroll
      for (int g = 0; g < 8; g++) tsum += part[g * 32 + tid];
      sums[k * NE_TERMS + tid] = tsum;
    }
  }
  __syncthreads();
  if (tid == 0) {
    J.ticketB = 0;
    gn_step(J, views, sums, p, level, trace ? trace + (size_t)V.job * traceJobStride : nullptr);
  }
}

void launch_accumulate(JobState * jobs, ViewState * views, const int * activeViews, int nActiveViews, int maxChunks,
                       int chunk, double * partials, int partialStride, SolverParams p, int level, double * trace,
                       size_t traceJobStride, cudaStream_t s)
{
  if (nActiveViews <= 0 || maxChunks <= 0) return;
  dim3 grid(maxChunks, nActiveViews);
  switch (chunk / THREADS) {
    case 2: accumulate_kernel<2><<<grid, THREADS, 0, s>>>(jobs, views, activeViews, partials, partialStride, p, level, trace, traceJobStride); break;
    case 4: accumulate_kernel<4><<<grid, THREADS, 0, s>>>(jobs, views, activeViews, partials, partialStride, p, level, trace, traceJobStride); break;
    case 8: accumulate_kernel<8><<<grid, THREADS, 0, s>>>(jobs, views, activeViews, partials, partialStride, p, level, trace, traceJobStride); break;
    default: accumulate_kernel<16><<<grid, THREADS, 0, s>>>(jobs, views, activeViews, partials, partialStride, p, level, trace, traceJobStride); break;
  }
}

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
  const FrameDev * __restrict__ frames, JobState * __restrict__ jobs, ViewState * __restrict__ views, int nViews,
  int level)
{
  const int vi = blockIdx.x * blockDim.x + threadIdx.x;
  if (vi >= nViews) return;
  ViewState & V = views[vi];
  JobState & J = jobs[V.job];
  const FrameDev & ff = frames[V.fromSlot];
  const FrameDev & ft = frames[V.toSlot];
  V.W = ff.W[level];
  V.H = ff.H[level];
  V.Wt = ft.W[level];
  V.Ht = ft.H[level];
  V.I = ft.I[level];
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
  V.nChunks = 0;
  V.ticketA = 0;
  V.scale = 1.0;
  V.offset = 0.0;
  V.se3 = se3_mul(J.pose, se3_inv(V.poseRef));
  view_set_warp(V);
  se3_log(se3_mul(J.toPose, se3_inv(V.poseRef)), V.xPred);  // MotionPrior ctor (SE3Alignment.cpp:32)
  if (vi == J.view0) {
    J.iter = 0;
    J.nPushed = 0;
    J.chi2Prev = 0.0;
    J.ticketB = 0;
    J.active = J.dead ? 0 : 1;
    for (int k = 0; k < 6; k++) J.xPrev[k] = 0.0;
  }
}

// one thread per job, after the lists are built: total chunk count; a job without a single interest
// point ends the level at once (NormalEquations all zero, n = 0 < 6: GaussNewton.cpp:53-58)
__global__ void level_begin_kernel(JobState * __restrict__ jobs, ViewState * __restrict__ views, int nJobs, int level)
{
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nJobs) return;
  JobState & J = jobs[j];
  int tot = 0;
  for (int k = 0; k < J.nViews; k++) tot += views[J.view0 + k].nChunks;
  J.totalChunks = tot;
  if (J.active && tot == 0) finish_level(J, views, level, VSLAMB200_STOP_CONSTRAINTS);
}

void launch_level_setup(const FrameDev * frames, JobState * jobs, ViewState * views, int nViews, int level,
                        cudaStream_t s)
{
  level_setup_kernel<<<(nViews + 127) / 128, 128, 0, s>>>(frames, jobs, views, nViews, level);
}
void launch_level_begin(JobState * jobs, ViewState * views, int nJobs, int level, cudaStream_t s)
{
  level_begin_kernel<<<(nJobs + 127) / 128, 128, 0, s>>>(jobs, views, nJobs, level);
}

// single block: ordered list of the views whose job is still iterating.
// counts: [0] active views, [1] active jobs, [2] max nChunks over active views, [3] sum nIp (low), [4] sum nIp (high)
__global__ void __launch_bounds__(1024) compact_active_kernel(
  const JobState * __restrict__ jobs, const ViewState * __restrict__ views, int nViews, int * __restrict__ activeViews,
  int * __restrict__ counts)
{
  __shared__ int wsum[32];
  __shared__ int carry;
  __shared__ int s_jobs, s_maxc;
  __shared__ unsigned long long s_ip;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) {
    carry = 0;
    s_jobs = 0;
    s_maxc = 0;
    s_ip = 0ull;
  }
  __syncthreads();
  for (int b = 0; b < nViews; b += 1024) {
    const int i = b + tid;
    bool act = false;
    if (i < nViews) {
      const JobState & J = jobs[views[i].job];
      act = J.active != 0;
      if (act) {
        if (i == J.view0) atomicAdd(&s_jobs, 1);
        atomicMax(&s_maxc, views[i].nChunks);
        atomicAdd(&s_ip, (unsigned long long)views[i].nIp);
      }
    }
    const unsigned m = __ballot_sync(0xffffffffu, act);
    if (lane == 0) wsum[warp] = __popc(m);
    __syncthreads();
    int woff = 0, tot = 0;
    for (int w = 0; w < 32; w++) {
      if (w < warp) woff += wsum[w];
      tot += wsum[w];
    }
    if (act) activeViews[carry + woff + __popc(m & ((1u << lane) - 1u))] = i;
    __syncthreads();
    if (tid == 0) carry += tot;
    __syncthreads();
  }
  if (tid == 0) {
    counts[0] = carry;
    counts[1] = s_jobs;
    counts[2] = s_maxc;
    counts[3] = (int)(s_ip & 0xffffffffull);
    counts[4] = (int)(s_ip >> 32);
  }
}

void launch_compact_active(const JobState * jobs, const ViewState * views, int nViews, int * activeViews, int * counts,
                           cudaStream_t s)
{
  compact_active_kernel<<<1, 1024, 0, s>>>(jobs, views, nViews, activeViews, counts);
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
    const uint32_t pk = V.uvt[i];
    plane[(size_t)((pk >> 12) & 0xfff) * V.W + (pk & 0xfff)] = V.r[i];
  }
}
void launch_scatter_rcode(const ViewState * views, int view, int16_t * plane, cudaStream_t s)
{
  scatter_rcode_kernel<<<296, 256, 0, s>>>(views, view, plane);
  scatter_rcode_kernel2<<<296, 256, 0, s>>>(views, view, plane);
}
}  // namespace vb
