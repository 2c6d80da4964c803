// EXTENSION: forward formulation of the alignment - bilinear sampling of the target image, Jacobians
// evaluated at the warped point in every iteration, optional depth residual.  The reference's shipped aligner is
// inverse compositional with nearest-pixel sampling (gn_level.cu); this mode follows the structure of the reference's
// lukas_kanade/src/ForwardAdditive.cpp:52-137 (warp -> sample image and gradient -> J = grad * dW/dx, every iteration)
// and is DEFINED by oracle/vslam_oracle.cpp: view_normal_equations_ext (parity unpinned: no reference golden vectors).
//
// Per view with warp se3 = (R, t), interest point p (back-projected template pixel), q = R p + t, (u', v') = pi(q):
//   r_I = S(I_to)(u', v') - T            S = true bilinear interpolation in fp64
//   g   = S(dIx, dIy)(u', v') / 8        (the Sobel planes carry 8 x the derivative)
//   G_I = g^T dpi/dq = (g_x fx / q_z,  g_y fy / q_z,  -(g_x fx q_x + g_y fy q_y) / q_z^2)
//   J_I = (G_I R,  p x (G_I R))          right perturbation se3 * exp(xi)
//   r_Z = S(Z_to)(u', v') - q_z          only where all four neighbours carry depth
//   J_Z = (G_Z R,  p x (G_Z R)),  G_Z = g_Z^T dpi/dq - (0, 0, 1)
//   A = (sum w_I J_I^T J_I + depthWeight sum w_Z J_Z^T J_Z) / n, b and chi2 alike, n = all interest points of the view
// The residuals are not integers here, so the histogram machinery of the nearest-pixel kernels does not apply: the
// t-distribution scale is a fixed point over the stored fp64 residuals.
//
// One evaluation = three launches over all active jobs of the batch:
//   ext_res_kernel    grid (chunks, views): warp + bilinear samples -> {r_I, r_Z} per list entry (NaN = absent)
//   ext_scale_kernel  one 8-CTA cluster per view: Scaler.cpp:49-67 on r_I and r_Z (sums combined over DSMEM in rank order)
//   ext_acc_kernel    grid (chunks, views): Jacobians, weights, 28 fp64 sums per block; the LAST block of a job to
//                     finish adds the blocks' partial sums in block order and runs the solver step (solve.cuh), so
//                     the result does not depend on the order in which blocks ran.
// This path is HBM/L2-bound like the dense kernel but far from tuned: it exists for parity of the extension row.
#include <cooperative_groups.h>

#include "robust.cuh"
#include "solve.cuh"

namespace cg = cooperative_groups;

namespace vb
{
namespace
{
constexpr int EXT_NT = 256;                   // threads per block of the residual / accumulation kernels
constexpr int EXT_PPT = EXT_CHUNK / EXT_NT;   // list entries per thread
constexpr int EXT_SC = 8;                     // CTAs per view of the scale kernel
constexpr int EXT_SNT = 512;

struct ExtPoint
{
  double X, Y, d;       // p
  double qx, qy, qz;    // R p + t
  double a, b;          // bilinear fractions
  int i00;              // index of the top-left neighbour in the target planes
};

__device__ __forceinline__ double bilinear(double f00, double f10, double f01, double f11, double a, double b)
{
  return (1.0 - b) * ((1.0 - a) * f00 + a * f10) + b * ((1.0 - a) * f01 + a * f11);
}

// image2camera, se3 * p, camera2image and the visibility test, operation for operation as warp_exact() (warp_filter.cuh)
__device__ __forceinline__ bool ext_warp(const ViewState & V, uint32_t pk, float df, ExtPoint & g)
{
  const int u = pk & 0xfff, v = (pk >> 12) & 0xfff;
  const double d = (double)df;
  g.d = d;
  g.X = __fma_rn(V.kx, d, __dmul_rn(V.ifx, __dmul_rn((double)u, d)));
  g.Y = __fma_rn(V.ky, d, __dmul_rn(V.ify, __dmul_rn((double)v, d)));
  g.qx = __fma_rn(V.R[2], d, __fma_rn(V.R[1], g.Y, __fma_rn(V.R[0], g.X, V.t[0])));
  g.qy = __fma_rn(V.R[5], d, __fma_rn(V.R[4], g.Y, __fma_rn(V.R[3], g.X, V.t[1])));
  g.qz = __fma_rn(V.R[8], d, __fma_rn(V.R[7], g.Y, __fma_rn(V.R[6], g.X, V.t[2])));
  if (!(d > 0.0) || !(g.qz > 0.0)) return false;
  const double up = __ddiv_rn(__fma_rn(V.fx, g.qx, __dmul_rn(V.cx, g.qz)), g.qz);
  const double vp = __ddiv_rn(__fma_rn(V.fy, g.qy, __dmul_rn(V.cy, g.qz)), g.qz);
  if (!(1.0 < up && up < (double)(V.Wt - 1) && 1.0 < vp && vp < (double)(V.Ht - 1))) return false;
  const int x0 = __double2int_rd(up), y0 = __double2int_rd(vp);
  g.a = up - (double)x0;
  g.b = vp - (double)y0;
  g.i00 = y0 * V.Wt + x0;
  return true;
}

__device__ __forceinline__ void list_entry(const ViewState & V, int i, uint32_t & pk, float & d)
{
  const unsigned char * tile = V.tiles + (size_t)(i / LIST_TILE) * LIST_TILE_BYTES;
  const int j = i % LIST_TILE;
  d = reinterpret_cast<const float *>(tile + TILE_OFF_D)[j];
  pk = reinterpret_cast<const uint32_t *>(tile + TILE_OFF_UVT)[j];
}

__device__ __forceinline__ bool depth_ok(float z00, float z10, float z01, float z11)
{
  return z00 > 0.f && z10 > 0.f && z01 > 0.f && z11 > 0.f && isfinite((double)z00 + (double)z10 + (double)z01 + (double)z11);
}
}  // namespace

// ------------------------------------------------------------------------------------------------
// residuals of the current warp
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(EXT_NT) ext_res_kernel(
  const JobState * __restrict__ jobs, const ViewState * __restrict__ views, int view0, double depthWeight)
{
  const ViewState & V = views[view0 + blockIdx.y];
  if (!jobs[V.job].active) return;
  const int n = V.nIp, base = blockIdx.x * EXT_CHUNK;
  if (base >= n) return;
  const double NaN = __longlong_as_double(0x7ff8000000000000ll);
#pragma unroll 2
  for (int k = 0; k < EXT_PPT; k++) {
    const int i = base + k * EXT_NT + threadIdx.x;
    if (i >= n) break;
    uint32_t pk;
    float d;
    list_entry(V, i, pk, d);
    ExtPoint g;
    double rI = NaN, rZ = NaN;
    if (ext_warp(V, pk, d, g)) {
      const uint8_t * I = V.I + g.i00;
      rI = bilinear((double)I[0], (double)I[1], (double)I[V.Wt], (double)I[V.Wt + 1], g.a, g.b) - (double)(pk >> 24);
      if (depthWeight > 0.0) {
        const float * Z = V.Zt + g.i00;
        const float z00 = Z[0], z10 = Z[1], z01 = Z[V.Wt], z11 = Z[V.Wt + 1];
        if (depth_ok(z00, z10, z01, z11)) rZ = bilinear((double)z00, (double)z10, (double)z01, (double)z11, g.a, g.b) - g.qz;
      }
    }
    V.extR[i] = make_double2(rI, rZ);
  }
}

// ------------------------------------------------------------------------------------------------
// ScalerTDistribution::compute (Scaler.cpp:49-67) over the stored residuals, photometric and depth side by side:
//   sigma^2 <- mean(r^2 (nu+1) / (nu + (r/sigma)^2)) from sigma = 1, at most 30 steps, until |step| <= 1e-5
// One cluster per view; every CTA sums a strided share, the shares are combined over distributed shared memory in
// rank order, so every CTA carries the same sigma and the result is reproducible.
// ------------------------------------------------------------------------------------------------
__global__ void __cluster_dims__(EXT_SC, 1, 1) __launch_bounds__(EXT_SNT) ext_scale_kernel(
  const JobState * __restrict__ jobs, ViewState * __restrict__ views, int view0, double nu)
{
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned rank = cluster.block_rank();
  ViewState & V = views[view0 + blockIdx.x / EXT_SC];
  if (!jobs[V.job].active) return;  // the same decision in all CTAs of the cluster
  __shared__ double slot[2][4];
  __shared__ double wsum[EXT_SNT / 32][4];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int n = V.nIp;
  const double2 * __restrict__ R = V.extR;
  double xI = 1.0, xZ = 1.0, sI = 1.0, sZ = 1.0, nI = 0.0, nZ = 0.0;
  bool doneI = false, doneZ = false;
  for (int it = 0; it < 30 && !(doneI && doneZ); it++) {
    double v[4] = {0.0, 0.0, 0.0, 0.0};
    const double vI = nu * xI, vZ = nu * xZ, kI = (nu + 1.0) * xI, kZ = (nu + 1.0) * xZ;
    for (int i = rank * EXT_SNT + tid; i < n; i += EXT_SC * EXT_SNT) {
      const double2 r = R[i];
      if (!doneI && r.x == r.x) {
        const double r2 = r.x * r.x;
        v[0] = fma(r2 * kI, rcp_newton(vI + r2), v[0]);
        v[2] += 1.0;
      }
      if (!doneZ && r.y == r.y) {
        const double r2 = r.y * r.y;
        v[1] = fma(r2 * kZ, rcp_newton(vZ + r2), v[1]);
        v[3] += 1.0;
      }
    }
#pragma unroll
    for (int q = 0; q < 4; q++) {
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) v[q] += __shfl_xor_sync(0xffffffffu, v[q], o);
      if (lane == 0) wsum[warp][q] = v[q];
    }
    __syncthreads();
    if (tid < 4) {
      double s = 0.0;
      for (int w = 0; w < EXT_SNT / 32; w++) s += wsum[w][tid];
      slot[it & 1][tid] = s;
    }
    cluster.sync();
    double tot[4] = {0.0, 0.0, 0.0, 0.0};
    for (unsigned rk = 0; rk < EXT_SC; rk++) {
      const double * remote = cluster.map_shared_rank(&slot[it & 1][0], rk);
#pragma unroll
      for (int q = 0; q < 4; q++) tot[q] += remote[q];
    }
    if (it == 0) {
      nI = tot[2];
      nZ = tot[3];
      if (nZ == 0.0) doneZ = true;  // no depth residuals: sigma_Z stays 1 (compute_scale is not called)
    }
    if (!doneI) {
      xI = tot[0] / nI;
      const double s = sqrt(xI), step = fabs(sI - s);
      sI = s;
      doneI = !(step > 1e-5);
    }
    if (!doneZ) {
      xZ = tot[1] / nZ;
      const double s = sqrt(xZ), step = fabs(sZ - s);
      sZ = s;
      doneZ = !(step > 1e-5);
    }
  }
  cluster.sync();  // nobody leaves while its slots may still be read
  if (rank == 0 && tid == 0) {
    V.scale = sI;
    V.offset = 0.0;
    V.scaleZ = sZ;
  }
}

// ------------------------------------------------------------------------------------------------
// Jacobians, weights and the normal equations; the last block of a job runs the solver step
// ------------------------------------------------------------------------------------------------
namespace
{
// rows of J = (H, p x H) with H = G R, added with weight w and residual r to the thread's 28 running sums
__device__ __forceinline__ void ext_add(
  double * acc, const ViewState & V, const ExtPoint & g, double g0, double g1, double g2, double r, double w)
{
  const double * R = V.R;
  const double H0 = g0 * R[0] + g1 * R[3] + g2 * R[6];
  const double H1 = g0 * R[1] + g1 * R[4] + g2 * R[7];
  const double H2 = g0 * R[2] + g1 * R[5] + g2 * R[8];
  double J[6];
  J[0] = H0;
  J[1] = H1;
  J[2] = H2;
  J[3] = g.Y * H2 - g.d * H1;
  J[4] = g.d * H0 - g.X * H2;
  J[5] = g.X * H1 - g.Y * H0;
  int idx = 0;
#pragma unroll
  for (int i = 0; i < 6; i++) {
    const double wj = w * J[i];
#pragma unroll
    for (int j = i; j < 6; j++) acc[idx++] += wj * J[j];
    acc[21 + i] += wj * r;
  }
  acc[27] += r * w * r;
}

__device__ __forceinline__ int ext_blocks_of(int nIp) { return nIp > 0 ? (nIp + EXT_CHUNK - 1) / EXT_CHUNK : 1; }
}  // namespace

__global__ void __launch_bounds__(EXT_NT) ext_acc_kernel(
  JobState * jobs, ViewState * views, int view0, SolverParams p, int level, int forceStop, double * trace,
  size_t traceJobStride)
{
  __shared__ double red[EXT_NT / 32][NE_TERMS];
  __shared__ double sums[MAX_VIEWS * NE_TERMS];
  __shared__ SolveScratch X;
  __shared__ int isLast;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  ViewState & V = views[view0 + blockIdx.y];
  JobState & J = jobs[V.job];
  if (!J.active) return;
  const int n = V.nIp, base = blockIdx.x * EXT_CHUNK;
  if (base >= n && blockIdx.x > 0) return;  // block 0 of every view always takes part (a view may be empty)
  double acc[NE_TERMS];
#pragma unroll
  for (int t = 0; t < NE_TERMS; t++) acc[t] = 0.0;
  const bool robust = p.lossKind != VSLAMB200_LOSS_NONE;
  const double fx = V.fx, fy = V.fy, isI = 1.0 / V.scale, isZ = 1.0 / V.scaleZ;
  for (int k = 0; k < EXT_PPT; k++) {
    const int i = base + k * EXT_NT + tid;
    if (i >= n) break;
    const double2 r = V.extR[i];
    if (!(r.x == r.x)) continue;  // not visible (then there is no depth residual either)
    uint32_t pk;
    float d;
    list_entry(V, i, pk, d);
    ExtPoint g;
    ext_warp(V, pk, d, g);
    const double iz = 1.0 / g.qz;
    {
      const int16_t * gxp = V.gxt + g.i00;
      const int16_t * gyp = V.gyt + g.i00;
      const double gx = bilinear((double)gxp[0], (double)gxp[1], (double)gxp[V.Wt], (double)gxp[V.Wt + 1], g.a, g.b) / 8.0;
      const double gy = bilinear((double)gyp[0], (double)gyp[1], (double)gyp[V.Wt], (double)gyp[V.Wt + 1], g.a, g.b) / 8.0;
      const double w = robust ? loss_weight(p, (r.x - V.offset) * isI) : 1.0;
      ext_add(acc, V, g, gx * fx * iz, gy * fy * iz, -(gx * fx * g.qx + gy * fy * g.qy) * iz * iz, r.x, w);
    }
    if (r.y == r.y) {
      const float * Z = V.Zt + g.i00;
      const double z00 = Z[0], z10 = Z[1], z01 = Z[V.Wt], z11 = Z[V.Wt + 1];
      const double gzx = (1.0 - g.b) * (z10 - z00) + g.b * (z11 - z01), gzy = (1.0 - g.a) * (z01 - z00) + g.a * (z11 - z10);
      const double w = p.depthWeight * (robust ? loss_weight(p, r.y * isZ) : 1.0);
      ext_add(acc, V, g, gzx * fx * iz, gzy * fy * iz, -(gzx * fx * g.qx + gzy * fy * g.qy) * iz * iz - 1.0, r.y, w);
    }
  }
  // block sum in a fixed order: lanes by butterfly, then warps 0..7
#pragma unroll
  for (int t = 0; t < NE_TERMS; t++) {
    double v = acc[t];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0) red[warp][t] = v;
  }
  __syncthreads();
  if (tid < NE_TERMS) {
    double v = 0.0;
#pragma unroll
    for (int w = 0; w < EXT_NT / 32; w++) v += red[w][tid];
    V.extPart[(size_t)blockIdx.x * NE_TERMS + tid] = v;
    __threadfence();
  }
  __syncthreads();
  ViewState * jv = views + J.view0;
  const int K = J.nViews;
  if (tid == 0) {
    int expected = 0;
    for (int k = 0; k < K; k++) expected += ext_blocks_of(jv[k].nIp);
    isLast = atomicAdd(&J.extCount, 1) == expected - 1;
  }
  __syncthreads();
  if (!isLast) return;
  __threadfence();
  for (int i = tid; i < K * NE_TERMS; i += EXT_NT) {
    const int k = i / NE_TERMS, t = i - k * NE_TERMS;
    const int nb = ext_blocks_of(jv[k].nIp);
    const double * part = jv[k].extPart;
    double v = 0.0;
    for (int b = 0; b < nb; b++) v += __ldcg(part + (size_t)b * NE_TERMS + t);
    sums[i] = v;
  }
  if (tid == 0) {
    J.extCount = 0;
    J.evals[level]++;
    J.pointIters += (unsigned long long)J.totalPoints;
  }
  __syncthreads();
  solver_step_block(J, jv, sums, X, p, level, trace ? trace + (size_t)V.job * traceJobStride : nullptr, forceStop != 0);
}

__global__ void ext_count_active_kernel(const JobState * __restrict__ jobs, int job0, int nJobs, int * __restrict__ out)
{
  __shared__ int tot;
  if (threadIdx.x == 0) tot = 0;
  __syncthreads();
  int c = 0;
  for (int j = threadIdx.x; j < nJobs; j += blockDim.x) c += jobs[job0 + j].active ? 1 : 0;
  if (c) atomicAdd(&tot, c);
  __syncthreads();
  if (threadIdx.x == 0) *out = tot;
}

void launch_ext_iteration(JobState * jobs, ViewState * views, int job0, int nJobs, int view0, int nViews, int maxPoints,
                          SolverParams p, int level, bool forceStop, double * trace, size_t traceJobStride, int * activeCount,
                          cudaStream_t s)
{
  (void)job0;
  const dim3 grid((maxPoints + EXT_CHUNK - 1) / EXT_CHUNK, nViews);
  ext_res_kernel<<<grid, EXT_NT, 0, s>>>(jobs, views, view0, p.depthWeight);
  if (p.lossKind != VSLAMB200_LOSS_NONE && p.scalerKind == VSLAMB200_SCALER_TDIST)
    ext_scale_kernel<<<nViews * EXT_SC, EXT_SNT, 0, s>>>(jobs, views, view0, p.scalerV);
  ext_acc_kernel<<<grid, EXT_NT, 0, s>>>(jobs, views, view0, p, level, forceStop ? 1 : 0, trace, traceJobStride);
  if (activeCount) ext_count_active_kernel<<<1, 256, 0, s>>>(jobs, job0, nJobs, activeCount);
}
}  // namespace vb
