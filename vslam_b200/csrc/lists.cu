// Interest-point lists: the compacted, row-major-ordered set of template pixels that take part in the
// alignment of one (keyframe, level), built once per level like the reference does
// (SE3Alignment.cpp:83-94 / :123-130 candidate mask; InverseCompositional.cpp:63-93 drops z <= 0).
// Three passes: per-chunk count, per-view exclusive scan, ordered scatter.
#include "common.cuh"

namespace vb
{
constexpr int LTHREADS = 256;
constexpr int LPER = 8;                       // pixels per thread
constexpr int LCHUNK = LTHREADS * LPER;       // pixels per block

__device__ __forceinline__ bool is_interest(
  const FrameDev & f, int level, int multi, size_t i, double minGradient2, float & dOut, uint32_t & gOut)
{
  const int gx = f.gx[level][i], gy = f.gy[level][i];
  const float d = multi ? f.Z[level][i] : f.Zsf[level][i];
  dOut = d;
  gOut = (uint32_t)(uint16_t)gx | ((uint32_t)(uint16_t)gy << 16);
  const int g2 = gx * gx + gy * gy;
  return ((double)g2 >= minGradient2) && (d > 0.f) && isfinite(d);
}

__global__ void __launch_bounds__(LTHREADS) list_count_kernel(
  const FrameDev * __restrict__ frames, const ViewState * __restrict__ views, const JobState * __restrict__ jobs,
  int level, double minGradient2, int * __restrict__ chunkCounts, int maxChunks, int view0)
{
  const int vi = view0 + blockIdx.y;
  const ViewState & V = views[vi];
  const int npx = V.W * V.H;
  const int base = blockIdx.x * LCHUNK;
  if (base >= npx) return;
  const FrameDev & f = frames[V.fromSlot];
  const int multi = jobs[V.job].multi;
  int cnt = 0;
#pragma unroll
  for (int k = 0; k < LPER; k++) {
    const int i = base + k * LTHREADS + threadIdx.x;
    float d;
    uint32_t g;
    if (i < npx && is_interest(f, level, multi, i, minGradient2, d, g)) cnt++;
  }
  cnt = __reduce_add_sync(0xffffffffu, cnt);
  __shared__ int wsum[LTHREADS / 32];
  if ((threadIdx.x & 31) == 0) wsum[threadIdx.x >> 5] = cnt;
  __syncthreads();
  if (threadIdx.x == 0) {
    int s = 0;
    for (int w = 0; w < LTHREADS / 32; w++) s += wsum[w];
    chunkCounts[vi * maxChunks + blockIdx.x] = s;
  }
}

// one block per view: exclusive scan of the chunk counts (in place), nIp and nChunks of the view
__global__ void __launch_bounds__(256) list_scan_kernel(
  ViewState * __restrict__ views, int * __restrict__ chunkCounts, int maxChunks, int view0)
{
  const int vi = view0 + blockIdx.x;
  ViewState & V = views[vi];
  const int npx = V.W * V.H;
  const int n = (npx + LCHUNK - 1) / LCHUNK;
  int * c = chunkCounts + vi * maxChunks;
  __shared__ int carry;
  __shared__ int wsum[8];
  if (threadIdx.x == 0) carry = 0;
  __syncthreads();
  for (int b = 0; b < n; b += 256) {
    const int i = b + threadIdx.x;
    const int v = i < n ? c[i] : 0;
    int incl = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if ((threadIdx.x & 31) >= o) incl += t;
    }
    if ((threadIdx.x & 31) == 31) wsum[threadIdx.x >> 5] = incl;
    __syncthreads();
    int woff = 0;
    for (int w = 0; w < (int)(threadIdx.x >> 5); w++) woff += wsum[w];
    const int excl = carry + woff + incl - v;
    if (i < n) c[i] = excl;
    __syncthreads();
    if (threadIdx.x == 255) carry = excl + v;
    __syncthreads();
  }
  if (threadIdx.x == 0) V.nIp = carry;
  // Padding of the last tile: zero gradient (hence a zero Jacobian row, which adds exact zeros whatever the weight),
  // finite 1/depth, depth 1.  The Jacobian phase of the level kernel then needs no "is this entry real" test.
  const int nIp = carry, padEnd = (nIp + LIST_TILE - 1) / LIST_TILE * LIST_TILE;
  for (int o = nIp + (int)threadIdx.x; o < padEnd; o += 256) {
    unsigned char * tile = V.tiles + (size_t)(o / LIST_TILE) * LIST_TILE_BYTES;
    const int j = o % LIST_TILE;
    reinterpret_cast<float *>(tile + TILE_OFF_D)[j] = 1.f;
    reinterpret_cast<uint32_t *>(tile + TILE_OFF_UVT)[j] = 0u;
    ListRec z;
    z.uvt = 0u;
    z.g = 0u;
    z.zi = 0.0;
    V.rec[o] = z;
  }
}

__global__ void __launch_bounds__(LTHREADS) list_write_kernel(
  const FrameDev * __restrict__ frames, ViewState * __restrict__ views, const JobState * __restrict__ jobs,
  int level, double minGradient2, const int * __restrict__ chunkOffsets, int maxChunks, int view0)
{
  const int vi = view0 + blockIdx.y;
  ViewState & V = views[vi];
  const int npx = V.W * V.H;
  const int base = blockIdx.x * LCHUNK;
  if (base >= npx) return;
  const FrameDev & f = frames[V.fromSlot];
  const int multi = jobs[V.job].multi;
  const int W = V.W;
  __shared__ int wsum[LTHREADS / 32];
  __shared__ unsigned jmax[JSTAT];
  if (threadIdx.x < JSTAT) jmax[threadIdx.x] = 0u;
  int running = chunkOffsets[vi * maxChunks + blockIdx.x];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // bounds of the Jacobian components J_i = fxJ dIx P_i + fyJ dIy Q_i over the list (the moments kernel derives its
  // fixed-point scales from them): max |dIx P_i|, max |dIy Q_i| in fp32
  const float ifx = (float)V.ifx, ify = (float)V.ify, kx = (float)V.kx, ky = (float)V.ky;
  float jm[JSTAT];
#pragma unroll
  for (int i = 0; i < JSTAT; i++) jm[i] = 0.f;
  for (int k = 0; k < LPER; k++) {
    const int i = base + k * LTHREADS + threadIdx.x;
    float d = 0.f;
    uint32_t g = 0;
    const bool keep = i < npx && is_interest(f, level, multi, i, minGradient2, d, g);
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) wsum[warp] = __popc(m);
    __syncthreads();
    int woff = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < LTHREADS / 32; w++) {
      const int s = wsum[w];
      if (w < warp) woff += s;
      tot += s;
    }
    if (keep) {
      const int o = running + woff + __popc(m & ((1u << lane) - 1u));
      const int v = i / W, u = i - v * W;
      unsigned char * tile = V.tiles + (size_t)(o / LIST_TILE) * LIST_TILE_BYTES;
      const int j = o % LIST_TILE;
      const uint32_t uvt = (uint32_t)u | ((uint32_t)v << 12) | ((uint32_t)f.I[level][i] << 24);
      reinterpret_cast<float *>(tile + TILE_OFF_D)[j] = d;
      reinterpret_cast<uint32_t *>(tile + TILE_OFF_UVT)[j] = uvt;
      ListRec rc;
      rc.uvt = uvt;
      rc.g = g;
      rc.zi = 1.0 / (double)d;  // Warp.cpp:170
      V.rec[o] = rc;
      const float ax = fabsf((float)(int16_t)(g & 0xffff)), ay = fabsf((float)(int16_t)(g >> 16));
      const float xn = fabsf(fmaf(ifx, (float)u, kx)), yn = fabsf(fmaf(ify, (float)v, ky)), zi = 1.f / d;
      jm[0] = fmaxf(jm[0], ax * zi);
      jm[2] = fmaxf(jm[2], ax * xn * zi);
      jm[3] = fmaxf(jm[3], ax * xn * yn);
      jm[4] = fmaxf(jm[4], ax * fmaf(xn, xn, 1.f));
      jm[5] = fmaxf(jm[5], ax * yn);
      jm[7] = fmaxf(jm[7], ay * zi);
      jm[8] = fmaxf(jm[8], ay * yn * zi);
      jm[9] = fmaxf(jm[9], ay * fmaf(yn, yn, 1.f));
      jm[10] = fmaxf(jm[10], ay * xn * yn);
      jm[11] = fmaxf(jm[11], ay * xn);
    }
    running += tot;
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < JSTAT; i++) {
    const unsigned m = __reduce_max_sync(0xffffffffu, __float_as_uint(jm[i]));  // non-negative floats order like their bits
    if (lane == 0 && m) atomicMax(&jmax[i], m);
  }
  __syncthreads();
  if (threadIdx.x < JSTAT && jmax[threadIdx.x]) atomicMax(&V.jstat[threadIdx.x], jmax[threadIdx.x]);
}

void launch_build_lists(const FrameDev * frames, ViewState * views, const JobState * jobs, int view0, int nViews, int level,
                        int maxPixels, double minGradient2, int * chunkCounts, int maxChunksPerView, cudaStream_t s)
{
  dim3 grid((maxPixels + LCHUNK - 1) / LCHUNK, nViews);
  list_count_kernel<<<grid, LTHREADS, 0, s>>>(frames, views, jobs, level, minGradient2, chunkCounts, maxChunksPerView, view0);
  list_scan_kernel<<<nViews, 256, 0, s>>>(views, chunkCounts, maxChunksPerView, view0);
  list_write_kernel<<<grid, LTHREADS, 0, s>>>(frames, views, jobs, level, minGradient2, chunkCounts, maxChunksPerView, view0);
}

int list_chunks_for(int pixels) { return (pixels + LCHUNK - 1) / LCHUNK; }
}  // namespace vb
