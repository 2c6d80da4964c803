// Interest-point lists: the compacted, row-major-ordered set of template pixels that take part in the
// alignment of one (keyframe, level), built once per level like the reference does
// (SE3Alignment.cpp:83-94 / :123-130 candidate mask; InverseCompositional.cpp:63-93 drops z <= 0).
// One pass per level: count, look-back over the view's earlier chunks, ordered write-out (list_build_kernel).
#include "common.cuh"

namespace vb
{
constexpr int LTHREADS = 256;
constexpr int LPER = 8;                       // pixels per thread
constexpr int LCHUNK = LTHREADS * LPER;       // pixels per block

constexpr unsigned LREADY = 0x80000000u;

// Single pass: a block takes the next chunk of ITS view from the view's ticket counter (so every chunk in front of it
// belongs to a block that is already running: the look-back below cannot wait for a block that was never scheduled),
// reads its 2048 pixels once with 128-bit loads, publishes its count, sums the counts in front of it, compacts its
// points into shared memory in pixel order and writes tiles and records with consecutive threads on consecutive
// entries.  chunkFlags (count | LREADY per view and chunk) starts at zero (launch_build_lists clears it).
// WANT_JSTAT: also the bounds of the Jacobian components J_i = fxJ dIx P_i + fyJ dIy Q_i over the list (max |dIx P_i|,
// max |dIy Q_i| in fp32), from which the moments kernel derives its fixed-point scales.
template <bool WANT_JSTAT>
__global__ void __launch_bounds__(LTHREADS) list_build_kernel(
  const FrameDev * __restrict__ frames, ViewState * __restrict__ views, const JobState * __restrict__ jobs, int level,
  double minGradient2, unsigned * chunkFlags, int maxChunks, int view0)
{
  __shared__ float sD[LCHUNK];
  __shared__ uint32_t sUvt[LCHUNK];
  __shared__ uint32_t sG[LCHUNK];
  __shared__ int wsum[LTHREADS / 32];
  __shared__ unsigned jmax[JSTAT];
  __shared__ int sChunk, sBase;
  const int vi = view0 + blockIdx.y;
  ViewState & V = views[vi];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int npx = V.W * V.H, nChunks = (npx + LCHUNK - 1) / LCHUNK;
  if (tid == 0) sChunk = atomicAdd(&V.listTicket, 1);
  if (WANT_JSTAT && tid < JSTAT) jmax[tid] = 0u;
  __syncthreads();
  const int chunk = sChunk;
  if (chunk >= nChunks) return;
  const FrameDev & f = frames[V.fromSlot];
  const int16_t * __restrict__ gxp = f.gx[level];
  const int16_t * __restrict__ gyp = f.gy[level];
  const float * __restrict__ dp = jobs[V.job].multi ? f.Z[level] : f.Zsf[level];
  const uint8_t * __restrict__ ip = f.I[level];
  const int first = chunk * LCHUNK + tid * LPER;
  alignas(16) int16_t gx[LPER], gy[LPER];
  alignas(16) float d[LPER];
  alignas(8) uint8_t t[LPER];
  const bool vec = (npx % LPER) == 0 &&
                   (((uintptr_t)gxp | (uintptr_t)gyp | (uintptr_t)dp | (uintptr_t)ip) & 15) == 0;
  if (vec) {
    if (first < npx) {
      *reinterpret_cast<uint4 *>(gx) = __ldg(reinterpret_cast<const uint4 *>(gxp + first));
      *reinterpret_cast<uint4 *>(gy) = __ldg(reinterpret_cast<const uint4 *>(gyp + first));
      *reinterpret_cast<float4 *>(d) = __ldg(reinterpret_cast<const float4 *>(dp + first));
      *reinterpret_cast<float4 *>(d + 4) = __ldg(reinterpret_cast<const float4 *>(dp + first + 4));
      *reinterpret_cast<uint2 *>(t) = __ldg(reinterpret_cast<const uint2 *>(ip + first));
    } else {
#pragma unroll
      for (int k = 0; k < LPER; k++) d[k] = 0.f, gx[k] = gy[k] = 0, t[k] = 0;
    }
  } else {
#pragma unroll
    for (int k = 0; k < LPER; k++) {
      const int i = first + k;
      const bool in = i < npx;
      gx[k] = in ? gxp[i] : (int16_t)0;
      gy[k] = in ? gyp[i] : (int16_t)0;
      d[k] = in ? dp[i] : 0.f;
      t[k] = in ? ip[i] : (uint8_t)0;
    }
  }
  unsigned keep = 0;
#pragma unroll
  for (int k = 0; k < LPER; k++) {
    const int g2 = (int)gx[k] * gx[k] + (int)gy[k] * gy[k];
    if ((double)g2 >= minGradient2 && d[k] > 0.f && isfinite(d[k]) && first + k < npx) keep |= 1u << k;
  }
  const int cnt = __popc(keep);
  int incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int x = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += x;
  }
  if (lane == 31) wsum[warp] = incl;
  __syncthreads();
  int woff = 0, total = 0;
#pragma unroll
  for (int w = 0; w < LTHREADS / 32; w++) {
    if (w < warp) woff += wsum[w];
    total += wsum[w];
  }
  unsigned * flags = chunkFlags + (size_t)vi * maxChunks;
  if (tid == 0) {
    __threadfence();
    atomicExch(&flags[chunk], (unsigned)total | LREADY);
  }
  if (warp == 0) {  // look-back over the view's chunks in front of this one
    unsigned sum = 0;
    for (int j = lane; j < chunk; j += 32) {
      unsigned x;
      do {
        x = *reinterpret_cast<volatile unsigned *>(&flags[j]);
      } while (!(x & LREADY));
      sum += x & ~LREADY;
    }
    sum = __reduce_add_sync(0xffffffffu, sum);
    if (lane == 0) sBase = (int)sum;
  }
  // compaction into shared memory, pixel order
  float jm[JSTAT];
  if (WANT_JSTAT) {
#pragma unroll
    for (int i = 0; i < JSTAT; i++) jm[i] = 0.f;
  }
  const int W = V.W;
  int loc = woff + incl - cnt;
  int v = first / W, u = first - v * W;
  const float ifx = (float)V.ifx, ify = (float)V.ify, kx = (float)V.kx, ky = (float)V.ky;
#pragma unroll
  for (int k = 0; k < LPER; k++) {
    if (keep & (1u << k)) {
      sD[loc] = d[k];
      sUvt[loc] = (uint32_t)u | ((uint32_t)v << 12) | ((uint32_t)t[k] << 24);
      sG[loc] = (uint32_t)(uint16_t)gx[k] | ((uint32_t)(uint16_t)gy[k] << 16);
      loc++;
      if (WANT_JSTAT) {
        const float ax = fabsf((float)gx[k]), ay = fabsf((float)gy[k]);
        const float xn = fabsf(fmaf(ifx, (float)u, kx)), yn = fabsf(fmaf(ify, (float)v, ky)), zi = 1.f / d[k];
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
    }
    if (++u == W) u = 0, v++;
  }
  if (WANT_JSTAT) {
#pragma unroll
    for (int i = 0; i < JSTAT; i++) {
      const unsigned m = __reduce_max_sync(0xffffffffu, __float_as_uint(jm[i]));  // non-negative floats order like their bits
      if (lane == 0 && m) atomicMax(&jmax[i], m);
    }
  }
  __syncthreads();
  if (WANT_JSTAT && tid < JSTAT && jmax[tid]) atomicMax(&V.jstat[tid], jmax[tid]);
  // write-out: consecutive threads, consecutive list entries
  const int base = sBase;
  unsigned char * tiles = V.tiles;
  ListRec * rec = V.rec;
  for (int j = tid; j < total; j += LTHREADS) {
    const int o = base + j;
    unsigned char * tile = tiles + (size_t)(o / LIST_TILE) * LIST_TILE_BYTES;
    const int q = o % LIST_TILE;
    const float dj = sD[j];
    const uint32_t uvt = sUvt[j];
    reinterpret_cast<float *>(tile + TILE_OFF_D)[q] = dj;
    reinterpret_cast<uint32_t *>(tile + TILE_OFF_UVT)[q] = uvt;
    ListRec rc;
    rc.uvt = uvt;
    rc.g = sG[j];
    rc.zi = 1.0 / (double)dj;  // Warp.cpp:170
    rec[o] = rc;
  }
  if (chunk == nChunks - 1) {
    // the view's last chunk: nIp, and the padding of the last tile - zero gradient (hence a zero Jacobian row, which adds
    // exact zeros whatever the weight), finite 1/depth, depth 1.  The Jacobian phase of the level kernel then needs no
    // "is this entry real" test.
    const int nIp = base + total, padEnd = (nIp + LIST_TILE - 1) / LIST_TILE * LIST_TILE;
    if (tid == 0) V.nIp = nIp;
    for (int o = nIp + tid; o < padEnd; o += LTHREADS) {
      unsigned char * tile = tiles + (size_t)(o / LIST_TILE) * LIST_TILE_BYTES;
      const int q = o % LIST_TILE;
      reinterpret_cast<float *>(tile + TILE_OFF_D)[q] = 1.f;
      reinterpret_cast<uint32_t *>(tile + TILE_OFF_UVT)[q] = 0u;
      ListRec z;
      z.uvt = 0u;
      z.g = 0u;
      z.zi = 0.0;
      rec[o] = z;
    }
  }
}

void launch_build_lists(const FrameDev * frames, ViewState * views, const JobState * jobs, int view0, int nViews, int level,
                        int maxPixels, double minGradient2, int * chunkCounts, int maxChunksPerView, int wantJstat,
                        cudaStream_t s)
{
  dim3 grid((maxPixels + LCHUNK - 1) / LCHUNK, nViews);
  unsigned * flags = reinterpret_cast<unsigned *>(chunkCounts);
  cudaMemsetAsync(flags + (size_t)view0 * maxChunksPerView, 0, (size_t)nViews * maxChunksPerView * sizeof(unsigned), s);
  if (wantJstat)
    list_build_kernel<true><<<grid, LTHREADS, 0, s>>>(frames, views, jobs, level, minGradient2, flags, maxChunksPerView, view0);
  else
    list_build_kernel<false><<<grid, LTHREADS, 0, s>>>(frames, views, jobs, level, minGradient2, flags, maxChunksPerView, view0);
}

int list_chunks_for(int pixels) { return (pixels + LCHUNK - 1) / LCHUNK; }
}  // namespace vb
