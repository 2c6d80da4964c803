// Frame preprocessing on the device: ingest, image pyramid, depth pyramid, gradients.
// Replaces Frame::computePyramid / computeDerivatives (core/src/Frame.cpp:215-275) for a batch of frames.
#include <algorithm>

#include "common.cuh"

namespace vb
{
// ------------------------------------------------------------------------------------------------
// ingest: staged host bytes -> level-0 planes (row-major u8 / f32).  Handles Eigen column-major input
// (core/src/types.h:33-35) with a tiled transpose and f64 -> f32 depth narrowing.
// ------------------------------------------------------------------------------------------------
template <typename DepthT>
__global__ void __launch_bounds__(256) ingest_kernel(
  const uint8_t * __restrict__ gray, const DepthT * __restrict__ depth, int layout, int W, int H,
  const FrameDev * __restrict__ frames, const int * __restrict__ slots)
{
  __shared__ uint8_t tg[32][33];
  __shared__ float td[32][33];
  const FrameDev & f = frames[slots[blockIdx.z]];
  const size_t npx = (size_t)W * H;
  const uint8_t * g = gray ? gray + blockIdx.z * npx : nullptr;
  const DepthT * d = depth ? depth + blockIdx.z * npx : nullptr;
  const int x0 = blockIdx.x * 32, y0 = blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 32 x 8
  if (layout == VSLAMB200_ROW_MAJOR) {
    for (int j = ty; j < 32; j += 8) {
      const int x = x0 + tx, y = y0 + j;
      if (x < W && y < H) {
        if (g) f.I[0][(size_t)y * W + x] = g[(size_t)y * W + x];
        if (d) f.Z[0][(size_t)y * W + x] = (float)d[(size_t)y * W + x];
      }
    }
  } else {
    // column-major source: element (y, x) at x*H + y; read coalesced along y, write coalesced along x
    for (int j = ty; j < 32; j += 8) {
      const int x = x0 + j, y = y0 + tx;
      if (x < W && y < H) {
        if (g) tg[j][tx] = g[(size_t)x * H + y];
        if (d) td[j][tx] = (float)d[(size_t)x * H + y];
      }
    }
    __syncthreads();
    for (int j = ty; j < 32; j += 8) {
      const int x = x0 + tx, y = y0 + j;
      if (x < W && y < H) {
        if (g) f.I[0][(size_t)y * W + x] = tg[tx][j];
        if (d) f.Z[0][(size_t)y * W + x] = td[tx][j];
      }
    }
  }
}

void launch_ingest(const uint8_t * stagedGray, const void * stagedDepth, int depthType, int layout, int W, int H,
                   const FrameDev * frames, const int * slots, int n, cudaStream_t s)
{
  dim3 grid((W + 31) / 32, (H + 31) / 32, n);
  if (depthType == VSLAMB200_DEPTH_F64)
    ingest_kernel<double><<<grid, 256, 0, s>>>(stagedGray, (const double *)stagedDepth, layout, W, H, frames, slots);
  else
    ingest_kernel<float><<<grid, 256, 0, s>>>(stagedGray, (const float *)stagedDepth, layout, W, H, frames, slots);
}

// Node-side ingestion fused into the upload (src/ros/nodes/NodeRgbdAlignment.cpp:177-199): cv::cvtColor(RGB2GRAY) of
// the interleaved 8UC3 image (OpenCV's 8-bit integer formula, 15-bit coefficients), depth / divisor in double
// (loadDepth(...) / 5000.0, odometry/test/test_alignment_se3.cpp:99), NaN -> 0 (:190), narrowed to the f32 plane.
// Four pixels per thread: 12 colour bytes in, one 32-bit store of grey values and one 128-bit store of depths out.
template <typename DepthT>
__global__ void __launch_bounds__(256) ingest_rgbd_kernel(
  const uint8_t * __restrict__ rgb, const DepthT * __restrict__ depth, int bgr, double divisor, size_t npx,
  const FrameDev * __restrict__ frames, const int * __restrict__ slots)
{
  const FrameDev & f = frames[slots[blockIdx.y]];
  const uint8_t * c = rgb + (size_t)blockIdx.y * npx * 3;
  const DepthT * d = depth + (size_t)blockIdx.y * npx;
  for (size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 4; i < npx; i += (size_t)gridDim.x * blockDim.x * 4) {
    uint8_t g[4];
    float z[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      const size_t j = i + k < npx ? i + k : npx - 1;
      const int c0 = c[3 * j], c1 = c[3 * j + 1], c2 = c[3 * j + 2];
      const int r = bgr ? c2 : c0, b = bgr ? c0 : c2;
      g[k] = (uint8_t)((r * 9798 + c1 * 19235 + b * 3735 + (1 << 14)) >> 15);
      const double v = (double)d[j] / divisor;
      z[k] = isnan(v) ? 0.f : (float)v;
    }
    if (i + 3 < npx && (npx & 3) == 0) {
      *reinterpret_cast<uchar4 *>(f.I[0] + i) = make_uchar4(g[0], g[1], g[2], g[3]);
      *reinterpret_cast<float4 *>(f.Z[0] + i) = make_float4(z[0], z[1], z[2], z[3]);
    } else {
      for (int k = 0; k < 4 && i + k < npx; k++) {
        f.I[0][i + k] = g[k];
        f.Z[0][i + k] = z[k];
      }
    }
  }
}

void launch_ingest_rgbd(const uint8_t * stagedRgb, const void * stagedDepth, int depthType, int bgr, double divisor, int W, int H,
                        const FrameDev * frames, const int * slots, int n, cudaStream_t s)
{
  const size_t npx = (size_t)W * H;
  dim3 grid((unsigned)std::min<size_t>(148 * 4, (npx / 4 + 255) / 256), n);
  if (depthType == VSLAMB200_DEPTH_F64)
    ingest_rgbd_kernel<double><<<grid, 256, 0, s>>>(stagedRgb, (const double *)stagedDepth, bgr, divisor, npx, frames, slots);
  else if (depthType == VSLAMB200_DEPTH_U16)
    ingest_rgbd_kernel<uint16_t><<<grid, 256, 0, s>>>(stagedRgb, (const uint16_t *)stagedDepth, bgr, divisor, npx, frames, slots);
  else
    ingest_rgbd_kernel<float><<<grid, 256, 0, s>>>(stagedRgb, (const float *)stagedDepth, bgr, divisor, npx, frames, slots);
}

// 16-bit sensor depth -> level-0 depth plane: k / divisor in double, narrowed to the f32 plane (the depth half of the
// ingestion above; loadDepth(...) / 5000.0, odometry/test/test_alignment_se3.cpp:99).  Eight pixels per thread.
__global__ void __launch_bounds__(256) depth_u16_kernel(
  const uint16_t * __restrict__ staged, double divisor, size_t npx, const FrameDev * __restrict__ frames,
  const int * __restrict__ slots)
{
  const FrameDev & f = frames[slots[blockIdx.y]];
  const uint16_t * d = staged + (size_t)blockIdx.y * npx;
  float * __restrict__ z = f.Z[0];
  const bool vec = (npx & 7) == 0;
  for (size_t i = ((size_t)blockIdx.x * blockDim.x + threadIdx.x) * 8; i < npx; i += (size_t)gridDim.x * blockDim.x * 8) {
    if (vec) {
      const uint4 w = *reinterpret_cast<const uint4 *>(d + i);
      const uint32_t ww[4] = {w.x, w.y, w.z, w.w};
      float o[8];
#pragma unroll
      for (int k = 0; k < 4; k++) {
        o[2 * k] = (float)((double)(ww[k] & 0xffff) / divisor);
        o[2 * k + 1] = (float)((double)(ww[k] >> 16) / divisor);
      }
      *reinterpret_cast<float4 *>(z + i) = make_float4(o[0], o[1], o[2], o[3]);
      *reinterpret_cast<float4 *>(z + i + 4) = make_float4(o[4], o[5], o[6], o[7]);
    } else {
      for (int k = 0; k < 8 && i + k < npx; k++) z[i + k] = (float)((double)d[i + k] / divisor);
    }
  }
}

void launch_depth_u16(const uint16_t * staged, double divisor, int W, int H, const FrameDev * frames, const int * slots, int n,
                      cudaStream_t s)
{
  const size_t npx = (size_t)W * H;
  dim3 grid((unsigned)std::min<size_t>(148 * 2, (npx / 8 + 255) / 256), n);
  depth_u16_kernel<<<grid, 256, 0, s>>>(staged, divisor, npx, frames, slots);
}

// ------------------------------------------------------------------------------------------------
// K1a + K1b: one pyramid step for intensity (cv::pyrDown: 5x5 binomial, reflect-101, (s+128)>>8, even
// rows/cols; Frame.cpp:264) and depth (3x3 masked median at (2i,2j) with zero border, even-count rule
// (v[n/2-1]+v[n/2+1])/2, n==2 -> v[0]; Frame.cpp:269-274, algorithm.h:157-184, algorithm.cpp:109-120).
// The intensity tile is staged through shared memory (each source pixel is read once per block).
// ------------------------------------------------------------------------------------------------
constexpr int PT = 32;  // output tile edge of the pyrDown kernel; source tile (2*PT+3)^2

__global__ void __launch_bounds__(256) pyr_down_u8_kernel(
  const FrameDev * __restrict__ frames, const int * __restrict__ slots, int level, int Wp, int Hp, int W, int H)
{
  __shared__ uint8_t src[2 * PT + 3][2 * PT + 4];
  __shared__ uint16_t hrow[2 * PT + 3][PT];
  const FrameDev & f = frames[slots[blockIdx.z]];
  const uint8_t * __restrict__ in = f.I[level - 1];
  uint8_t * __restrict__ out = f.I[level];
  const int ox0 = blockIdx.x * PT, oy0 = blockIdx.y * PT;
  const int sx0 = 2 * ox0 - 2, sy0 = 2 * oy0 - 2;
  for (int i = threadIdx.x; i < (2 * PT + 3) * (2 * PT + 3); i += 256) {
    const int r = i / (2 * PT + 3), c = i - r * (2 * PT + 3);
    const int y = reflect101(sy0 + r, Hp), x = reflect101(sx0 + c, Wp);
    src[r][c] = in[(size_t)y * Wp + x];
  }
  __syncthreads();
  // horizontal pass at even columns
  for (int i = threadIdx.x; i < (2 * PT + 3) * PT; i += 256) {
    const int r = i / PT, c = i - r * PT;
    const uint8_t * p = &src[r][2 * c];
    hrow[r][c] = (uint16_t)(p[0] + 4 * p[1] + 6 * p[2] + 4 * p[3] + p[4]);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < PT * PT; i += 256) {
    const int r = i / PT, c = i - r * PT;
    const int ox = ox0 + c, oy = oy0 + r;
    if (ox < W && oy < H) {
      const int sacc = hrow[2 * r][c] + 4 * hrow[2 * r + 1][c] + 6 * hrow[2 * r + 2][c] + 4 * hrow[2 * r + 3][c] +
                       hrow[2 * r + 4][c];
      out[(size_t)oy * W + ox] = (uint8_t)((sacc + 128) >> 8);
    }
  }
}

__device__ __forceinline__ void cswap(float & a, float & b)
{
  const float lo = fminf(a, b), hi = fmaxf(a, b);
  a = lo;
  b = hi;
}

__global__ void __launch_bounds__(256) depth_down_kernel(
  const FrameDev * __restrict__ frames, const int * __restrict__ slots, int level, int Wdp, int Hdp, int Wd, int Hd)
{
  const FrameDev & f = frames[slots[blockIdx.z]];
  const float * __restrict__ in = f.Z[level - 1];
  const int j = blockIdx.x * 32 + (threadIdx.x & 31);
  const int i = blockIdx.y * 8 + (threadIdx.x >> 5);
  if (i >= Hd || j >= Wd) return;
  const int y = 2 * i, x = 2 * j;
  float out = 0.f;
  if (y >= 1 && y < Hdp - 1 && x >= 1 && x < Wdp - 1) {
    const float INF = __int_as_float(0x7f800000);
    float v[9];
    int n = 0;
#pragma unroll
    for (int dy = -1; dy <= 1; dy++)
#pragma unroll
      for (int dx = -1; dx <= 1; dx++) {
        const float z = in[(size_t)(y + dy) * Wdp + (x + dx)];
        const bool keep = !(z <= 0.f);  // mask v <= 0 (NaN passes, as in the reference)
        v[(dy + 1) * 3 + dx + 1] = keep ? z : INF;
        n += keep;
      }
    // 9-input sorting network (25 compare-exchanges); masked samples sort to the end as +inf.
    // NaN inputs are undefined here exactly as std::sort with NaN is in the reference.
    cswap(v[0], v[1]); cswap(v[3], v[4]); cswap(v[6], v[7]);
    cswap(v[1], v[2]); cswap(v[4], v[5]); cswap(v[7], v[8]);
    cswap(v[0], v[1]); cswap(v[3], v[4]); cswap(v[6], v[7]);
    cswap(v[0], v[3]); cswap(v[3], v[6]); cswap(v[0], v[3]);
    cswap(v[1], v[4]); cswap(v[4], v[7]); cswap(v[1], v[4]);
    cswap(v[2], v[5]); cswap(v[5], v[8]); cswap(v[2], v[5]);
    cswap(v[1], v[3]); cswap(v[5], v[7]);
    cswap(v[2], v[6]); cswap(v[4], v[6]); cswap(v[2], v[4]);
    cswap(v[2], v[3]); cswap(v[5], v[6]);
    float a = 0.f, b = 0.f;
    bool even = false;
    switch (n) {
      case 0: a = 0.f; break;
      case 1: a = v[0]; break;
      case 2: a = v[0]; break;                       // reference reads out of bounds here; documented rule
      case 3: a = v[1]; break;
      case 4: a = v[1]; b = v[3]; even = true; break;
      case 5: a = v[2]; break;
      case 6: a = v[2]; b = v[4]; even = true; break;
      case 7: a = v[3]; break;
      case 8: a = v[3]; b = v[5]; even = true; break;
      default: a = v[4]; break;
    }
    out = even ? (float)(((double)a + (double)b) / 2.0) : a;
  }
  f.Z[level][(size_t)i * Wd + j] = out;
  // single-frame candidate mask folded into a second depth plane: depth(0)(int(v*2^l), int(u*2^l)) > 0
  const int W0 = f.Wd[0];
  const float z0 = f.Z[0][(size_t)(i << level) * W0 + (j << level)];
  f.Zsf[level][(size_t)i * Wd + j] = (z0 > 0.f) ? out : 0.f;
}

// cv::pyrDown for source widths that are a multiple of 8, register-resident: a lane owns four consecutive output
// columns and marches down a band of output rows.  Per source row it loads the 11 bytes 2x-2 .. 2x+8 (four aligned
// words), forms the horizontal [1 4 6 4 1] sums at the four even centres, and keeps the sums of five source rows; every
// output row consumes two new source rows.  Same integers as the tiled kernel: (sum + 128) >> 8, reflect-101.
constexpr int PB_ROWS = 16;  // output rows per warp

__device__ __forceinline__ void pyr_hrow(const uint8_t * __restrict__ in, int Wp, int Hp, int sy, int sx, int (&h)[4])
{
  const uint8_t * row = in + (size_t)reflect101(sy, Hp) * Wp;
  const uint32_t w1 = *reinterpret_cast<const uint32_t *>(row + sx), w2 = *reinterpret_cast<const uint32_t *>(row + sx + 4);
  int p[11];
  p[2] = w1 & 0xff; p[3] = (w1 >> 8) & 0xff; p[4] = (w1 >> 16) & 0xff; p[5] = w1 >> 24;
  p[6] = w2 & 0xff; p[7] = (w2 >> 8) & 0xff; p[8] = (w2 >> 16) & 0xff; p[9] = w2 >> 24;
  if (sx >= 4) {
    const uint32_t w0 = *reinterpret_cast<const uint32_t *>(row + sx - 4);
    p[0] = (w0 >> 16) & 0xff; p[1] = w0 >> 24;
  } else {  // left border (sx == 0): columns -2, -1 reflect to 2, 1
    p[0] = p[4]; p[1] = p[3];
  }
  if (sx + 8 < Wp) p[10] = row[sx + 8];
  else p[10] = p[8];  // right border (sx + 8 == Wp): column Wp reflects to Wp - 2
#pragma unroll
  for (int j = 0; j < 4; j++) h[j] = p[2 * j] + 4 * p[2 * j + 1] + 6 * p[2 * j + 2] + 4 * p[2 * j + 3] + p[2 * j + 4];
}

__global__ void __launch_bounds__(256) pyr_down_v4_kernel(
  const FrameDev * __restrict__ frames, const int * __restrict__ slots, int level, int Wp, int Hp, int W, int H)
{
  const FrameDev & f = frames[slots[blockIdx.z]];
  const uint8_t * __restrict__ in = f.I[level - 1];
  uint8_t * __restrict__ out = f.I[level];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int x = (blockIdx.x * 32 + lane) * 4;
  const int y0 = (blockIdx.y * 8 + warp) * PB_ROWS;
  if (x >= W || y0 >= H) return;
  const int y1 = min(y0 + PB_ROWS, H);
  int h0[4], h1[4], h2[4], h3[4], h4[4];  // source rows 2y-2 .. 2y+2
  pyr_hrow(in, Wp, Hp, 2 * y0 - 2, 2 * x, h0);
  pyr_hrow(in, Wp, Hp, 2 * y0 - 1, 2 * x, h1);
  pyr_hrow(in, Wp, Hp, 2 * y0, 2 * x, h2);
  for (int y = y0; y < y1; y++) {
    pyr_hrow(in, Wp, Hp, 2 * y + 1, 2 * x, h3);
    pyr_hrow(in, Wp, Hp, 2 * y + 2, 2 * x, h4);
    uint32_t o = 0;
#pragma unroll
    for (int j = 0; j < 4; j++) {
      const int sacc = h0[j] + 4 * h1[j] + 6 * h2[j] + 4 * h3[j] + h4[j];
      o |= (uint32_t)((sacc + 128) >> 8) << (8 * j);
      h0[j] = h2[j];
      h1[j] = h3[j];
      h2[j] = h4[j];
    }
    *reinterpret_cast<uint32_t *>(out + (size_t)y * W + x) = o;
  }
}

void launch_pyramid_level(const FrameDev * frames, const int * slots, int n, int level, int W, int H, int Wd, int Hd,
                          int Wp, int Hp, int Wdp, int Hdp, cudaStream_t s)
{
  if (Wp % 8 == 0 && Wp >= 16 && Hp >= 3) {
    dim3 g1((W / 4 + 31) / 32, (H + 8 * PB_ROWS - 1) / (8 * PB_ROWS), n);
    pyr_down_v4_kernel<<<g1, 256, 0, s>>>(frames, slots, level, Wp, Hp, W, H);
  } else {
    dim3 g1((W + PT - 1) / PT, (H + PT - 1) / PT, n);
    pyr_down_u8_kernel<<<g1, 256, 0, s>>>(frames, slots, level, Wp, Hp, W, H);
  }
  if (Wd > 0 && Hd > 0) {
    dim3 g2((Wd + 31) / 32, (Hd + 7) / 8, n);
    depth_down_kernel<<<g2, 256, 0, s>>>(frames, slots, level, Wdp, Hdp, Wd, Hd);
  }
}

// ------------------------------------------------------------------------------------------------
// K2: cv::GaussianBlur 3x3 ((s+8)>>4, reflect-101) fused with cv::Sobel CV_16S dx, dy (reflect-101 on the
// blurred image) — Frame.cpp:225-228.  Blurred halo tile lives in shared memory only.
// ------------------------------------------------------------------------------------------------
constexpr int GT_X = 64, GT_Y = 16;

__global__ void __launch_bounds__(256) blur_sobel_kernel(
  const FrameDev * __restrict__ frames, const int * __restrict__ slots, int level, int W, int H)
{
  __shared__ uint8_t blur[GT_Y + 2][GT_X + 2];
  const FrameDev & f = frames[slots[blockIdx.z]];
  const uint8_t * __restrict__ in = f.I[level];
  const int x0 = blockIdx.x * GT_X, y0 = blockIdx.y * GT_Y;
  for (int i = threadIdx.x; i < (GT_Y + 2) * (GT_X + 2); i += 256) {
    const int r = i / (GT_X + 2), c = i - r * (GT_X + 2);
    // position in the blurred image, reflected into range (Sobel's border), then blur with its own reflect
    const int by = reflect101(y0 + r - 1, H), bx = reflect101(x0 + c - 1, W);
    const int ym = reflect101(by - 1, H), yp = reflect101(by + 1, H);
    const int xm = reflect101(bx - 1, W), xp = reflect101(bx + 1, W);
    const uint8_t * r0 = in + (size_t)ym * W;
    const uint8_t * r1 = in + (size_t)by * W;
    const uint8_t * r2 = in + (size_t)yp * W;
    const int sacc = (r0[xm] + 2 * r0[bx] + r0[xp]) + 2 * (r1[xm] + 2 * r1[bx] + r1[xp]) + (r2[xm] + 2 * r2[bx] + r2[xp]);
    blur[r][c] = (uint8_t)((sacc + 8) >> 4);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < GT_X * GT_Y; i += 256) {
    const int r = i / GT_X, c = i - r * GT_X;
    const int x = x0 + c, y = y0 + r;
    if (x < W && y < H) {
      const int a00 = blur[r][c], a01 = blur[r][c + 1], a02 = blur[r][c + 2];
      const int a10 = blur[r + 1][c], a12 = blur[r + 1][c + 2];
      const int a20 = blur[r + 2][c], a21 = blur[r + 2][c + 1], a22 = blur[r + 2][c + 2];
      f.gx[level][(size_t)y * W + x] = (int16_t)((a02 - a00) + 2 * (a12 - a10) + (a22 - a20));
      f.gy[level][(size_t)y * W + x] = (int16_t)((a20 + 2 * a21 + a22) - (a00 + 2 * a01 + a02));
    }
  }
}

// The same operation for image widths that are a multiple of 4 (every level of the usual sizes), register-resident:
// a lane owns four consecutive columns and marches down a band of rows.  Per source row it loads the 8 bytes
// x-2 .. x+5 (three aligned words), forms the horizontal [1 2 1] sums of 6 columns, completes the blurred row of the
// row above from three such rows (the 3x3 binomial is separable and the rounding (s+8)>>4 is applied once), and from
// three blurred rows the Sobel pair of the row above that.  Reflect-101 of the raw image with a halo of 2 equals
// reflect-101 of the blurred image with a halo of 1 (the kernel is symmetric), so one border rule serves both steps.
// One 32-bit load per source byte quadruple instead of nine byte loads per pixel; 8-byte stores of gx and gy.
constexpr int GB_ROWS = 32;  // output rows per warp

__device__ __forceinline__ void load_row8(const uint8_t * __restrict__ in, int W, int H, int y, int x, int (&p)[8])
{
  const uint8_t * row = in + (size_t)reflect101(y, H) * W;
  const uint32_t w1 = *reinterpret_cast<const uint32_t *>(row + x);
  p[2] = w1 & 0xff; p[3] = (w1 >> 8) & 0xff; p[4] = (w1 >> 16) & 0xff; p[5] = w1 >> 24;
  if (x >= 4) {
    const uint32_t w0 = *reinterpret_cast<const uint32_t *>(row + x - 4);
    p[0] = (w0 >> 16) & 0xff; p[1] = w0 >> 24;
  } else {  // left border (x == 0): columns -2, -1 reflect to 2, 1
    p[0] = p[4]; p[1] = p[3];
  }
  if (x + 8 <= W) {
    const uint32_t w2 = *reinterpret_cast<const uint32_t *>(row + x + 4);
    p[6] = w2 & 0xff; p[7] = (w2 >> 8) & 0xff;
  } else {  // right border (x == W-4): columns W, W+1 reflect to W-2, W-3
    p[6] = p[4]; p[7] = p[3];
  }
}

__global__ void __launch_bounds__(256) blur_sobel_v4_kernel(
  const FrameDev * __restrict__ frames, const int * __restrict__ slots, int level, int W, int H)
{
  const FrameDev & f = frames[slots[blockIdx.z]];
  const uint8_t * __restrict__ in = f.I[level];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int x = (blockIdx.x * 32 + lane) * 4;
  const int y0 = (blockIdx.y * 8 + warp) * GB_ROWS;
  if (x >= W || y0 >= H) return;
  const int y1 = min(y0 + GB_ROWS, H);
  int hA[6], hB[6], hC[6];  // horizontal sums of source rows s-2, s-1, s
  int bA[6], bB[6];         // blurred rows s-3, s-2 (bC, row s-1, is formed from hA, hB, hC)
#pragma unroll
  for (int k = 0; k < 6; k++) hA[k] = hB[k] = hC[k] = bA[k] = bB[k] = 0;
  // source rows y0-2 .. y1+1; after source row s: blurred row s-1, output row s-2
  for (int sr = y0 - 2; sr <= y1 + 1; sr++) {
    int p[8];
    load_row8(in, W, H, sr, x, p);
#pragma unroll
    for (int k = 0; k < 6; k++) {
      hA[k] = hB[k];
      hB[k] = hC[k];
      hC[k] = p[k] + 2 * p[k + 1] + p[k + 2];
    }
    if (sr < y0) continue;  // fewer than three source rows yet
    int bC[6];
#pragma unroll
    for (int k = 0; k < 6; k++) bC[k] = (hA[k] + 2 * hB[k] + hC[k] + 8) >> 4;
    const int oy = sr - 2;
    if (oy >= y0) {  // blurred rows oy-1, oy, oy+1 = bA, bB, bC
      int gx[4], gy[4];
#pragma unroll
      for (int k = 0; k < 4; k++) {
        gx[k] = (bA[k + 2] - bA[k]) + 2 * (bB[k + 2] - bB[k]) + (bC[k + 2] - bC[k]);
        gy[k] = (bC[k] + 2 * bC[k + 1] + bC[k + 2]) - (bA[k] + 2 * bA[k + 1] + bA[k + 2]);
      }
      const size_t o = (size_t)oy * W + x;
      *reinterpret_cast<uint2 *>(f.gx[level] + o) =
        make_uint2(((uint32_t)gx[0] & 0xffffu) | ((uint32_t)gx[1] << 16), ((uint32_t)gx[2] & 0xffffu) | ((uint32_t)gx[3] << 16));
      *reinterpret_cast<uint2 *>(f.gy[level] + o) =
        make_uint2(((uint32_t)gy[0] & 0xffffu) | ((uint32_t)gy[1] << 16), ((uint32_t)gy[2] & 0xffffu) | ((uint32_t)gy[3] << 16));
    }
#pragma unroll
    for (int k = 0; k < 6; k++) {
      bA[k] = bB[k];
      bB[k] = bC[k];
    }
  }
}

void launch_gradients(const FrameDev * frames, const int * slots, int n, int level, int W, int H, cudaStream_t s)
{
  if (W % 4 == 0 && W >= 8 && H >= 4) {
    dim3 g((W / 4 + 31) / 32, (H + 8 * GB_ROWS - 1) / (8 * GB_ROWS), n);
    blur_sobel_v4_kernel<<<g, 256, 0, s>>>(frames, slots, level, W, H);
    return;
  }
  dim3 g((W + GT_X - 1) / GT_X, (H + GT_Y - 1) / GT_Y, n);
  blur_sobel_kernel<<<g, 256, 0, s>>>(frames, slots, level, W, H);
}
}  // namespace vb
