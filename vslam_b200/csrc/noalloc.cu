// vslamb200_interest_points_noalloc: the interest-point list of one pyramid level (the candidate mask of
// SE3Alignment.cpp:83-94 / :123-130, compacted in row-major order like lists.cu does) for callers that cannot allocate:
// code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc.  No counterpart in the
// reference.
//
// The three list kernels of lists.cu keep their per-chunk counts in a workspace that vslamb200_align allocates.  This
// variant is ONE kernel whose scratch - a ticket counter and one record per 2048-pixel chunk - lives in a file-scope
// static device array, g_noalloc_scratch, sized for VSLAMB200_NOALLOC_MAX_PIXELS (4096 x 4096): about 33 KB that every
// GPU which loads the library holds until exit.  The entry point calls no cudaMalloc / cudaFree and no allocating
// library routine.  Every call shares the array and clears the part it uses first, so calls that overlap on the GPU
// (two streams, two threads) race: serialise them.
//
// In host code the array's name is a host-side stand-in, not a device address.  The address the kernel gets comes from
// cudaGetSymbolAddress, looked up once per device with that device current (scratch_for_current_device).
#include "common.cuh"

namespace vb
{
namespace na
{
constexpr int THREADS = 256;
constexpr int PER = 8;                                  // consecutive pixels per thread
constexpr int CHUNK = THREADS * PER;                    // pixels per block
constexpr int MAX_CHUNKS = VSLAMB200_NOALLOC_MAX_PIXELS / CHUNK;
constexpr int HEAD = 64;                                // words in front of the chunk records: [0] is the ticket counter
constexpr unsigned READY = 0x80000000u;
static_assert(VSLAMB200_NOALLOC_MAX_PIXELS % CHUNK == 0, "whole chunks");

static __device__ __align__(256) unsigned g_noalloc_scratch[HEAD + MAX_CHUNKS];

// Blocks take their chunk from a ticket counter, so a block's predecessors are always already running: the look-back
// below cannot wait for a block that has not been scheduled.
__global__ void __launch_bounds__(THREADS) interest_points_noalloc_kernel(
  const int16_t * __restrict__ gx, const int16_t * __restrict__ gy, const float * __restrict__ depth, int npx, int nChunks,
  double minGradient2, uint32_t * __restrict__ points, int * __restrict__ count, unsigned * scratch)
{
  __shared__ unsigned sChunk, sBase;
  __shared__ int wsum[THREADS / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) sChunk = atomicAdd(&scratch[0], 1u);
  __syncthreads();
  const int chunk = (int)sChunk;
  if (chunk >= nChunks) return;
  const int first = chunk * CHUNK + tid * PER;
  unsigned keep = 0;
#pragma unroll
  for (int k = 0; k < PER; k++) {
    const int i = first + k;
    if (i < npx) {
      const int x = gx[i], y = gy[i];
      const float d = depth[i];
      if ((double)(x * x + y * y) >= minGradient2 && d > 0.f && isfinite(d)) keep |= 1u << k;
    }
  }
  const int cnt = __popc(keep);
  int incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) wsum[warp] = incl;
  __syncthreads();
  int woff = 0, total = 0;
#pragma unroll
  for (int w = 0; w < THREADS / 32; w++) {
    if (w < warp) woff += wsum[w];
    total += wsum[w];
  }
  if (tid == 0) {
    __threadfence();
    atomicExch(&scratch[HEAD + chunk], (unsigned)total | READY);
  }
  if (warp == 0) {  // look-back over the chunks in front of this one
    unsigned sum = 0;
    for (int j = lane; j < chunk; j += 32) {
      unsigned v;
      do {
        v = *reinterpret_cast<volatile unsigned *>(&scratch[HEAD + j]);
      } while (!(v & READY));
      sum += v & ~READY;
    }
    sum = __reduce_add_sync(0xffffffffu, sum);
    if (lane == 0) sBase = sum;
  }
  __syncthreads();
  uint32_t o = sBase + (unsigned)(woff + incl - cnt);
#pragma unroll
  for (int k = 0; k < PER; k++)
    if (keep & (1u << k)) points[o++] = (uint32_t)(first + k);
  if (chunk == nChunks - 1 && tid == 0) *count = (int)(sBase + (unsigned)total);
}

// Device address of the scratch array on the current device, resolved once per device.
cudaError_t scratch_for_current_device(unsigned ** out)
{
  constexpr int kMaxDevices = 64;
  static unsigned * addr[kMaxDevices] = {};
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
  if (!addr[dev]) {
    unsigned * p = nullptr;
    e = cudaGetSymbolAddress((void **)&p, g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    addr[dev] = p;
  }
  *out = addr[dev];
  return cudaSuccess;
}
}  // namespace na
}  // namespace vb

extern "C" {

int vslamb200_interest_points_noalloc(
  const int16_t * dGx, const int16_t * dGy, const float * dDepth, int W, int H, double minGradient, uint32_t * dPoints,
  int * dCount, void * cudaStream)
{
  auto fail = [](int code, const char * msg) { return vb::set_last_error(code, msg); };  // read with vslamb200_last_error()
  if (!dGx || !dGy || !dDepth || !dPoints || !dCount) return fail(VSLAMB200_EINVAL, "interest_points_noalloc: null argument");
  if (W < 1 || H < 1) return fail(VSLAMB200_EINVAL, "interest_points_noalloc: bad image size");
  if ((long long)W * H > (long long)VSLAMB200_NOALLOC_MAX_PIXELS)
    return fail(VSLAMB200_EINVAL,
                "interest_points_noalloc: W * H exceeds VSLAMB200_NOALLOC_MAX_PIXELS (the static scratch array is sized for "
                "4096 x 4096 pixels); use vslamb200_align, which allocates its workspace");
  unsigned * scratch = nullptr;
  cudaError_t e = vb::na::scratch_for_current_device(&scratch);
  if (e != cudaSuccess) return fail(VSLAMB200_ECUDA, cudaGetErrorString(e));
  const int npx = W * H, nChunks = (npx + vb::na::CHUNK - 1) / vb::na::CHUNK;
  cudaStream_t s = (cudaStream_t)cudaStream;
  // the array keeps the last call's contents: the ticket and this call's chunk records start at zero
  e = cudaMemsetAsync(scratch, 0, (size_t)(vb::na::HEAD + nChunks) * sizeof(unsigned), s);
  if (e == cudaSuccess) {
    vb::na::interest_points_noalloc_kernel<<<nChunks, vb::na::THREADS, 0, s>>>(dGx, dGy, dDepth, npx, nChunks,
                                                                               minGradient * minGradient, dPoints, dCount, scratch);
    e = cudaGetLastError();
  }
  if (e != cudaSuccess) return fail(VSLAMB200_ECUDA, cudaGetErrorString(e));
  return VSLAMB200_OK;
}

}  // extern "C"
