// Micro-benchmarks that calibrate the design of the Gauss-Newton level kernel on B200:
//   hist_atoms   : shared-memory atomicAdd histogram of int residuals (clustered around 0)
//   hist_private : thread-private byte sub-histograms (no atomics) + overflow atomics
//   dfma28       : 28 dependent-chain-free DFMA accumulations per point (J^T W J terms)
//   ffma28       : same in fp32
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gpurun_out/microbench tools/microbench.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

constexpr int T = 256;
constexpr int PPT = 128;  // points per thread

__device__ __forceinline__ uint32_t hash32(uint32_t x)
{
  x ^= x >> 16; x *= 0x7feb352dU; x ^= x >> 15; x *= 0x846ca68bU; x ^= x >> 16;
  return x;
}
// pseudo residual: sum of 4 uniform [-s, s] -> roughly normal, sigma ~ s*1.15
__device__ __forceinline__ int fake_r(uint32_t i, int s)
{
  uint32_t h = hash32(i);
  int a = (int)(h & 0xff) % (2 * s + 1) - s, b = (int)((h >> 8) & 0xff) % (2 * s + 1) - s;
  int c = (int)((h >> 16) & 0xff) % (2 * s + 1) - s, d = (int)(h >> 24) % (2 * s + 1) - s;
  int r = a + b + c + d;
  return max(-255, min(255, r));
}

__global__ void __launch_bounds__(T) gen_kernel(int16_t * r, size_t n, int s)
{
  for (size_t i = blockIdx.x * (size_t)T + threadIdx.x; i < n; i += (size_t)gridDim.x * T) r[i] = (int16_t)fake_r((uint32_t)i, s);
}

__global__ void __launch_bounds__(T, 2) hist_atoms_kernel(const int16_t * __restrict__ r, int * __restrict__ out)
{
  __shared__ int hist[512];
  hist[threadIdx.x] = 0; hist[threadIdx.x + 256] = 0;
  __syncthreads();
  const uint2 * p = reinterpret_cast<const uint2 *>(r) + (size_t)blockIdx.x * (T * PPT / 4);
#pragma unroll 4
  for (int k = 0; k < PPT / 4; k++) {
    const uint2 v = __ldg(p + k * T + threadIdx.x);
    atomicAdd(&hist[(int)(int16_t)(v.x & 0xffff) + 255], 1);
    atomicAdd(&hist[(int)(int16_t)(v.x >> 16) + 255], 1);
    atomicAdd(&hist[(int)(int16_t)(v.y & 0xffff) + 255], 1);
    atomicAdd(&hist[(int)(int16_t)(v.y >> 16) + 255], 1);
  }
  __syncthreads();
  if (hist[threadIdx.x]) atomicAdd(out + threadIdx.x, hist[threadIdx.x]);
  if (hist[threadIdx.x + 256]) atomicAdd(out + threadIdx.x + 256, hist[threadIdx.x + 256]);
}

// thread-private byte counters for r in [-NB/2, NB/2): word j of thread t at (j*T + t) -> bank == lane
template <int NB>
__global__ void __launch_bounds__(T, 2) hist_private_kernel(const int16_t * __restrict__ r, int * __restrict__ out)
{
  __shared__ int hist[512];
  __shared__ uint32_t priv[NB / 4 * T];
  hist[threadIdx.x] = 0; hist[threadIdx.x + 256] = 0;
  for (int j = 0; j < NB / 4; j++) priv[j * T + threadIdx.x] = 0;
  __syncthreads();
  const uint2 * p = reinterpret_cast<const uint2 *>(r) + (size_t)blockIdx.x * (T * PPT / 4);
  uint8_t * mine = reinterpret_cast<uint8_t *>(priv) + threadIdx.x * 4;
#pragma unroll 2
  for (int k = 0; k < PPT / 4; k++) {
    const uint2 v = __ldg(p + k * T + threadIdx.x);
    int rr[4] = {(int)(int16_t)(v.x & 0xffff), (int)(int16_t)(v.x >> 16), (int)(int16_t)(v.y & 0xffff), (int)(int16_t)(v.y >> 16)};
#pragma unroll
    for (int q = 0; q < 4; q++) {
      const int b = rr[q] + NB / 2;
      if ((unsigned)b < (unsigned)NB) {
        uint8_t * a = mine + (b >> 2) * (T * 4) + (b & 3);
        *a = *a + 1;
      } else {
        atomicAdd(&hist[rr[q] + 255], 1);
      }
    }
  }
  __syncthreads();
  // merge: thread t sums bin (t % NB) over a slice of threads
  for (int b = threadIdx.x; b < NB; b += T) {
    int s = 0;
    const uint8_t * col = reinterpret_cast<const uint8_t *>(priv) + (b >> 2) * (T * 4) + (b & 3);
    for (int t = 0; t < T; t++) s += col[t * 4];
    hist[b - NB / 2 + 255] += s;
  }
  __syncthreads();
  if (hist[threadIdx.x]) atomicAdd(out + threadIdx.x, hist[threadIdx.x]);
  if (hist[threadIdx.x + 256]) atomicAdd(out + threadIdx.x + 256, hist[threadIdx.x + 256]);
}

template <typename F>
__global__ void __launch_bounds__(T, 2) fma28_kernel(const int16_t * __restrict__ r, F * __restrict__ out)
{
  F acc[28];
#pragma unroll
  for (int s = 0; s < 28; s++) acc[s] = 0;
  const uint2 * p = reinterpret_cast<const uint2 *>(r) + (size_t)blockIdx.x * (T * PPT / 4);
#pragma unroll 1
  for (int k = 0; k < PPT / 4; k++) {
    const uint2 v = __ldg(p + k * T + threadIdx.x);
    int rr[4] = {(int)(int16_t)(v.x & 0xffff), (int)(int16_t)(v.x >> 16), (int)(int16_t)(v.y & 0xffff), (int)(int16_t)(v.y >> 16)};
#pragma unroll
    for (int q = 0; q < 4; q++) {
      const F x = (F)rr[q];
      F J[6] = {x, x + F(1), x * F(0.5), x - F(3), x * x, x + F(7)};
      const F w = F(1) / (F(5) + x * x * F(0.01));
      int s = 0;
#pragma unroll
      for (int i = 0; i < 6; i++) {
        const F wj = w * J[i];
#pragma unroll
        for (int j = i; j < 6; j++) { acc[s] = fma(wj, J[j], acc[s]); s++; }
        acc[21 + i] = fma(wj, x, acc[21 + i]);
      }
      acc[27] = fma(w * x, x, acc[27]);
    }
  }
  F t = 0;
#pragma unroll
  for (int s = 0; s < 28; s++) t += acc[s];
  out[(size_t)blockIdx.x * T + threadIdx.x] = t;
}

template <typename K>
float time_it(K launch)
{
  cudaEvent_t a, b;
  cudaEventCreate(&a); cudaEventCreate(&b);
  launch(); launch();
  cudaEventRecord(a);
  for (int i = 0; i < 5; i++) launch();
  cudaEventRecord(b);
  cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  return ms / 5;
}

int main()
{
  const int blocks = 148 * 2 * 16;
  const size_t n = (size_t)blocks * T * PPT;
  int16_t * r; int * out; double * dout;
  cudaMalloc(&r, n * 2); cudaMalloc(&out, 512 * 4); cudaMalloc(&dout, (size_t)blocks * T * 8);
  cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
  const double clk = 1.965e9;
  for (int s : {2, 6, 20}) {
    gen_kernel<<<2048, T>>>(r, n, s);
    cudaMemset(out, 0, 2048);
    auto rep = [&](const char * name, float ms) {
      printf("spread %2d %-16s %8.3f ms  %7.1f Gpt/s  %6.2f clk/pt/SM\n", s, name, ms, n / ms * 1e-6, ms * 1e-3 * clk * prop.multiProcessorCount / n);
    };
    rep("hist_atoms", time_it([&] { hist_atoms_kernel<<<blocks, T>>>(r, out); }));
    rep("hist_private64", time_it([&] { hist_private_kernel<64><<<blocks, T>>>(r, out); }));
    rep("hist_private128", time_it([&] { hist_private_kernel<128><<<blocks, T>>>(r, out); }));
    rep("dfma28", time_it([&] { fma28_kernel<double><<<blocks, T>>>(r, dout); }));
    rep("ffma28", time_it([&] { fma28_kernel<float><<<blocks, T>>>(r, (float *)dout); }));
  }
  // correctness of the private histogram vs atomics
  int h1[512], h2[512];
  cudaMemset(out, 0, 2048); hist_atoms_kernel<<<blocks, T>>>(r, out); cudaMemcpy(h1, out, 2048, cudaMemcpyDeviceToHost);
  cudaMemset(out, 0, 2048); hist_private_kernel<64><<<blocks, T>>>(r, out); cudaMemcpy(h2, out, 2048, cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int i = 0; i < 512; i++) bad += h1[i] != h2[i];
  printf("private vs atoms mismatching bins: %d, err=%s\n", bad, cudaGetErrorString(cudaGetLastError()));
  return 0;
}
