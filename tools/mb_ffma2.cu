// Micro-benchmark: issue rate of FFMA against FFMA2 (packed fp32 pair, sm_100) - is the pair one issue slot?
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/bin/mb_ffma2 tools/mb_ffma2.cu
#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned long long pk(float a, float b){ unsigned long long r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r;}
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c){ unsigned long long r; asm volatile("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r;}
template <int MODE>
__global__ void k(float * out, float c0, float c1, int iters)
{
  float x[8];
  for (int i = 0; i < 8; i++) x[i] = threadIdx.x * 0.001f + i;
  unsigned long long y[4];
  for (int i = 0; i < 4; i++) y[i] = pk(x[2 * i], x[2 * i + 1]);
  const unsigned long long C0 = pk(c0, c0), C1 = pk(c1, c1);
  for (int it = 0; it < iters; it++) {
    if (MODE == 0) {
#pragma unroll
      for (int r = 0; r < 8; r++)
#pragma unroll
        for (int i = 0; i < 8; i++) asm volatile("fma.rn.f32 %0, %0, %1, %2;" : "+f"(x[i]) : "f"(c0), "f"(c1));
    } else {
#pragma unroll
      for (int r = 0; r < 8; r++)
#pragma unroll
        for (int i = 0; i < 4; i++) y[i] = fma2(y[i], C0, C1);
    }
  }
  float s = 0;
  for (int i = 0; i < 8; i++) s += x[i];
  for (int i = 0; i < 4; i++) { float lo, hi; asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(y[i])); s += lo + hi; }
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
int main()
{
  float * out; cudaMalloc(&out, 148 * 8 * 256 * 4);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  const int iters = 20000;
  for (int mode = 0; mode < 2; mode++) {
    for (int rep = 0; rep < 2; rep++) {
      cudaEventRecord(a);
      if (mode == 0) k<0><<<148 * 8, 256>>>(out, 0.999f, 0.001f, iters); else k<1><<<148 * 8, 256>>>(out, 0.999f, 0.001f, iters);
      cudaEventRecord(b); cudaEventSynchronize(b);
      float ms; cudaEventElapsedTime(&ms, a, b);
      const double fmas = 148.0 * 8 * 256 * iters * 64;  // scalar FMAs in both modes
      if (rep) printf("%s: %.3f ms, %.1f TFMA/s (%.1f per clk per SM at 1.965 GHz)\n", mode ? "FFMA2" : "FFMA ", ms, fmas / ms / 1e9, fmas / ms / 1e6 / 148 / 1965);
    }
  }
  return 0;
}
