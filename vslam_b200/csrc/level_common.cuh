// Helpers shared by the Gauss-Newton level kernels (gn_level.cu, gn_mom.cu): mbarrier / TMA bulk-copy wrappers and
// small block utilities.
#pragma once
#include "common.cuh"

namespace vb
{
__device__ __forceinline__ uint32_t smem_u32(const void * p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t * bar, uint32_t count)
{
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t * bar, uint32_t bytes)
{
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t * bar)
{
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t * bar, uint32_t parity)
{
  const uint32_t a = smem_u32(bar);
  uint32_t done;
  do {
    asm volatile(
      "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n\tselp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(done)
      : "r"(a), "r"(parity), "r"(0x989680u)  // suspend-time hint: sleep in hardware instead of re-polling
      : "memory");
  } while (!done);
}
// L2 policy for data that is streamed once per phase (the static interest lists): evict first, so that the lines
// which are re-used within an iteration (the gathered target image, the residual codes) stay in L2
__device__ __forceinline__ uint64_t l2_evict_first_policy()
{
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void bulk_g2s_stream(void * dst, const void * src, uint32_t bytes, uint64_t * bar, uint64_t pol)
{
  asm volatile(
    "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
      smem_u32(dst)),
    "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(pol)
    : "memory");
}
// global -> own shared memory, completion counted in bytes on an mbarrier of this CTA
__device__ __forceinline__ void bulk_g2s(void * dst, const void * src, uint32_t bytes, uint64_t * bar)
{
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                 smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}


template <int NT, typename T>
__device__ __forceinline__ void block_copy_n(T * dst, const T * src)
{
  static_assert(sizeof(T) % 4 == 0, "word copy");
  const uint32_t * s = reinterpret_cast<const uint32_t *>(src);
  uint32_t * d = reinterpret_cast<uint32_t *>(dst);
  for (int i = threadIdx.x; i < (int)(sizeof(T) / 4); i += NT) d[i] = s[i];
}

// slice of a view's interest list owned by one of C co-operating CTAs: whole list tiles
__device__ __forceinline__ void slice_of(int n, unsigned C, unsigned rank, int & p0, int & p1)
{
  const int per = (((n + (int)C - 1) / (int)C) + LIST_TILE - 1) / LIST_TILE * LIST_TILE;
  p0 = min(n, (int)rank * per);
  p1 = min(n, p0 + per);
}
}  // namespace vb
