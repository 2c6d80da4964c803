// Moment-based Gauss-Newton level kernel: nearest-pixel sampling with a scaler without offset (identity, t-distribution),
// i.e. the reference's LossTDistribution(ScalerTDistribution) configuration.  The default for STACKED alignment
// (SE3Alignment.cpp:106-146): the cluster's CTAs are split among the job's keyframes, so the views run side by side
// (1 frame against 8 keyframes at 1280x720: 478 alignments/s against 185 with the dense kernel, which walks the views one
// after the other).  For one keyframe per job the dense kernel (gn_level.cu) is faster and stays
// the default; vslamb200_config.level_kernel = VSLAMB200_KERNEL_MOMENTS selects this kernel there too.
//
// One launch per pyramid level; a thread-block CLUSTER per alignment job runs the whole GaussNewton::solve loop of that
// level on device (GaussNewton.cpp:33-102 over InverseCompositional::computeNormalEquations,
// InverseCompositional.cpp:105-154, stacked over the job's keyframes, InverseCompositionalStacked.cpp:39-48).
//
// The inverse-compositional Jacobian of a point is fixed for the level (InverseCompositional.cpp:63-93) and the target is
// sampled at the NEAREST pixel (:119-120), so the residual of a point is an integer in [-255, 255] and, between two
// iterations, only the few points whose rounded target moved change it (1 - 3 % per iteration), while the robust scale -
// and with it every weight - changes every time.  Both facts are used exactly:
//   * MOMENTS.  Per residual magnitude c = |r| the CTA keeps, for its slice of the interest list,
//         M[c][i,j] = sum_{k: |r_k| = c} J_ki J_kj  (21 terms)       M[c][21+i] = sum_{k: |r_k| = c} sign(r_k) J_ki  (6 terms)
//     so that for any weight that is a function of |r| (NormalEquations.cpp:53-61)
//         A = sum_c w(c) M[c][i,j]          b = sum_c w(c) c M[c][21+i]
//     costs 256 x 27 multiply-adds per iteration instead of 27 per point.  The moments are 64-bit FIXED-POINT integers:
//     integer addition is exact and associative, so the sums do not depend on the order in which threads reach them
//     (bit-identical run to run, no floating-point atomics) and a point's old contribution is removed exactly.
//   * FIRST EVALUATION of a level: dense warp + sample of every point (residual codes, histogram), then the moments are
//     built by a counting sort of the slice by residual code and a pass in which every thread accumulates a run of
//     equal-code points in registers (one flush per code run instead of 54 atomics per point).
//   * LATER EVALUATIONS: warp + sample of every point (filtered fp32 evaluation, exact fp64 fallback: warp_filter.cuh),
//     compare with the stored code; the points that changed are queued per warp and move their moments from the old to
//     the new cell, 32 points at a time (one per lane).
//   * EVERY EVALUATION: the view's histogram (summed over its CTAs through distributed shared memory) gives the scale
//     (Scaler.cpp:49-67), the weight table (Loss.cpp:51-92) and chi2; every CTA folds its moments into 27 partial sums;
//     the cluster adds them in rank order and every CTA runs the 6x6 solve redundantly (solve.cuh).
// Two cluster barriers per iteration.  The interest list streams through per-warp shared-memory rings filled by TMA
// bulk copies (cp.async.bulk + mbarrier complete_tx), 10 B per point and iteration: d, (u, v, T) and the old code; the
// target pixels are gathered with asynchronous copies (cp.async) into per-lane shared-memory slots.
#include <cooperative_groups.h>

#include "robust.cuh"
#include "solve.cuh"
#include "warp_filter.cuh"
#include "level_common.cuh"

namespace cg = cooperative_groups;

namespace vb
{
namespace mom
{
constexpr int MT = 256;            // threads per CTA
constexpr int MW = MT / 32;        // warps per CTA
constexpr int CTAS_PER_SM = 2;
constexpr int NC = 256;            // residual magnitudes 0..255
constexpr int NTERM = 27;          // moments per magnitude: 21 (upper A) + 6 (b)
constexpr int CH = 4 * MT;         // interest points per chunk (4 per thread); a warp's share is one list tile
constexpr int NS = 3;              // ring stages per warp (4 measured the same)
constexpr int ST_D = 0, ST_UVT = 512, ST_R = 1024, ST_TILE_BYTES = 1024, ST_R_BYTES = 256, ST_BYTES = 1280;
static_assert(CH / MW == LIST_TILE && TILE_OFF_D == ST_D && TILE_OFF_UVT == ST_UVT && LIST_TILE_BYTES == ST_TILE_BYTES, "tile layout");
constexpr int QB = 32;             // changed points whose moments move together: one per lane
constexpr int QN = 64;            // queue entries per warp (circular): < QB left over + 32 new ones (more: see finish)
constexpr int NROW = NC + 1;       // moment rows: magnitudes 0..255 and a row that absorbs the not-visible code
constexpr double MAGIC = 6755399441055744.0;          // 1.5 * 2^52: x + MAGIC has the integer round(x) in its low mantissa bits
constexpr long long MAGIC_BITS = 0x4338000000000000ll;
constexpr double BSCALE = 1048576.0;                  // 2^20: extra scale of the b moments

// constants of the scaled fp64 Jacobian row (jac_row)
struct JacConsts
{
  double ifx, ify, kx, ky, fxJ, fyJ, s0, s1, s2, s3, s4, s5;
};

struct Shared
{
  JobState J;
  SolveScratch X;
  int hist[HIST_BINS];             // residual histogram of this CTA's slice, kept up to date by the changed points
  int tot[HIST_BINS];              // the view's histogram (sum over its CTAs); during the build: bucket starts
  double red[MW];
  double part[MW][NTERM + 1];
  double scaleJ[6];                // 2^e_i
  double unscale[NTERM];           // 2^-(e_i + e_j), 2^-(e_i + 20)
  double sigma, chi2;
  JacConsts jc;
  alignas(8) uint64_t full[MW][NS];  // per warp and stage: the TMA bytes have landed
  unsigned long long nChanged;
};
// dynamic shared memory: Mlo u32[NROW][NTERM] | Mhi u32[NROW][NTERM] | ring | scratch | queue | gather slots | ViewState[K] | partial[32] | sums[K][28]
constexpr size_t OFF_MHI = (size_t)NROW * NTERM * 4;
constexpr size_t OFF_RING = (2 * OFF_MHI + 127) / 128 * 128;
constexpr size_t OFF_SCRATCH = OFF_RING + (size_t)MW * NS * ST_BYTES;
constexpr size_t SCRATCH_BYTES = 2 * NC * 8;   // bucket cursors (build) | weight tables (fold)
static_assert(SCRATCH_BYTES >= HIST_BINS * 4, "scratch aliases");
constexpr size_t OFF_QUEUE = OFF_SCRATCH + SCRATCH_BYTES;
constexpr size_t OFF_SLOTS = OFF_QUEUE + (size_t)MW * QN * 8;   // gathered words: [warp][register set][lane] x 16 bytes
constexpr size_t OFF_VIEWS = OFF_SLOTS + (size_t)MW * 2 * 32 * 16;
constexpr int PARTIAL_DOUBLES = 32;  // 28 terms + {scale, offset}
__host__ __device__ inline size_t dyn_smem(int K)
{
  return OFF_VIEWS + (size_t)K * sizeof(ViewState) + (size_t)(PARTIAL_DOUBLES + K * NE_TERMS) * sizeof(double);
}

template <typename T>
__device__ __forceinline__ void block_copy(T * dst, const T * src)
{
  block_copy_n<MT>(dst, src);
}

__device__ __forceinline__ double block_sum(double v, double * sm /* >= MW doubles */)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0;
#pragma unroll
  for (int w = 0; w < MW; w++) s += sm[w];
  return s;
}

// which Jacobian components term t multiplies: 21 upper-triangle pairs (row-major, as NE_TERMS orders them), then the 6
// b terms (component x the b scale: j == 6)
__device__ __forceinline__ void term_pair(int t, int & i, int & j)
{
  i = 0;
  j = 0;
  int idx = 0;
#pragma unroll
  for (int a = 0; a < 6; a++)
#pragma unroll
    for (int b = a; b < 6; b++) {
      if (idx == t) {
        i = a;
        j = b;
      }
      idx++;
    }
  if (t >= 21) {
    i = t - 21;
    j = 6;
  }
}

// 64-bit two's-complement add on a {lo, hi} pair of 32-bit shared-memory words, the carry handed over explicitly
// (shared memory has no native 64-bit atomic add).  Exact whatever the interleaving with other adds to the cell.
__device__ __forceinline__ void cell_add(uint32_t * Mlo, uint32_t * Mhi, int cell, long long v)
{
  const uint32_t lo = (uint32_t)(unsigned long long)v, hi = (uint32_t)((unsigned long long)v >> 32);
  const uint32_t old = atomicAdd(&Mlo[cell], lo);
  atomicAdd(&Mhi[cell], hi + ((uint32_t)(old + lo) < old ? 1u : 0u));
}

// Scaled Jacobian row of one list entry in fp64 (WarpSE3::J, Warp.cpp:166-201, contracted with the template gradient,
// InverseCompositional.cpp:79-80), in normalised coordinates xn = x/z, yn = y/z
__device__ __forceinline__ void jac_row(const JacConsts & c, const uint4 rc, double (&J)[6])
{
  const uint32_t pk = rc.x;
  // integer -> double through the 2^52 trick (one full-rate add instead of a conversion); gradients biased by 2^15
  const uint32_t gb = rc.y ^ 0x80008000u;
  const double u = __hiloint2double(0x43300000, (int)(pk & 0xfff)) - 4503599627370496.0;
  const double v = __hiloint2double(0x43300000, (int)((pk >> 12) & 0xfff)) - 4503599627370496.0;
  const double a = (__hiloint2double(0x43300000, (int)(gb & 0xffff)) - 4503599627403264.0) * c.fxJ;
  const double b = (__hiloint2double(0x43300000, (int)(gb >> 16)) - 4503599627403264.0) * c.fyJ;
  const double zi = __hiloint2double((int)rc.w, (int)rc.z);
  const double xn = fma(c.ifx, u, c.kx), yn = fma(c.ify, v, c.ky);
  const double cc = fma(a, xn, b * yn);
  J[0] = (a * zi) * c.s0;
  J[1] = (b * zi) * c.s1;
  J[2] = (-cc * zi) * c.s2;
  J[3] = (-fma(cc, yn, b)) * c.s3;
  J[4] = fma(cc, xn, a) * c.s4;
  J[5] = fma(b, xn, -a * yn) * c.s5;
}

// A warp's stream through its slice of the list: tile c of the warp is entries [w0 + c CH, w0 + c CH + 128) with
// w0 = p0 + warp * 128; lane 0 brings it (and, from the second evaluation on, its residual codes) into stage
// (item + c) % NS of the warp's private ring.  A warp refills a stage right after it has copied it into registers, so no
// warp ever waits for another one (no "empty" barriers, no producer thread that can fall behind).
struct Stream
{
  const unsigned char * tiles;  // first tile of the warp
  const int16_t * codes;        // its residual codes
  unsigned char * ring;         // the warp's NS stages
  uint64_t * bars;              // the warp's NS barriers
  int w0, p1;
};
__device__ __forceinline__ Stream make_stream(Shared & S, unsigned char * ring, const ViewState & V, int p0, int p1, int warp)
{
  Stream t;
  t.w0 = p0 + warp * LIST_TILE;
  t.p1 = p1;
  t.tiles = V.tiles + (size_t)(t.w0 / LIST_TILE) * LIST_TILE_BYTES;
  t.codes = V.r + t.w0;
  t.ring = ring + (size_t)warp * NS * ST_BYTES;
  t.bars = &S.full[warp][0];
  return t;
}
__device__ __forceinline__ void issue_tile(const Stream & t, bool withCodes, int c, int stage, uint64_t pol)
{
  uint64_t * bar = t.bars + stage;
  if (t.w0 + c * CH >= t.p1) {  // nothing of the chunk belongs to this warp: keep the barrier's phase in step
    mbar_arrive(bar);
    return;
  }
  unsigned char * st = t.ring + stage * ST_BYTES;
  mbar_expect_tx(bar, withCodes ? ST_TILE_BYTES + ST_R_BYTES : ST_TILE_BYTES);
  bulk_g2s_stream(st, t.tiles + (size_t)c * (MW * LIST_TILE_BYTES), ST_TILE_BYTES, bar, pol);
  if (withCodes) bulk_g2s(st + ST_R, t.codes + (size_t)c * CH, ST_R_BYTES, bar);
}

// What a chunk leaves in flight between its two pipeline halves.  The gathered target pixels are NOT held in registers:
// each lane gathers the aligned 32-bit words that hold its four target pixels into a 16-byte shared-memory slot with
// asynchronous copies (cp.async, LDGSTS), so nothing waits for them while the next chunk's arithmetic runs - a register
// with a load in flight would be spilled by the compiler under this loop's register pressure, and a spill waits.
struct Pend
{
  uint32_t T;    // the four template values, one byte each
  uint32_t sel;  // per point q, bits 4q..4q+2: byte of the gathered word | 4 when the point is visible
  uint2 old;     // the four stored residual codes
};
__device__ __forceinline__ void cp_async4(void * smemDst, const void * gsrc)
{
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(smemDst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
// all but the most recently committed group of this thread have landed
__device__ __forceinline__ void cp_async_wait_but_one() { asm volatile("cp.async.wait_group 1;" ::: "memory"); }

// first half of a chunk's pipeline: list entries out of the ring, warp, gathers issued
template <bool FIRST>
__device__ __forceinline__ void start_chunk(
  const Stream & t, const ViewState & V, const WarpConsts & c, const uint8_t * __restrict__ I, int nCh, int ch, int stage,
  uint32_t parity, int lane, uint64_t pol, Pend & P, uint32_t * slot /* this lane's four words for this register set */)
{
  const unsigned char * st = t.ring + stage * ST_BYTES;
  mbar_wait(t.bars + stage, parity);
  const int g = t.w0 + ch * CH + 4 * lane;
  if (t.w0 + ch * CH >= t.p1) {  // warp-uniform: nothing of this chunk belongs to the warp
    if (lane == 0 && ch + NS < nCh) issue_tile(t, !FIRST, ch + NS, stage, pol);
    P.T = 0u;
    P.sel = 0u;
    P.old = make_uint2(R_CODE_INVISIBLE * 0x10001u, R_CODE_INVISIBLE * 0x10001u);
    cp_async_commit();  // an empty group: the group count stays in step with the chunks
    return;
  }
  const uint4 pk4 = *reinterpret_cast<const uint4 *>(st + ST_UVT + lane * 16);
  const float4 dd4 = *reinterpret_cast<const float4 *>(st + ST_D + lane * 16);
  if (!FIRST) P.old = *reinterpret_cast<const uint2 *>(st + ST_R + lane * 8);
  // The refill below is an async-proxy write into the stage these generic-proxy loads read, and nothing orders the two
  // proxies: the loads can sit in the shared-memory queue behind another warp's conflicted atomics for longer than the
  // bulk copy takes to arrive (seen: wrong residual codes in a few chunks per thousand jobs).  So the refill is made to
  // depend on the loaded registers: `zero` is 0 (p1 >= 0), but the hardware has to wait for all three loads to form it.
  // (A fence.proxy.async here would also wait for the gathers in flight and serialise the software pipeline: +15 %.)
  const uint32_t zero = (pk4.x ^ __float_as_uint(dd4.x) ^ (FIRST ? 0u : P.old.x)) & (uint32_t)(t.p1 >> 31);
  __syncwarp();
  if (lane == 0 && ch + NS < nCh) issue_tile(t, !FIRST, ch + NS, stage + (int)zero, pol);
  const uint32_t pk[4] = {pk4.x, pk4.y, pk4.z, pk4.w};
  const float dd[4] = {dd4.x, dd4.y, dd4.z, dd4.w};
  bool good[4];
  int pix[4];
  WarpEval ev[4];
#pragma unroll
  for (int q = 0; q < 4; q++) ev[q] = warp_eval(c, pk[q], dd[q]);  // padding lanes compute on stale ring data: harmless
  if (__all_sync(0xffffffffu, ev[0].interior && ev[1].interior && ev[2].interior && ev[3].interior)) {
#pragma unroll
    for (int q = 0; q < 4; q++) pix[q] = warp_pick_interior(c, ev[q], good[q]);
  } else {
#pragma unroll
    for (int q = 0; q < 4; q++) pix[q] = warp_pick_ring(c, ev[q], good[q]);
  }
  if (!(good[0] && good[1] && good[2] && good[3])) {
#pragma unroll
    for (int q = 0; q < 4; q++)
      if (!good[q]) pix[q] = warp_exact(V, pk[q] & 0xfff, (pk[q] >> 12) & 0xfff, dd[q]);
  }
  if (g + 3 >= t.p1) {  // only in the last tile of the slice: entries past the end of the list take no part
#pragma unroll
    for (int q = 0; q < 4; q++) pix[q] = g + q < t.p1 ? pix[q] : -1;
  }
  P.T = __byte_perm(__byte_perm(pk[0], pk[1], 0x0073), __byte_perm(pk[2], pk[3], 0x0073), 0x5410);
  uint32_t sel = 0u;
#pragma unroll
  for (int q = 0; q < 4; q++)
    if (pix[q] >= 0) {
      cp_async4(slot + q, I + (pix[q] & ~3));  // the plane is 256-byte aligned (api.cu, layout_frame)
      sel |= (uint32_t)((pix[q] & 3) | 4) << (4 * q);
    }
  cp_async_commit();
  P.sel = sel;
}
// Second half: the residual codes of a chunk whose gathers have landed (r + 255, or 511 when the point is not visible).
// Call after cp_async_wait_but_one() with exactly one younger group committed.
__device__ __forceinline__ void codes_of(const Pend & P, const uint32_t * slot, uint32_t (&nc)[4])
{
  const uint4 w4 = *reinterpret_cast<const uint4 *>(slot);
  const uint32_t w[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
  for (int q = 0; q < 4; q++) {
    const uint32_t s4 = P.sel >> (4 * q);
    const uint32_t val = (w[q] >> (8 * (s4 & 3))) & 0xff;
    const uint32_t T = (P.T >> (8 * q)) & 0xff;
    nc[q] = (s4 & 4) ? val - T + 255u : (uint32_t)R_CODE_INVISIBLE;
  }
}

__device__ __forceinline__ int thread_id()
{
  unsigned t;
  asm volatile("mov.u32 %0, %%tid.x;" : "=r"(t));  // read once (S2R has a long latency; see gn_level.cu)
  return (int)t;
}

__device__ __forceinline__ WarpConsts warp_consts(const ViewState & V)
{
  WarpConsts c;
  c.r00 = V.rI[0]; c.r01 = V.rI[1]; c.r02 = V.rI[2]; c.r10 = V.rI[3]; c.r11 = V.rI[4]; c.r12 = V.rI[5];
  c.r20 = V.rI[6]; c.r21 = V.rI[7]; c.r22 = V.rI[8];
  c.t0 = V.t32[0]; c.t1 = V.t32[1]; c.t2 = V.t32[2];
  c.fx = V.fx32; c.fy = V.fy32; c.ifx = V.ifxH; c.ify = V.ifyH; c.kx = V.kxH; c.ky = V.kyH;
  c.epsA = V.epsA; c.epsB = V.epsB; c.Wt = V.Wt; c.Ht = V.Ht;
  if (V.Wt < 4 || V.Ht < 4) c.epsB = __int_as_float(0x7f800000);  // no strict interior: always the exact path
  return c;
}

// Every phase of the level loop is its own function: each gets its own register allocation, so the 54 registers of the
// build's accumulators do not spill the warp loop and vice versa.

// ---------------------------------------------------------------------------------------------------------------------
// first evaluation of a level: every point of the slice, dense; residual codes and histogram
// ---------------------------------------------------------------------------------------------------------------------
static __device__ __noinline__ void phase_a_first(
  Shared & S, unsigned char * ring, uint32_t * slots, const ViewState & V, int p0, int p1, int nCh, uint32_t item, uint64_t pol)
{
  const int tid = thread_id(), lane = tid & 31;
  const WarpConsts c = warp_consts(V);
  const Stream t = make_stream(S, ring, V, p0, p1, tid >> 5);
  const uint8_t * __restrict__ I = V.I;
  uint2 * __restrict__ r4 = reinterpret_cast<uint2 *>(V.r);
  uint32_t * slotA = slots + ((tid >> 5) * 2 * 32 + lane) * 4, * slotB = slotA + 32 * 4;
  int stage = (int)(item % NS);
  uint32_t parity = (item / NS) & 1;
  auto start = [&](int ch, Pend & P, uint32_t * slot) {
    start_chunk<true>(t, V, c, I, nCh, ch, stage, parity, lane, pol, P, slot);
    if (++stage == NS) stage = 0, parity ^= 1;
  };
  auto finish = [&](int ch, const Pend & P, const uint32_t * slot) {
    cp_async_wait_but_one();
    uint32_t rr[4];
    codes_of(P, slot, rr);
#pragma unroll
    for (int q = 0; q < 4; q++) atomicAdd(&S.hist[rr[q]], 1);  // branch-free: invisible points land in bin 511, which nobody reads
    // every entry of a tile that holds part of the slice gets a code (entries past the end: not visible), so the
    // later evaluations never see an unwritten code
    const int g = t.w0 + ch * CH;
    if (g < p1) r4[(g >> 2) + lane] = make_uint2(rr[0] | (rr[1] << 16), rr[2] | (rr[3] << 16));
  };
  // Software pipeline over two register sets / gather slots: a chunk's gathers are consumed after the next chunk has
  // been started (every start commits one cp.async group, an empty one when there is nothing left to start).
  // (Measured: one instance of each half with a register copy between them is 14 % slower.)
  Pend A, B;
#pragma unroll 1
  for (int ch = 0; ch <= nCh; ch += 2) {
    if (ch < nCh) start(ch, A, slotA);
    else cp_async_commit();
    if (ch > 0) finish(ch - 1, B, slotB);
    if (ch + 1 < nCh) start(ch + 1, B, slotB);
    else cp_async_commit();
    if (ch < nCh) finish(ch, A, slotA);
  }
}

// ---------------------------------------------------------------------------------------------------------------------
// moments of every visible point of the slice (after the first evaluation)
// ---------------------------------------------------------------------------------------------------------------------
static __device__ __noinline__ void build_moments(
  Shared & S, unsigned char * scratch, uint32_t * Mlo, uint32_t * Mhi, const ViewState & V, int p0, int p1)
{
  const int tid = thread_id(), lane = tid & 31, warp = tid >> 5;
  // counting sort of the slice by residual code: bucket starts from the histogram ...
  int * start = S.tot;                                 // [512]: start[c] = #visible points with a code < c
  int * cursor = reinterpret_cast<int *>(scratch);     // [512]
  if (warp == 0) {
    int loc[16], s = 0;
#pragma unroll
    for (int j = 0; j < 16; j++) {
      const int b = lane * 16 + j;
      loc[j] = b < R_CODE_INVISIBLE ? S.hist[b] : 0;
      s += loc[j];
    }
    int incl = s;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    int run = incl - s;
#pragma unroll
    for (int j = 0; j < 16; j++) {
      start[lane * 16 + j] = run;
      cursor[lane * 16 + j] = run;
      run += loc[j];
    }
  }
  __syncthreads();
  const int nv = start[R_CODE_INVISIBLE];  // visible points of the slice
  // ... list indices scattered into their buckets (the order inside a bucket does not matter: integer sums) ...
  uint32_t * perm = V.perm + p0;
  const uint2 * __restrict__ r4 = reinterpret_cast<const uint2 *>(V.r);
  for (int g = p0 + 4 * tid; g < p1; g += CH) {
    const uint2 cw = r4[g >> 2];
    const uint32_t cd[4] = {cw.x & 0xffff, cw.x >> 16, cw.y & 0xffff, cw.y >> 16};
#pragma unroll
    for (int q = 0; q < 4; q++)
      if (cd[q] != (uint32_t)R_CODE_INVISIBLE) perm[atomicAdd(&cursor[cd[q]], 1)] = (uint32_t)(g + q);
  }
  __syncthreads();
  // ... and every thread accumulates a run of consecutive sorted points in registers, flushing once per code
  const int L = (((nv + MT - 1) / MT) + 3) & ~3;
  int pos = tid * L;
  const int end = min(nv, pos + L);
  if (pos >= end) return;
  int lo = 0, hi = R_CODE_INVISIBLE;  // largest code with start[code] <= pos
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (start[mid] <= pos) lo = mid;
    else hi = mid;
  }
  int code = lo, codeEnd = start[code + 1];
  long long acc[NTERM];
#pragma unroll
  for (int t = 0; t < NTERM; t++) acc[t] = 0ll;
  long long cnt = 0;
  double bs = code < 255 ? -BSCALE : BSCALE;
  auto flush = [&]() {
    const int cell = abs(code - 255) * NTERM;
    const long long bias = cnt * MAGIC_BITS;
#pragma unroll
    for (int t = 0; t < NTERM; t++) {
      cell_add(Mlo, Mhi, cell + t, acc[t] - bias);
      acc[t] = 0ll;
    }
    cnt = 0;
  };
#pragma unroll 1
  for (; pos < end; pos += 2) {
    const uint2 pi2 = *reinterpret_cast<const uint2 *>(perm + pos);
    const uint32_t pi[2] = {pi2.x, pi2.y};
    uint4 rc[2];
#pragma unroll
    for (int q = 0; q < 2; q++)
      if (pos + q < end) rc[q] = *reinterpret_cast<const uint4 *>(V.rec + pi[q]);
#pragma unroll
    for (int q = 0; q < 2; q++) {
      if (pos + q < end) {
        if (pos + q >= codeEnd) {
          flush();
          do code++;
          while (start[code + 1] <= pos + q);
          codeEnd = start[code + 1];
          bs = code < 255 ? -BSCALE : BSCALE;
        }
        double J[6];
        jac_row(S.jc, rc[q], J);
        int s = 0;
#pragma unroll
        for (int i = 0; i < 6; i++) {
#pragma unroll
          for (int j = i; j < 6; j++) {
            acc[s] += __double_as_longlong(fma(J[i], J[j], MAGIC));
            s++;
          }
          acc[21 + i] += __double_as_longlong(fma(J[i], bs, MAGIC));
        }
        cnt++;
      }
    }
  }
  flush();
}

// ---------------------------------------------------------------------------------------------------------------------
// later evaluations: every point is warped and sampled again; the points whose residual changed move their moments
// ---------------------------------------------------------------------------------------------------------------------
// Moves the moments of the first `cnt` (<= 32) queued points of the warp from their old to their new residual cell:
// lane l owns point l - its scaled fp64 Jacobian row, then 27 products, each subtracted from the old cell and added to
// the new one (the not-visible code has a row of its own that nobody reads, so there is no branch).
// Queue entry: {list index, old code | new code << 16}.
__device__ __forceinline__ int row_of(int code) { return (code == R_CODE_INVISIBLE ? NC : abs(code - 255)) * NTERM; }
static __device__ __noinline__ void move_moments(
  const Shared & S, const ViewState & V, uint32_t * Mlo, uint32_t * Mhi, const uint2 * queue, int qh, int cnt)
{
  const int lane = threadIdx.x & 31;
  __syncwarp();
  if (lane < cnt) {
    const uint2 e = queue[(qh + lane) & (QN - 1)];
    const uint4 rc = *reinterpret_cast<const uint4 *>(V.rec + e.x);
    const int o = (int)(e.y & 0xffff), n = (int)(e.y >> 16);
    const int co = row_of(o), cn = row_of(n);
    // b moments carry the sign of the residual (code < 255: negative)
    const double bo = o < 255 ? BSCALE : -BSCALE;  // removes the old contribution
    const double bn = n < 255 ? -BSCALE : BSCALE;  // adds the new one
    double J[6];
    jac_row(S.jc, rc, J);
    int t = 0;
#pragma unroll
    for (int i = 0; i < 6; i++) {
#pragma unroll
      for (int j = i; j < 6; j++) {
        const long long v = __double_as_longlong(fma(J[i], J[j], MAGIC)) - MAGIC_BITS;
        cell_add(Mlo, Mhi, co + t, -v);
        cell_add(Mlo, Mhi, cn + t, v);
        t++;
      }
      cell_add(Mlo, Mhi, co + 21 + i, __double_as_longlong(fma(J[i], bo, MAGIC)) - MAGIC_BITS);
      cell_add(Mlo, Mhi, cn + 21 + i, __double_as_longlong(fma(J[i], bn, MAGIC)) - MAGIC_BITS);
    }
  }
  __syncwarp();
}

// returns the number of points of this warp whose residual changed
static __device__ __noinline__ unsigned phase_a_inc(
  Shared & S, unsigned char * ring, uint32_t * slots, uint2 * queue, uint32_t * Mlo, uint32_t * Mhi, const ViewState & V, int p0,
  int p1, int nCh, uint32_t item, uint64_t pol)
{
  const int tid = thread_id(), lane = tid & 31;
  const WarpConsts c = warp_consts(V);
  const Stream t = make_stream(S, ring, V, p0, p1, tid >> 5);
  const uint8_t * __restrict__ I = V.I;
  int16_t * __restrict__ rcode = V.r;
  uint32_t * slotA = slots + ((tid >> 5) * 2 * 32 + lane) * 4, * slotB = slotA + 32 * 4;
  const unsigned ltmask = (1u << lane) - 1u;
  int stage = (int)(item % NS);
  uint32_t parity = (item / NS) & 1;
  int qh = 0, qe = 0;  // queue head and fill (warp-uniform)
  unsigned nChanged = 0;
  auto start = [&](int ch, Pend & P, uint32_t * slot) {
    start_chunk<false>(t, V, c, I, nCh, ch, stage, parity, lane, pol, P, slot);
    if (++stage == NS) stage = 0, parity ^= 1;
  };
  auto flush = [&](int m) {
    move_moments(S, V, Mlo, Mhi, queue, qh, m);
    qh = (qh + m) & (QN - 1);
    qe -= m;
  };
  auto record = [&](int g, uint32_t o, uint32_t n, int pos) {
    queue[pos & (QN - 1)] = make_uint2((uint32_t)g, o | (n << 16));
    rcode[g] = (int16_t)n;
    atomicSub(&S.hist[o], 1);  // bin 511 (not visible) collects garbage and is never read
    atomicAdd(&S.hist[n], 1);
  };
  auto finish = [&](int ch, const Pend & P, const uint32_t * slot) {
    cp_async_wait_but_one();
    // entries past the end of the slice are not visible now and were stored as not visible by the first
    // evaluation: they never change
    const uint32_t oc[4] = {P.old.x & 0xffff, P.old.x >> 16, P.old.y & 0xffff, P.old.y >> 16};
    uint32_t nc[4];
    codes_of(P, slot, nc);
    int cn = 0;
#pragma unroll
    for (int q = 0; q < 4; q++) cn += nc[q] != oc[q] ? 1 : 0;
    if (__any_sync(0xffffffffu, cn != 0)) {
      const int g = t.w0 + ch * CH + 4 * lane;
      // per-lane counts 0..4 -> warp prefix through three ballots
      const unsigned b0 = __ballot_sync(0xffffffffu, cn & 1), b1 = __ballot_sync(0xffffffffu, cn & 2),
                     b2 = __ballot_sync(0xffffffffu, cn & 4);
      const int tot = __popc(b0) + 2 * __popc(b1) + 4 * __popc(b2);
      if (qe + tot <= QN) {
        if (cn) {
          int pos = qh + qe + __popc(b0 & ltmask) + 2 * __popc(b1 & ltmask) + 4 * __popc(b2 & ltmask);
#pragma unroll
          for (int q = 0; q < 4; q++)
            if (nc[q] != oc[q]) record(g + q, oc[q], nc[q], pos++);
        }
        qe += tot;
      } else {
        // more changes than the queue has room for (a large step): one point per lane at a time
#pragma unroll 1
        for (int q = 0; q < 4; q++) {
          const uint32_t o = q == 0 ? oc[0] : q == 1 ? oc[1] : q == 2 ? oc[2] : oc[3];
          const uint32_t n = q == 0 ? nc[0] : q == 1 ? nc[1] : q == 2 ? nc[2] : nc[3];
          const unsigned m = __ballot_sync(0xffffffffu, n != o);
          if (qe + __popc(m) > QN) flush(qe >= QB ? QB : qe);
          if (n != o) record(g + q, o, n, qh + qe + __popc(m & ltmask));
          qe += __popc(m);
        }
      }
      nChanged += tot;
      while (qe >= QB) flush(QB);
    }
  };
  // software pipeline over two gather slots (see phase_a_first)
  Pend A, B;
#pragma unroll 1
  for (int ch = 0; ch <= nCh; ch += 2) {
    if (ch < nCh) start(ch, A, slotA);
    else cp_async_commit();
    if (ch > 0) finish(ch - 1, B, slotB);
    if (ch + 1 < nCh) start(ch + 1, B, slotB);
    else cp_async_commit();
    if (ch < nCh) finish(ch, A, slotA);
  }
  if (qe > 0) flush(qe);
  return nChanged;
}

// the view's histogram (summed over its CTAs through distributed shared memory), scale, weight tables, chi2, then this
// CTA's 27 partial sums from its moments
static __device__ __noinline__ void scale_and_fold(
  Shared & S, cg::cluster_group & cluster, unsigned char * scratch, const uint32_t * Mlo, const uint32_t * Mhi, ViewState & V,
  const SolverParams & p, double * partial, int myView, int cpv, int srank, int level, long long & tc)
{
  const int tid = thread_id(), lane = tid & 31, warp = tid >> 5;
  auto lap = [&](int slot) {
    if (tid == 0) {
      const long long now = clock64();
      S.J.clk[level][slot] += now - tc;
      tc = now;
    }
  };
  const unsigned r0 = (unsigned)(myView * cpv);
  for (int b = tid; b < HIST_BINS; b += MT) {
    int a = 0;
    for (unsigned rk = r0; rk < r0 + (unsigned)cpv; rk++) a += cluster.map_shared_rank(S.hist, rk)[b];
    S.tot[b] = b == R_CODE_INVISIBLE ? 0 : a;
  }
  __syncthreads();
  const long n = V.nIp;
  if (warp == 0) {
    int vis = 0;
#pragma unroll
    for (int j = 0; j < 16; j++) vis += S.tot[lane * 16 + j];
    vis = __reduce_add_sync(0xffffffffu, vis);
    double sigma = 1.0;
    // invisible points count as residual 0 (InverseCompositional.cpp:111,130)
    if (p.scalerKind == VSLAMB200_SCALER_TDIST) sigma = tdist_sigma_warp(S.tot, (int)(n - vis), n, p.scalerV);
    if (lane == 0) {
      S.sigma = sigma;
      V.scale = sigma;
      V.offset = 0.0;
    }
  }
  __syncthreads();
  double * wA = reinterpret_cast<double *>(scratch);  // w(c)
  double * wB = wA + NC;                              // w(c) c
  double c2 = 0.0;
  {
    const double r = (double)tid;  // MT == NC: one magnitude per thread
    const double w = p.lossKind == VSLAMB200_LOSS_NONE ? 1.0 : loss_weight(p, r / S.sigma);
    const double wr = w * r;
    wA[tid] = w;
    wB[tid] = wr;
    const int cnt = tid == 0 ? 0 : S.tot[255 + tid] + S.tot[255 - tid];
    c2 = (double)cnt * (p.chi2Mode == VSLAMB200_CHI2_WEIGHTED ? wr * r : r * r);
  }
  double chi2 = block_sum(c2, S.red);  // contains the barriers that publish the tables
  if (p.chi2Mode == VSLAMB200_CHI2_REFERENCE && srank == 0) {
    // NormalEquations.cpp:58 as the reference's Release build evaluates it (vslam_b200.h): the weight of the view's first
    // interest point times the plain sum of squares.  This CTA owns the first slice and wrote the code itself.
    double w0 = 0.0;
    if (n > 0) {
      const int code = *reinterpret_cast<const volatile int16_t *>(V.r);
      if (code != R_CODE_INVISIBLE)
        w0 = p.lossKind == VSLAMB200_LOSS_NONE ? 1.0 : loss_weight(p, (double)(code - 255) / S.sigma);
    }
    chi2 *= w0;
  }
  lap(2);
  // A, b of the slice from its moments: warp g folds the magnitudes [32 g, 32 g + 32), lane t owns term t; fixed order
  if (lane < NTERM) {
    const double * w = lane < 21 ? wA : wB;
    double acc = 0.0;
#pragma unroll 4
    for (int cI = warp * (NC / MW); cI < (warp + 1) * (NC / MW); cI++) {
      const long long mv = (long long)(((unsigned long long)Mhi[cI * NTERM + lane] << 32) | (unsigned long long)Mlo[cI * NTERM + lane]);
      acc = fma(w[cI], (double)mv, acc);
    }
    S.part[warp][lane] = acc;
  }
  __syncthreads();
  if (tid < NTERM) {
    double v = 0.0;
#pragma unroll
    for (int g = 0; g < MW; g++) v += S.part[g][tid];
    partial[tid] = v * S.unscale[tid];
  }
  if (tid == NTERM) partial[NTERM] = srank == 0 ? chi2 : 0.0;  // every CTA of the view holds the view's chi2
  if (tid == 28) partial[28] = S.sigma;
  if (tid == 29) partial[29] = 0.0;
  lap(3);
}

__global__ void __launch_bounds__(MT, CTAS_PER_SM) gn_mom_kernel(
  JobState * __restrict__ jobs, ViewState * __restrict__ views, SolverParams p, int level, int maxEvals,
  double * __restrict__ trace, size_t traceJobStride, double * __restrict__ viewSums, int job0, int cpv)
{
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned C = cluster.num_blocks(), rank = cluster.block_rank();
  const int job = job0 + blockIdx.x / C;
  const int tid = thread_id(), lane = tid & 31, warp = tid >> 5;
  __shared__ Shared S;
  extern __shared__ __align__(128) unsigned char dyn[];
  uint32_t * Mlo = reinterpret_cast<uint32_t *>(dyn);
  uint32_t * Mhi = reinterpret_cast<uint32_t *>(dyn + OFF_MHI);
  unsigned char * ring = dyn + OFF_RING;
  unsigned char * scratch = dyn + OFF_SCRATCH;
  uint2 * queue = reinterpret_cast<uint2 *>(dyn + OFF_QUEUE) + warp * QN;
  uint32_t * slots = reinterpret_cast<uint32_t *>(dyn + OFF_SLOTS);

  block_copy(&S.J, jobs + job);
  __syncthreads();
  if (!S.J.active) return;  // uniform over the cluster: nobody writes jobs[job] before the final barrier
  const int K = S.J.nViews;  // the launcher guarantees C == K * cpv
  ViewState * Vs = reinterpret_cast<ViewState *>(dyn + OFF_VIEWS);
  double * partial = reinterpret_cast<double *>(Vs + K);  // [32]: this CTA's 28 terms + {scale, offset} of its view
  double * sums = partial + PARTIAL_DOUBLES;              // [K][28]: view totals
  for (int k = 0; k < K; k++) block_copy(&Vs[k], views + S.J.view0 + k);
  for (int i = tid; i < 2 * NROW * NTERM; i += MT) Mlo[i] = 0u;  // Mlo and Mhi are adjacent
  for (int i = tid; i < HIST_BINS; i += MT) S.hist[i] = 0;
  if (tid == 0) {
    for (int w = 0; w < MW; w++)
      for (int s = 0; s < NS; s++) mbar_init(&S.full[w][s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    S.nChanged = 0ull;
  }
  __syncthreads();
  const int myView = (int)rank / cpv, srank = (int)rank % cpv;
  ViewState & V = Vs[myView];
  int p0, p1;
  slice_of(V.nIp, (unsigned)cpv, (unsigned)srank, p0, p1);
  const int nCh = (p1 - p0 + CH - 1) / CH;
  // Fixed-point scales 2^e_i of the Jacobian components: with |J_i| <= jmax_i over the list (lists.cu) and n points in
  // the slice, |J_i 2^e_i| <= min(2^24, 2^30 / sqrt(n)) keeps every product below 2^48 (the MAGIC trick holds 2^51) and
  // every moment below n (2^30 / sqrt(n))^2 = 2^60; the b moments carry 2^20 more: <= sqrt(n) 2^50.
  if (tid < 6) {
    const double jm = (V.fxJd * (double)__uint_as_float(V.jstat[tid]) + V.fyJd * (double)__uint_as_float(V.jstat[6 + tid])) * 1.001;
    int e = 0;
    if (jm > 0.0 && isfinite(jm)) {
      const double lim = fmin(16777216.0, 1073741824.0 / sqrt((double)max(p1 - p0, 1))) / jm;
      e = ilogb(lim);  // floor(log2(lim))
    }
    e = max(-480, min(480, e));
    S.scaleJ[tid] = ldexp(1.0, e);
  }
  __syncthreads();
  if (tid < NTERM) {
    int i, j;
    term_pair(tid, i, j);
    S.unscale[tid] = 1.0 / (S.scaleJ[i] * (j < 6 ? S.scaleJ[j] : BSCALE));
  }
  if (tid == 0) {
    JacConsts & jc = S.jc;
    jc.ifx = V.ifx; jc.ify = V.ify; jc.kx = V.kx; jc.ky = V.ky; jc.fxJ = V.fxJd; jc.fyJ = V.fyJd;
    jc.s0 = S.scaleJ[0]; jc.s1 = S.scaleJ[1]; jc.s2 = S.scaleJ[2]; jc.s3 = S.scaleJ[3]; jc.s4 = S.scaleJ[4]; jc.s5 = S.scaleJ[5];
  }
  __syncthreads();
  const uint64_t pol = l2_evict_first_policy();
  uint32_t item = 0;  // chunks consumed so far == ring position of the next chunk (uniform over the CTA)
  int evals = 0;
  unsigned myChanged = 0;  // warp-uniform

  const Stream strm = make_stream(S, ring, V, p0, p1, warp);
  if (lane == 0)  // prologue of the first phase A
    for (int c = 0; c < min(NS, nCh); c++) issue_tile(strm, false, c, (int)((item + c) % NS), pol);

  long long tc = clock64();
  auto lap = [&](int slot) {
    if (tid == 0) {
      const long long now = clock64();
      S.J.clk[level][slot] += now - tc;
      tc = now;
    }
  };

  for (;;) {
    if (evals == 0) {
      phase_a_first(S, ring, slots, V, p0, p1, nCh, item, pol);
      __syncthreads();
      lap(7);
      build_moments(S, scratch, Mlo, Mhi, V, p0, p1);
      lap(8);
    } else {
      myChanged += phase_a_inc(S, ring, slots, queue, Mlo, Mhi, V, p0, p1, nCh, item, pol);
      lap(0);
    }
    item += nCh;
    // the changed residual codes were written through the generic proxy and are read back by TMA (async proxy); a warp
    // reads back exactly the entries it wrote, so its prologue of the next phase A can start as soon as the warp is done:
    // the list does not depend on the pose, so it streams in behind the barriers, the scale and the solve
    asm volatile("fence.proxy.async;" ::: "memory");
    __syncwarp();
    if (lane == 0)
      for (int cc = 0; cc < min(NS, nCh); cc++) issue_tile(strm, true, cc, (int)((item + cc) % NS), pol);
    // ------------------------------------------------------------------ the view's histogram, scale, weights, chi2
    __syncthreads();
    cluster.sync();
    lap(1);
    scale_and_fold(S, cluster, scratch, Mlo, Mhi, V, p, partial, myView, cpv, srank, level, tc);
    // ------------------------------------------------------------------ cluster reduction + solve
    __syncthreads();
    cluster.sync();
    lap(4);
    for (int i = tid; i < K * NE_TERMS; i += MT) {
      const int k = i / NE_TERMS, t = i - k * NE_TERMS;
      double v = 0.0;
      for (int rk = k * cpv; rk < (k + 1) * cpv; rk++) v += cluster.map_shared_rank(partial, (unsigned)rk)[t];
      sums[i] = v;
      if (viewSums && rank == 0) viewSums[(size_t)(S.J.view0 + k) * NE_TERMS + t] = v;
    }
    if (tid >= 32 && tid < 32 + K && tid - 32 != myView) {
      const double * pr = cluster.map_shared_rank(partial, (unsigned)((tid - 32) * cpv));
      Vs[tid - 32].scale = pr[28];
      Vs[tid - 32].offset = pr[29];
    }
    __syncthreads();
    lap(5);
    evals++;
    if (tid == 0) {
      S.J.evals[level]++;
      S.J.pointIters += (unsigned long long)S.J.totalPoints;
    }
    solver_step_block(S.J, Vs, sums, S.X, p, level, (trace && rank == 0) ? trace + (size_t)job * traceJobStride : nullptr,
                      evals >= maxEvals);
    lap(6);
    if (!S.J.active) break;
  }
  // drain the speculative prologue of the phase A that will not run: no bulk copy may be in flight at exit
  for (int cc = 0; cc < min(NS, nCh); cc++) mbar_wait(&S.full[warp][(item + cc) % NS], ((item + cc) / NS) & 1);
  if (lane == 0) atomicAdd(&S.nChanged, (unsigned long long)myChanged);
  __syncthreads();
  cluster.sync();  // nobody leaves while its shared memory may still be read remotely
  if (rank == 0 && tid == 0) {
    unsigned long long v = 0ull;
    for (unsigned rk = 0; rk < C; rk++) v += cluster.map_shared_rank(&S.nChanged, rk)[0];
    S.J.clk[level][9] += (long long)v;
  }
  cluster.sync();
  if (rank == 0) {
    block_copy(jobs + job, &S.J);
    for (int k = 0; k < K; k++) block_copy(views + S.J.view0 + k, &Vs[k]);
  }
}
}  // namespace mom

// Moment-based level kernel: K keyframes per job (the same K for every job of the launch), cpv CTAs per keyframe,
// cluster = K * cpv CTAs; scalers without offset only.
cudaError_t launch_gn_mom(
  JobState * jobs, ViewState * views, int job0, int nJobs, int K, int cpv, SolverParams p, int level, int maxEvals,
  double * trace, size_t traceJobStride, double * viewSums, cudaStream_t s)
{
  if (nJobs <= 0) return cudaSuccess;
  const size_t dyn = mom::dyn_smem(K);
  cudaError_t e = cudaFuncSetAttribute(mom::gn_mom_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
  if (e != cudaSuccess) return e;
  int clusterSize = K * cpv;
  if (clusterSize > 8) {
    e = cudaFuncSetAttribute(mom::gn_mom_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e != cudaSuccess) return e;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(mom::MT, 1, 1);
  cfg.dynamicSmemBytes = dyn;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  if (clusterSize > 8) {
    // clusters beyond the portable size need a GPC that can co-schedule them: asked once per configuration
    static int okBig = -1, okSize = 0;
    static size_t okDyn = 0;
    if (okBig < 0 || okDyn != dyn || okSize != clusterSize) {
      cfg.gridDim = dim3(clusterSize, 1, 1);
      at[0].val.clusterDim.x = clusterSize;
      int n = 0;
      const cudaError_t q = cudaOccupancyMaxActiveClusters(&n, mom::gn_mom_kernel, &cfg);
      if (q != cudaSuccess) cudaGetLastError();
      okBig = (q == cudaSuccess && n >= 1) ? 1 : 0;
      okDyn = dyn;
      okSize = clusterSize;
    }
    if (!okBig) {
      while (K * cpv > 8 && cpv > 1) cpv >>= 1;
      clusterSize = K * cpv;
      if (clusterSize > 8) return cudaErrorInvalidConfiguration;  // the caller only asks for K <= 8
    }
  }
  cfg.gridDim = dim3((unsigned)nJobs * clusterSize, 1, 1);
  at[0].val.clusterDim.x = clusterSize;
  return cudaLaunchKernelEx(&cfg, mom::gn_mom_kernel, jobs, views, p, level, maxEvals, trace, traceJobStride, viewSums, job0, cpv);
}
size_t gn_mom_smem(int K) { return mom::dyn_smem(K) + sizeof(mom::Shared); }
}  // namespace vb
