// One launch per pyramid level: a thread-block CLUSTER per alignment job runs the whole
// GaussNewton::solve loop of that level on device (GaussNewton.cpp:33-102 over
// InverseCompositional::computeNormalEquations, InverseCompositional.cpp:105-154, stacked over the job's
// keyframes, InverseCompositionalStacked.cpp:39-48).  No host round trip inside the loop.
//
// Per Gauss-Newton iteration and per keyframe ("view") every CTA of the cluster owns a contiguous slice
// of the view's interest list:
//   phase A  SE3 warp + nearest sample + integer residual r (2 B per point, written to the view's r array)
//            + shared-memory residual histogram
//   barrier  cluster barrier; every CTA sums the cluster's histograms through distributed shared memory and
//            derives the robust scale and the 511-entry weight table from it (Scaler.cpp:21-67,
//            Loss.cpp:51-92) - residuals are u8 - u8 integers, so every scaler is a function of the histogram
//   phase C  1x6 Jacobian (fp64), weight lookup, 21 + 6 + 1 products (fp64), fixed-order block reduction
//   barrier  cluster barrier; every CTA sums the cluster's block partials in rank order (deterministic) and
//            runs the 6x6 solve / pose update / termination logic redundantly (bitwise identical), so no
//            broadcast of the new pose is needed.
// The interest list is static for the level, so both phases stream it through a 4-stage shared-memory ring
// filled by TMA bulk copies (cp.async.bulk + mbarrier complete_tx): the first chunks of the next phase are
// issued before the barrier / scale / solve sections, which hides the HBM latency behind them.
// Everything that feeds A, b and chi2 is fp64: the nearest-neighbour sampling makes the iteration chaotic
// (one pixel switching its rounded target changes b by ~1 % near convergence), so matching the CPU
// aligner's iterates and stopping decisions needs ~1e-13 relative accuracy, not fp32's 1e-7.
#include <cooperative_groups.h>

#include "robust.cuh"
#include "solve.cuh"
#include "warp_filter.cuh"
#include "level_common.cuh"

namespace cg = cooperative_groups;

namespace vb
{
#ifndef DENSE_GT
#define DENSE_GT 128
#endif
constexpr int GT = DENSE_GT;         // threads per CTA (256 x 2 per SM measured the same as 128 x 4)
constexpr int BPT = HIST_BINS / GT;  // histogram bins per thread
constexpr int CTAS_PER_SM = 512 / GT;

__device__ __forceinline__ double block_sum(double v, double * sm /* >= GT/32 doubles */)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0;
#pragma unroll
  for (int w = 0; w < GT / 32; w++) s += sm[w];
  return s;
}


// bin b holds the residual value b - 255 (bin 511 is unused).  Thread t owns bins (t + j*GT + 192) mod 512:
// j = 0 covers the central residuals -63..64, where nearly all the mass is, so the other j are skipped warp-wide
__device__ __forceinline__ int bin_of(int j) { return ((int)threadIdx.x + j * GT + 192) & (HIST_BINS - 1); }

// Scale / offset from the residual histogram (h[b] = #visible points with r = b-255; invisible points
// count as r = 0, InverseCompositional.cpp:111,130).  Whole block (the leading CTA of the cluster runs it).
// nvis: this thread's bins before the zeros of the invisible points are added (for chi2).
__device__ void compute_scale(
  ViewState & V, const SolverParams & p, int * h /* smem 512 */, double * red, double (&nvis)[BPT], double & scaleOut,
  double & offsetOut)
{
  const int tid = threadIdx.x;
  __shared__ double s_scale, s_offset;
  int bin[BPT];
#pragma unroll
  for (int j = 0; j < BPT; j++) bin[j] = bin_of(j);
  int c0 = 0;
#pragma unroll
  for (int j = 0; j < BPT; j++) c0 += h[bin[j]];
  const double vis = block_sum((double)c0, red);
  const long n = V.nIp;
#pragma unroll
  for (int j = 0; j < BPT; j++) nvis[j] = (double)h[bin[j]];
  __syncthreads();
  if (tid == 0) h[255] += (int)(n - (long)vis);  // zeros of the invisible points
  __syncthreads();
  double scale = 1.0, offset = 0.0;
  double rv[BPT], nv[BPT];
#pragma unroll
  for (int j = 0; j < BPT; j++) {
    rv[j] = (double)(bin[j] - 255);
    nv[j] = (double)h[bin[j]];
  }
  if (p.scalerKind == VSLAMB200_SCALER_TDIST) {
    // sigma^2 <- mean(r^2 (nu+1) / (nu + (r/sigma)^2)), from sigma = 1, <= 30 steps, until |step| <= 1e-5
    // (Scaler.cpp:49-67).  The steps are strictly sequential, so one warp runs them without block barriers:
    // lane l owns the residual values whose offset from 0 is l + 32 m (both signs folded: the summand is even
    // in r), slot m = 0..7; slots nobody populates are skipped warp-wide.
    // r^2 (nu+1) / (nu + r^2/sigma^2) == r^2 (nu+1) sigma^2 / (nu sigma^2 + r^2): one division per slot.
    __syncthreads();
    if (tid < 32) scale = tdist_sigma_warp(h, 0, n, p.scalerV);
  } else if (p.scalerKind == VSLAMB200_SCALER_MEAN || p.scalerKind == VSLAMB200_SCALER_MEDIAN) {
    if (n > 0) {
      if (p.scalerKind == VSLAMB200_SCALER_MEAN) {
        double t = 0.0;
#pragma unroll
        for (int j = 0; j < BPT; j++) t += nv[j] * rv[j];
        const double sum = block_sum(t, red);  // exact: integers < 2^53
        offset = sum / (double)n;
      } else {
        if (tid == 0) {
          // order statistics of the sorted residual vector (algorithm.cpp:109-120)
          long ka, kb;
          if (n % 2 == 0) {
            ka = n / 2 - 1;
            kb = n / 2 + 1;
            if (n == 2) ka = kb = 0;
          } else {
            ka = kb = n / 2;
          }
          long cum = 0;
          double va = 0, vb2 = 0;
          bool fa = false, fb = false;
          for (int b = 0; b < 511 && !(fa && fb); b++) {
            cum += h[b];
            if (!fa && cum > ka) {
              va = b - 255;
              fa = true;
            }
            if (!fb && cum > kb) {
              vb2 = b - 255;
              fb = true;
            }
          }
          s_offset = (n % 2 == 0 && n != 2) ? (va + vb2) / 2 : va;
        }
        __syncthreads();
        offset = s_offset;
      }
      double t = 0.0;
#pragma unroll
      for (int j = 0; j < BPT; j++) t += nv[j] * fabs(rv[j] - offset);
      const double a = block_sum(t, red);
      scale = sqrt(a / (double)(n - 1));
    }
  }
  if (tid == 0) {
    s_scale = scale;
    s_offset = offset;
    V.scale = scale;
    V.offset = offset;
  }
  __syncthreads();
  scaleOut = s_scale;
  offsetOut = s_offset;
}

// w_0 of VSLAMB200_CHI2_REFERENCE: the weight of the first interest point under the evaluation's scale (0 when it is not
// visible or the list is empty).  V.r[0] was written in phase A by the cluster's leading CTA.
__device__ __forceinline__ double first_point_weight(const SolverParams & p, const ViewState & V, double scale, double offset)
{
  if (V.nIp < 1) return 0.0;
  const int code = *reinterpret_cast<const volatile int16_t *>(V.r);
  if (code == R_CODE_INVISIBLE) return 0.0;
  return p.lossKind == VSLAMB200_LOSS_NONE ? 1.0 : loss_weight(p, ((double)(code - 255) - offset) / scale);
}

// Weight table {sqrt(w(r)), sqrt(w(r)) r} per residual code (Loss.cpp:51-92); code 511 (not visible) -> {0, 0}.
// Returns this thread's share of sum_k w_k r_k^2 (VSLAMB200_CHI2_WEIGHTED) or of sum_k r_k^2 (VSLAMB200_CHI2_REFERENCE,
// multiplied by the first point's weight afterwards: NormalEquations.cpp:58, see vslam_b200.h) given its visible counts.
__device__ __forceinline__ double build_weights(
  const SolverParams & p, double scale, double offset, double2 * lut, const double (&nvis)[BPT])
{
  double c2 = 0.0;
#pragma unroll
  for (int j = 0; j < BPT; j++) {
    const int b = bin_of(j);
    const double r = (double)(b - 255);
    const double w = p.lossKind == VSLAMB200_LOSS_NONE ? 1.0 : loss_weight(p, (r - offset) / scale);
    const double wr = w * r;
    // the table holds {sqrt(w), sqrt(w) r}: phase C scales the template gradient by sqrt(w) once and forms
    // (sqrt(w) J)(sqrt(w) J)^T, which saves the six w * J_i products per point
    const double sw = sqrt(w);
    lut[b] = b == HIST_BINS - 1 ? make_double2(0.0, 0.0) : make_double2(sw, sw * r);
    c2 = fma(nvis[j], p.chi2Mode == VSLAMB200_CHI2_WEIGHTED ? wr * r : r * r, c2);
  }
  return c2;
}


// ------------------------------------------------------------------------------------------------
// shared-memory ring fed by TMA bulk copies
// ------------------------------------------------------------------------------------------------
constexpr int CH = 4 * GT;                  // interest points per chunk (4 per thread)
constexpr int NS = 4;                       // ring stages
constexpr int WCH = CH / (GT / 32);         // points of a chunk owned by one warp == one list tile
static_assert(WCH == LIST_TILE, "a warp's share of a chunk is one list tile");
// one stage of a warp's private ring.  Phase A: the tile [d | uvt] (1024 B).  Phase C: the tile's 128 records
// {uvt, g, zi} (2048 B) followed by its 128 residual codes.
constexpr int A_OFF_D = 0, A_OFF_UVT = 512, A_BYTES = 1024;
constexpr int C_OFF_R = 2048, C_TILE_BYTES = 2048, C_R_BYTES = 2 * WCH;
constexpr int WSTAGE_BYTES = C_OFF_R + C_R_BYTES;
constexpr int STAGE_BYTES = WSTAGE_BYTES * (GT / 32);
static_assert(TILE_OFF_D == A_OFF_D && TILE_OFF_UVT == A_OFF_UVT && LIST_TILE_BYTES == A_BYTES &&
              sizeof(ListRec) * LIST_TILE == C_TILE_BYTES, "tile layout");

struct LevelShared
{
  JobState J;
  int hist[2][HIST_BINS];
  double2 lut[HIST_BINS];  // {sqrt(w(r)), sqrt(w(r)) r} per residual code; code 511 (not visible) -> {0, 0}
  double red[GT / 32 * NE_TERMS];
  SolveScratch X;
  double so[2];  // {scale, offset} of the current view, written by the cluster's leading CTA
  alignas(8) uint64_t full[GT / 32][NS];  // per warp and stage: the TMA bytes have landed
};
// dynamic shared memory: ring | ViewState[K] | partial[2][K][28] | sums[K][28]
__host__ __device__ inline size_t level_dyn_smem(int K)
{
  return (size_t)NS * STAGE_BYTES + (size_t)K * (sizeof(ViewState) + 3 * NE_TERMS * sizeof(double));
}

template <typename T>
__device__ __forceinline__ void block_copy(T * dst, const T * src)
{
  block_copy_n<GT>(dst, src);
}

// Every warp owns a private ring: lane 0 streams the warp's quarter of chunk c of the slice [p0, p1) into stage
// (item % NS) of that ring.  A warp refills a stage right after it has copied it into registers, so no warp ever
// waits for another one (no "empty" barriers, no producer thread that can fall behind).
constexpr int CHUNK_A = 0, CHUNK_C = 1;  // what a ring stage is filled with
__device__ __forceinline__ void issue_chunk(
  LevelShared & S, unsigned char * ring, const ViewState & V, int kind, int p0, int p1, int c, uint32_t item, int warp,
  uint64_t pol)
{
  const int start = p0 + c * CH + warp * WCH;
  uint64_t * bar = &S.full[warp][item % NS];
  if (start >= p1) {  // nothing of the chunk belongs to this warp: keep the barrier's phase in step with `item`
    mbar_arrive(bar);
    return;
  }
  unsigned char * st = ring + (size_t)(warp * NS + item % NS) * WSTAGE_BYTES;
  if (kind == CHUNK_C) {
    mbar_expect_tx(bar, C_TILE_BYTES + C_R_BYTES);
    bulk_g2s_stream(st, V.rec + start, C_TILE_BYTES, bar, pol);
    bulk_g2s(st + C_OFF_R, V.r + start, C_R_BYTES, bar);
  } else {
    mbar_expect_tx(bar, A_BYTES);
    bulk_g2s_stream(st, V.tiles + (size_t)(start / LIST_TILE) * LIST_TILE_BYTES, A_BYTES, bar, pol);
  }
}

__global__ void __launch_bounds__(GT, CTAS_PER_SM) gn_level_kernel(
  JobState * __restrict__ jobs, ViewState * __restrict__ views, SolverParams p, int level, int maxEvals,
  double * __restrict__ trace, size_t traceJobStride, double * __restrict__ viewSums, int job0)
{
  cg::cluster_group cluster = cg::this_cluster();
  const unsigned C = cluster.num_blocks(), rank = cluster.block_rank();
  const int job = job0 + blockIdx.x / C;
  // read once through a volatile asm: the compiler otherwise re-reads the special register in every chunk (S2R has a
  // long latency and sits in the hot loops)
  unsigned tidv;
  asm volatile("mov.u32 %0, %%tid.x;" : "=r"(tidv));
  const int tid = (int)tidv, lane = tid & 31, warp = tid >> 5;
  __shared__ LevelShared S;
  extern __shared__ __align__(128) unsigned char ring[];

  block_copy(&S.J, jobs + job);
  __syncthreads();
  if (!S.J.active) return;  // uniform over the cluster: nobody writes jobs[job] before the final barrier
  const int K = S.J.nViews;
  ViewState * Vs = reinterpret_cast<ViewState *>(ring + (size_t)NS * STAGE_BYTES);
  double * partial = reinterpret_cast<double *>(Vs + K);       // [2][K][28]: this CTA's block sums, double buffered
  double * sums = partial + 2 * K * NE_TERMS;                  // [K][28]: cluster totals
  for (int k = 0; k < K; k++) block_copy(&Vs[k], views + S.J.view0 + k);
  if (tid == 0) {
    for (int w = 0; w < GT / 32; w++)
      for (int s = 0; s < NS; s++) mbar_init(&S.full[w][s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  int hb = 0, evals = 0;
  const uint64_t pol = l2_evict_first_policy();
  uint32_t item = 0;  // chunks consumed so far == ring position of the next chunk (uniform over the CTA)

  // prologue of the very first phase A
  {
    int p0, p1;
    slice_of(Vs[0].nIp, C, rank, p0, p1);
    const int nCh = (p1 - p0 + CH - 1) / CH;
    if (lane == 0)
      for (int c = 0; c < min(NS, nCh); c++) issue_chunk(S, ring, Vs[0], CHUNK_A, p0, p1, c, item + c, warp, pol);
  }

  long long tc = clock64();
  auto lap = [&](int slot) {
    if (tid == 0) {
      const long long now = clock64();
      S.J.clk[level][slot] += now - tc;
      tc = now;
    }
  };
  long long ts = 0;  // sub-laps of the scale section (do not reset the main lap clock)
  auto sublap = [&](int slot) {
    if (tid == 0) {
      const long long now = clock64();
      S.J.clk[level][slot] += now - ts;
      ts = now;
    }
  };
  for (;;) {
    const int pb = evals & 1;
    for (int k = 0; k < K; k++) {
      ViewState & V = Vs[k];
      int p0, p1;
      slice_of(V.nIp, C, rank, p0, p1);
      const int nCh = (p1 - p0 + CH - 1) / CH;
      int * hist = S.hist[hb];
#pragma unroll
      for (int j = 0; j < BPT; j++) hist[tid + j * GT] = 0;
      __syncthreads();
      // ---------------------------------------------------------------- phase A
      {
        WarpConsts c;
        c.r00 = V.rI[0]; c.r01 = V.rI[1]; c.r02 = V.rI[2]; c.r10 = V.rI[3]; c.r11 = V.rI[4]; c.r12 = V.rI[5];
        c.r20 = V.rI[6]; c.r21 = V.rI[7]; c.r22 = V.rI[8];
        c.t0 = V.t32[0]; c.t1 = V.t32[1]; c.t2 = V.t32[2];
        c.fx = V.fx32; c.fy = V.fy32; c.ifx = V.ifxH; c.ify = V.ifyH; c.kx = V.kxH; c.ky = V.kyH;
        c.epsA = V.epsA; c.epsB = V.epsB; c.Wt = V.Wt; c.Ht = V.Ht;
        if (V.Wt < 4 || V.Ht < 4) c.epsB = __int_as_float(0x7f800000);  // no strict interior: always the exact path
        const uint8_t * __restrict__ I = V.I;
        uint2 * __restrict__ r4 = reinterpret_cast<uint2 *>(V.r);
        // Software pipeline: the gathers of chunk ch are issued, then chunk ch-1 (whose gathers have had a whole
        // chunk of arithmetic to arrive) is finished: residual, histogram, store.
        struct Pend
        {
          int g, pix[4], val[4];
          uint32_t T[4];
        };
        auto start_chunk = [&](int ch, Pend & P) {
          const uint32_t it = item + ch;
          P.g = p0 + ch * CH + 4 * tid;
          const unsigned char * st = ring + (size_t)(warp * NS + it % NS) * WSTAGE_BYTES;
          mbar_wait(&S.full[warp][it % NS], (it / NS) & 1);
          if (p0 + ch * CH + warp * WCH >= p1) {  // warp-uniform: nothing of this chunk belongs to the warp
            if (lane == 0 && ch + NS < nCh) issue_chunk(S, ring, V, CHUNK_A, p0, p1, ch + NS, it + NS, warp, pol);
#pragma unroll
            for (int q = 0; q < 4; q++) P.pix[q] = -1, P.val[q] = 0, P.T[q] = 0;
            return;
          }
          const uint4 pk4 = *reinterpret_cast<const uint4 *>(st + A_OFF_UVT + lane * 16);
          const float4 dd4 = *reinterpret_cast<const float4 *>(st + A_OFF_D + lane * 16);
          // the refill (an async-proxy write into the stage just read) is made to depend on the loaded registers: zero
          // is 0, but the hardware has to wait for both loads to form it (see gn_mom.cu, start_chunk)
          const int zero = (int)((pk4.x ^ __float_as_uint(dd4.x)) & (uint32_t)(p1 >> 31));
          __syncwarp();
          if (lane == 0 && ch + NS < nCh) issue_chunk(S, ring, V, CHUNK_A, p0, p1 + zero, ch + NS, it + NS, warp, pol);
          const uint32_t pk[4] = {pk4.x, pk4.y, pk4.z, pk4.w};
          const float dd[4] = {dd4.x, dd4.y, dd4.z, dd4.w};
          bool good[4];
          WarpEval ev[4];
#pragma unroll
          for (int q = 0; q < 4; q++) {
            P.T[q] = pk[q] >> 24;
            ev[q] = warp_eval(c, pk[q], dd[q]);  // padding lanes compute on stale ring data: harmless
          }
          if (__all_sync(0xffffffffu, ev[0].interior && ev[1].interior && ev[2].interior && ev[3].interior)) {
#pragma unroll
            for (int q = 0; q < 4; q++) P.pix[q] = warp_pick_interior(c, ev[q], good[q]);
          } else {
#pragma unroll
            for (int q = 0; q < 4; q++) P.pix[q] = warp_pick_ring(c, ev[q], good[q]);
          }
          if (!(good[0] && good[1] && good[2] && good[3])) {
#pragma unroll
            for (int q = 0; q < 4; q++)
              if (!good[q]) P.pix[q] = warp_exact(V, pk[q] & 0xfff, (pk[q] >> 12) & 0xfff, dd[q]);
          }
          if (P.g + 3 >= p1) {  // only in the last tile of the slice: entries past the end of the list take no part
#pragma unroll
            for (int q = 0; q < 4; q++) P.pix[q] = P.g + q < p1 ? P.pix[q] : -1;
          }
#pragma unroll
          for (int q = 0; q < 4; q++) P.val[q] = P.pix[q] >= 0 ? (int)__ldg(I + P.pix[q]) : 0;
        };
        auto finish_chunk = [&](const Pend & P) {
          uint32_t rr[4];
#pragma unroll
          for (int q = 0; q < 4; q++) {
            const int r = P.val[q] - (int)P.T[q];
            rr[q] = P.pix[q] >= 0 ? (uint32_t)(r + 255) : (uint32_t)R_CODE_INVISIBLE;  // residual code
            atomicAdd(&hist[rr[q]], 1);  // branch-free: invisible points land in bin 511, which nobody reads
          }
          if (P.g < p1) r4[P.g >> 2] = make_uint2(rr[0] | (rr[1] << 16), rr[2] | (rr[3] << 16));
        };
        // one instance of each half of the pipeline (instruction-cache footprint): the chunk started in this turn
        // is finished in the next one
        Pend P, Pn;
#pragma unroll 1
        for (int ch = 0; ch <= nCh; ch++) {
          if (ch < nCh) start_chunk(ch, Pn);
          if (ch > 0) finish_chunk(P);
          P = Pn;
        }
        item += nCh;
      }
      // the residuals were written through the generic proxy and are read back by TMA (async proxy); a warp reads
      // back exactly the entries it wrote, so its phase-C prologue can start as soon as the warp itself is done
      asm volatile("fence.proxy.async;" ::: "memory");
      __syncwarp();
      if (lane == 0)
        for (int c = 0; c < min(NS, nCh); c++) issue_chunk(S, ring, V, CHUNK_C, p0, p1, c, item + c, warp, pol);
      lap(0);
      // ---------------------------------------------------------------- scale + weight table
      // The leading CTA sums the cluster's histograms, runs the scaler and publishes {scale, offset} into every
      // CTA's shared memory; the others only wait (they would repeat the same arithmetic, and the kernel is bound
      // by instruction issue).  Every CTA then builds its own weight table.
      double chi2View = 0.0;
      {
        cluster.sync();
        lap(1);
        if (tid == 0) ts = clock64();
        double nvis[BPT];
#pragma unroll
        for (int j = 0; j < BPT; j++) nvis[j] = 0.0;
        if (rank == 0) {
          int a[BPT];
#pragma unroll
          for (int j = 0; j < BPT; j++) a[j] = 0;
          for (unsigned rk = 0; rk < C; rk++) {
            const int * h = cluster.map_shared_rank(hist, rk);
#pragma unroll
            for (int j = 0; j < BPT; j++) a[j] += h[tid + j * GT];
          }
          // the other histogram buffer is idle (its last remote readers passed a cluster barrier since)
          int * tot = S.hist[hb ^ 1];
#pragma unroll
          for (int j = 0; j < BPT; j++) tot[tid + j * GT] = tid + j * GT == R_CODE_INVISIBLE ? 0 : a[j];
          __syncthreads();
          double sc, of;
          compute_scale(V, p, tot, S.red, nvis, sc, of);
          if (tid < (int)C) {
            double * dst = cluster.map_shared_rank(S.so, tid);
            dst[0] = sc;
            dst[1] = of;
          }
        }
        sublap(7);
        cluster.sync();
        sublap(8);
        const double c2 = build_weights(p, S.so[0], S.so[1], S.lut, nvis);
        if (rank == 0) {
          chi2View = block_sum(c2, S.red);  // contains the barrier that publishes the table
          if (p.chi2Mode == VSLAMB200_CHI2_REFERENCE)
            chi2View *= first_point_weight(p, V, S.so[0], S.so[1]);  // this CTA wrote the code in phase A
        } else {
          __syncthreads();
        }
        sublap(9);
        hb ^= 1;
        lap(2);
      }
      // ---------------------------------------------------------------- phase C
      double acc[NE_TERMS];
#pragma unroll
      for (int s = 0; s < NE_TERMS; s++) acc[s] = 0.0;
      {
        const double ifx = V.ifx, ify = V.ify, kx = V.kx, ky = V.ky, fxJ = V.fxJd, fyJ = V.fyJd;
        const double2 * __restrict__ lut = S.lut;
#pragma unroll 1
        for (int ch = 0; ch < nCh; ch++, item++) {
          const unsigned char * st = ring + (size_t)(warp * NS + item % NS) * WSTAGE_BYTES;
          mbar_wait(&S.full[warp][item % NS], (item / NS) & 1);
          if (p0 + ch * CH + warp * WCH >= p1) {  // warp-uniform: nothing of this chunk belongs to the warp
            if (lane == 0 && ch + NS < nCh) issue_chunk(S, ring, V, CHUNK_C, p0, p1, ch + NS, item + NS, warp, pol);
            continue;
          }
          // lane l takes the tile's records l, l + 32, l + 64, l + 96 (16-byte records: conflict-free 128-bit loads)
          uint4 rc[4];
          int rq[4];
#pragma unroll
          for (int q = 0; q < 4; q++) {
            rc[q] = *reinterpret_cast<const uint4 *>(st + (q * 32 + lane) * 16);
            // residual codes: r + 255, or 511 when not visible.  Entries past the end of the list (padding of the last
            // tile) may hold anything: masked to a valid table index, and their Jacobian is exactly zero (lists.cu)
            rq[q] = (int)(*reinterpret_cast<const uint16_t *>(st + C_OFF_R + (q * 32 + lane) * 2)) & 0x1ff;
          }
          const int zero = (int)((rc[0].x ^ rc[1].x ^ rc[2].x ^ rc[3].x ^ (uint32_t)(rq[0] ^ rq[1] ^ rq[2] ^ rq[3])) & (uint32_t)(p1 >> 31));
          __syncwarp();
          if (lane == 0 && ch + NS < nCh) issue_chunk(S, ring, V, CHUNK_C, p0, p1 + zero, ch + NS, item + NS, warp, pol);
#pragma unroll
          for (int q = 0; q < 4; q++) {
            // branch-free: a point that is not visible gets weight 0, a padding entry has a zero Jacobian; both add
            // exact zeros - the oracle skips w == 0 points the same way.
            // WarpSE3::J rows (Warp.cpp:184-199) contracted with the template gradient
            // (InverseCompositional.cpp:79-80), in normalised coordinates xn = x/z, yn = y/z.
            // Integer -> double through the 2^52 trick (one full-rate add instead of a quarter-rate conversion):
            // the low word of 2^52 + k is k, and the gradients are biased by 2^15 to make them non-negative.
            const uint32_t pk = rc[q].x;
            const uint32_t gb = rc[q].y ^ 0x80008000u;
            const double u = __hiloint2double(0x43300000, (int)(pk & 0xfff)) - 4503599627370496.0;
            const double v = __hiloint2double(0x43300000, (int)((pk >> 12) & 0xfff)) - 4503599627370496.0;
            const double2 swr = lut[rq[q]];  // {sqrt(w), sqrt(w) r}
            const double sw = swr.x, swr_ = swr.y;
            // sqrt(w) dIx fx, sqrt(w) dIy fy: every Jacobian component below carries the factor sqrt(w)
            const double a = (__hiloint2double(0x43300000, (int)(gb & 0xffff)) - 4503599627403264.0) * fxJ * sw;
            const double b = (__hiloint2double(0x43300000, (int)(gb >> 16)) - 4503599627403264.0) * fyJ * sw;
            const double zi = __hiloint2double((int)rc[q].w, (int)rc[q].z);
            const double xn = fma(ifx, u, kx), yn = fma(ify, v, ky);
            const double cc = fma(a, xn, b * yn);
            double J[6];
            J[0] = a * zi;
            J[1] = b * zi;
            J[2] = -cc * zi;
            J[3] = -fma(cc, yn, b);
            J[4] = fma(cc, xn, a);
            J[5] = fma(b, xn, -a * yn);
            int s = 0;
#pragma unroll
            for (int i = 0; i < 6; i++) {
#pragma unroll
              for (int j = i; j < 6; j++) {
                acc[s] = fma(J[i], J[j], acc[s]);
                s++;
              }
              acc[21 + i] = fma(J[i], swr_, acc[21 + i]);
            }
          }
        }
      }
      lap(3);
      // prologue of the next phase A (next keyframe, or keyframe 0 of the next iteration): the list does not
      // depend on the pose, so it streams in behind the reduction and the solve
      {
        const ViewState & Vn = Vs[k + 1 < K ? k + 1 : 0];
        int q0, q1;
        slice_of(Vn.nIp, C, rank, q0, q1);
        const int nChN = (q1 - q0 + CH - 1) / CH;
        if (lane == 0)  // the warp's own ring is free as soon as the warp leaves phase C
          for (int c = 0; c < min(NS, nChN); c++) issue_chunk(S, ring, Vn, CHUNK_A, q0, q1, c, item + c, warp, pol);
      }
      // fixed-order block reduction: lanes (xor tree), then warps in order
#pragma unroll
      for (int s = 0; s < NE_TERMS; s++) {
        double v = acc[s];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) S.red[warp * NE_TERMS + s] = v;
      }
      __syncthreads();
      if (tid < NE_TERMS) {
        double v = 0.0;
#pragma unroll
        for (int w = 0; w < GT / 32; w++) v += S.red[w * NE_TERMS + tid];
        // chi2 came out of the histogram (every CTA holds the view's total): rank 0 contributes it
        if (tid == NE_TERMS - 1) v = rank == 0 ? chi2View : 0.0;
        partial[(pb * K + k) * NE_TERMS + tid] = v;
      }
      __syncthreads();
    }
    lap(4);
    // ------------------------------------------------------------------ cluster reduction + solve
    cluster.sync();
    for (int i = tid; i < K * NE_TERMS; i += GT) {
      const int k = i / NE_TERMS, t = i - k * NE_TERMS;
      double v = 0.0;
      for (unsigned rk = 0; rk < C; rk++) v += cluster.map_shared_rank(&partial[(pb * K + k) * NE_TERMS + t], rk)[0];
      sums[i] = v;
      if (viewSums && rank == 0) viewSums[(size_t)(S.J.view0 + k) * NE_TERMS + t] = v;
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
  {
    int p0, p1;
    slice_of(Vs[0].nIp, C, rank, p0, p1);
    const int nCh = (p1 - p0 + CH - 1) / CH;
    for (int c = 0; c < min(NS, nCh); c++) mbar_wait(&S.full[warp][(item + c) % NS], ((item + c) / NS) & 1);
  }
  cluster.sync();  // nobody leaves while its shared memory may still be read remotely
  if (rank == 0) {
    block_copy(jobs + job, &S.J);
    for (int k = 0; k < K; k++) block_copy(views + S.J.view0 + k, &Vs[k]);
  }
}

size_t gn_level_static_smem() { return sizeof(LevelShared); }

// The dense level kernel: any number of keyframes per job, every loss / scaler, Gauss-Newton or Levenberg-Marquardt.
cudaError_t launch_gn_level(
  JobState * jobs, ViewState * views, int job0, int nJobs, int maxViewsPerJob, int clusterSize, SolverParams p, int level,
  int maxEvals, double * trace, size_t traceJobStride, double * viewSums, cudaStream_t s)
{
  if (nJobs <= 0) return cudaSuccess;
  const size_t dyn = level_dyn_smem(maxViewsPerJob);
  cudaError_t e = cudaFuncSetAttribute(gn_level_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn);
  if (e != cudaSuccess) return e;
  if (clusterSize > 8) {
    e = cudaFuncSetAttribute(gn_level_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e != cudaSuccess) return e;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(GT, 1, 1);
  cfg.dynamicSmemBytes = dyn;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  if (clusterSize > 8) {
    // 16-CTA clusters need a GPC with 16 SMs that can each take a CTA of this size: asked once per shared-memory size;
    // a device (or partition) that cannot co-schedule them falls back to the portable maximum
    static int ok16 = -1;
    static size_t dyn16 = 0;
    if (ok16 < 0 || dyn16 != dyn) {
      cfg.gridDim = dim3(16, 1, 1);
      at[0].val.clusterDim.x = 16;
      int n = 0;
      const cudaError_t q = cudaOccupancyMaxActiveClusters(&n, gn_level_kernel, &cfg);
      if (q != cudaSuccess) cudaGetLastError();
      ok16 = (q == cudaSuccess && n >= 1) ? 1 : 0;
      dyn16 = dyn;
    }
    if (!ok16) clusterSize = 8;
  }
  cfg.gridDim = dim3((unsigned)nJobs * clusterSize, 1, 1);
  at[0].val.clusterDim.x = clusterSize;
  return cudaLaunchKernelEx(&cfg, gn_level_kernel, jobs, views, p, level, maxEvals, trace, traceJobStride, viewSums, job0);
}
}  // namespace vb
