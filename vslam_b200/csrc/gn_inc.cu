// Incremental Gauss-Newton level kernel (single keyframe, zero-offset scalers).
//
// The inverse-compositional problem of the reference keeps the Jacobian of a point fixed for the whole level
// (InverseCompositional.cpp:63-93) and samples the target image at the NEAREST pixel
// (InverseCompositional.cpp:119-120), so between two Gauss-Newton iterations only the integer residual of the
// few points whose rounded target pixel moved can change (measured: 0.5 - 3 % per iteration at 640x480), while
// the robust scale - and with it every weight - changes every time.  Both facts are used exactly:
//
//   * MOMENTS.  Per residual magnitude c = |r| in 0..255 the CTA keeps 21 + 6 moments
//         M[c][i,j] = sum_{k: |r_k| = c} J_ki J_kj          M[c][21+i] = sum_{k: |r_k| = c} sign(r_k) J_ki
//     so that for any weight function of |r| (zero-offset scalers: identity and t-distribution)
//         A = sum_c w(c) M[c][i,j]      b = sum_c w(c) c M[c][21+i]            (NormalEquations.cpp:53-61)
//     costs 256 x 27 multiply-adds per iteration instead of 27 per point.  The moments are 64-bit FIXED-POINT
//     integers (two 32-bit words, the carry handed over explicitly) updated with integer atomics: integer addition
//     is exact and associative, so the sums do not depend on the order in which warps reach them (bit-identical
//     run to run) and removing a point's old contribution restores the cell exactly.  Each Jacobian component is
//     scaled by a power of two chosen from sum_k J_ki^2 (Cauchy-Schwarz bounds every partial sum by 2^61), which
//     leaves a quantum of ~2^-61 of the diagonal terms: finer than an fp64 accumulation.
//   * MOTION BOUNDS.  When a point is evaluated, the distance of its projection from the nearest decision boundary
//     (rounding to another pixel, the visibility ring of InverseCompositional.cpp:115-116) is known; it cannot
//     change its residual before the image motion accumulated since then exceeds that margin.  Per tile of 128
//     list entries a rigorous bound of the image motion of one Gauss-Newton step is computed from the ranges of
//     (u, v, 1/depth) in the tile (interval arithmetic on the exact displacement formula, see tile_motion_bound)
//     and accumulated; a point is looked at again only when the tile's accumulated bound reaches the point's
//     limit.  Everything else keeps its residual code, its histogram entry and its moments.
//
// One CTA per alignment job runs the whole GaussNewton::solve loop of the level (GaussNewton.cpp:33-102):
//   first evaluation   warp + sample every point (filtered fp32 evaluation, exact fp64 fallback: warp_filter.cuh),
//                      residual codes, limits, histogram, tile ranges, then the moments of all points
//   later evaluations  per tile: motion bound, compare with the points' limits; flagged points are queued per
//                      warp, re-evaluated 32 at a time, and those whose residual changed move their moments
//   every evaluation   scale from the histogram (Scaler.cpp:49-67), weight table (Loss.cpp:51-92), A / b / chi2
//                      from the moments, 6x6 solve and stop rules (solve.cuh)
#include "robust.cuh"
#include "solve.cuh"
#include "warp_filter.cuh"
#include "bounds.cuh"

namespace vb
{
namespace inc
{
constexpr int NT = 384;            // threads per CTA
constexpr int NW = NT / 32;        // warps per CTA
constexpr int CTAS_PER_SM = 2;
constexpr int NC = 256;            // residual magnitudes 0..255
constexpr int MT = 27;             // moments per magnitude: 21 (upper A) + 6 (b)
constexpr int JS = 7;              // doubles per stashed point: 6 scaled Jacobian components + the b-term scale
constexpr int EB = 4;              // re-evaluation super-batch: EB points per lane, all loads of a round trip in flight together
constexpr int QE = 256;            // evaluation queue per warp (circular): < 32 EB left over + one tile of 128 new entries
constexpr int QC = 160;            // change queue per warp: <= 31 left over + 32 EB new entries
constexpr double BSCALE = 1048576.0;  // 2^20: extra scale of the b moments (|sum_k J'_ki| <= 2^10 * 2^30.5)
constexpr int NCHUNK = 8;          // the A-from-moments pass splits the 256 magnitudes into 8 chunks of 32
constexpr unsigned FLAGGED = 0x80000000u;  // eval-queue entry: the motion bound asked for this point (check mode)

struct Shared
{
  JobState J;
  ViewState V;
  SolveScratch X;
  StepConsts stp;
  int hist[HIST_BINS];
  double wA[NC], wB[NC];           // w(c) and w(c) c
  double part[NCHUNK][MT + 1];
  double sums[NE_TERMS];
  double red[NW];
  float redf[NW][6];
  double scaleJ[6];                // 2^e_i
  double unscale[MT];              // 2^-(e_i + e_j), 2^-(e_i + 20)
  double sigma;
  int vis;
  int redi[NW];
  unsigned long long nFlagged, nChanged, nMissed;
};
// dynamic shared memory: Mlo u32[NC][MT] | Mhi u32[NC][MT] | stash double[NW][32][JS] | evalQ u32[NW][QE] |
// changeQ u32[NW][QC]
constexpr size_t OFF_MHI = (size_t)NC * MT * 4;
constexpr size_t OFF_STASH = OFF_MHI + (size_t)NC * MT * 4;
constexpr size_t OFF_EVALQ = OFF_STASH + (size_t)NW * 32 * JS * 8;
constexpr size_t OFF_CHGQ = OFF_EVALQ + (size_t)NW * QE * 4;
constexpr size_t DYN_BYTES = OFF_CHGQ + (size_t)NW * QC * 4;

template <typename T>
__device__ __forceinline__ void block_copy(T * dst, const T * src)
{
  static_assert(sizeof(T) % 4 == 0, "word copy");
  const uint32_t * s = reinterpret_cast<const uint32_t *>(src);
  uint32_t * d = reinterpret_cast<uint32_t *>(dst);
  for (int i = threadIdx.x; i < (int)(sizeof(T) / 4); i += NT) d[i] = s[i];
}

__device__ __forceinline__ double block_sum(double v, double * sm)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0;
#pragma unroll
  for (int w = 0; w < NW; w++) s += sm[w];
  return s;
}

// which Jacobian components term t multiplies: 21 upper-triangle pairs (row-major, as NE_TERMS orders them), then
// the 6 b terms (component x stash slot 6, which holds the b scale)
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

__device__ __forceinline__ void moment_add(uint32_t * Mlo, uint32_t * Mhi, int code, int lane, long long v, bool subtract)
{
  if (code == R_CODE_INVISIBLE) return;
  const int cs = code - 255;
  long long sv = (lane >= 21 && cs < 0) ? -v : v;
  sv = subtract ? -sv : sv;
  const int cell = abs(cs) * MT + lane;
  const uint32_t lo = (uint32_t)(unsigned long long)sv, hi = (uint32_t)((unsigned long long)sv >> 32);
  const uint32_t old = atomicAdd(&Mlo[cell], lo);
  atomicAdd(&Mhi[cell], hi + ((uint32_t)(old + lo) < old ? 1u : 0u));  // 64-bit add, carry handed over
}

// Moves the moments of `cnt` (<= 32) points from their old to their new residual cell.  Lane l < cnt describes one
// point: list index, old and new residual code; lanes >= cnt pass R_CODE_INVISIBLE twice.  First the lanes compute
// the fp64 Jacobian rows of their points (WarpSE3::J contracted with the template gradient, Warp.cpp:166-201,
// InverseCompositional.cpp:79-80) into the warp's stash; then lane t < 27 owns term t and walks over the points,
// four at a time (independent dependency chains).
__device__ __forceinline__ void move_moments(
  const ViewState & V, const double * scaleJ, uint32_t * Mlo, uint32_t * Mhi, double * stash, int cnt, int idx, int oldc,
  int newc, int ti, int tj)
{
  const int lane = threadIdx.x & 31;
  __syncwarp();
  {
    double * s = stash + lane * JS;
    if (lane < cnt) {
      const unsigned char * tile = V.tiles + (size_t)(idx / LIST_TILE) * LIST_TILE_BYTES;
      const int j = idx % LIST_TILE;
      const uint32_t pk = reinterpret_cast<const uint32_t *>(tile + TILE_OFF_UVT)[j];
      const uint32_t gg = reinterpret_cast<const uint32_t *>(tile + TILE_OFF_G)[j];
      const double zi = reinterpret_cast<const double *>(tile + TILE_OFF_ZI)[j];
      const double u = (double)(int)(pk & 0xfff), v = (double)(int)((pk >> 12) & 0xfff);
      const double a = (double)(int)(int16_t)(gg & 0xffff) * V.fxJd;  // dIx * fx
      const double b = (double)(int)(int16_t)(gg >> 16) * V.fyJd;     // dIy * fy
      const double xn = fma(V.ifx, u, V.kx), yn = fma(V.ify, v, V.ky);
      const double cc = fma(a, xn, b * yn);
      s[0] = (a * zi) * scaleJ[0];
      s[1] = (b * zi) * scaleJ[1];
      s[2] = (-cc * zi) * scaleJ[2];
      s[3] = (-fma(cc, yn, b)) * scaleJ[3];
      s[4] = fma(cc, xn, a) * scaleJ[4];
      s[5] = fma(b, xn, -a * yn) * scaleJ[5];
      s[6] = BSCALE;
    } else {
#pragma unroll
      for (int i = 0; i < JS; i++) s[i] = 0.0;
    }
  }
  __syncwarp();
  for (int pnt = 0; pnt < cnt; pnt += 4) {
    int o[4], n[4];
    long long v[4];
#pragma unroll
    for (int k = 0; k < 4; k++) {
      o[k] = __shfl_sync(0xffffffffu, oldc, pnt + k);
      n[k] = __shfl_sync(0xffffffffu, newc, pnt + k);
      v[k] = __double2ll_rn(stash[(pnt + k) * JS + ti] * stash[(pnt + k) * JS + tj]);
    }
    if (lane < MT) {
#pragma unroll
      for (int k = 0; k < 4; k++) {
        moment_add(Mlo, Mhi, o[k], lane, v[k], true);
        moment_add(Mlo, Mhi, n[k], lane, v[k], false);
      }
    }
  }
  __syncwarp();
}

__global__ void __launch_bounds__(NT, CTAS_PER_SM) gn_inc_kernel(
  JobState * __restrict__ jobs, ViewState * __restrict__ views, SolverParams p, int level, int maxEvals,
  double * __restrict__ trace, size_t traceJobStride, int job0, int checkMode)
{
  const int job = job0 + blockIdx.x;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  __shared__ Shared S;
  extern __shared__ __align__(16) unsigned char dyn[];
  uint32_t * Mlo = reinterpret_cast<uint32_t *>(dyn);
  uint32_t * Mhi = reinterpret_cast<uint32_t *>(dyn + OFF_MHI);
  double * stash = reinterpret_cast<double *>(dyn + OFF_STASH) + warp * 32 * JS;
  uint32_t * evalQ = reinterpret_cast<uint32_t *>(dyn + OFF_EVALQ) + warp * QE;
  uint32_t * chgQ = reinterpret_cast<uint32_t *>(dyn + OFF_CHGQ) + warp * QC;

  block_copy(&S.J, jobs + job);
  __syncthreads();
  if (!S.J.active) return;
  block_copy(&S.V, views + S.J.view0);
  for (int i = tid; i < 2 * NC * MT; i += NT) Mlo[i] = 0u;  // Mlo and Mhi are adjacent
  for (int i = tid; i < HIST_BINS; i += NT) S.hist[i] = 0;
  if (tid == 0) S.nFlagged = S.nChanged = S.nMissed = 0ull;
  __syncthreads();
  ViewState & V = S.V;
  const int n = V.nIp;
  const int nTiles = (n + LIST_TILE - 1) / LIST_TILE;
  const uint8_t * __restrict__ I = V.I;
  uint2 * lim = reinterpret_cast<uint2 *>(V.lim);  // per list entry: 4 x 16-bit limits {+u | -u, +v | -v}
  // per tile 8 words: ulo | uhi << 16, vlo | vhi << 16, zlo, zhi (1/depth range), accumulated bounds +u, -u, +v, -v
  float * tileState = V.tileState;
  int16_t * rcode = V.r;
  int ti, tj;
  term_pair(lane, ti, tj);
  int evals = 0;
  unsigned myFlagged = 0, myChanged = 0, myMissed = 0;  // lane 0 of every warp counts

  long long tc = clock64();
  auto lap = [&](int slot) {
    if (tid == 0) {
      const long long now = clock64();
      S.J.clk[level][slot] += now - tc;
      tc = now;
    }
  };

  for (;;) {
    const bool first = evals == 0;
    WarpConsts c;
    c.r00 = V.rI[0]; c.r01 = V.rI[1]; c.r02 = V.rI[2]; c.r10 = V.rI[3]; c.r11 = V.rI[4]; c.r12 = V.rI[5];
    c.r20 = V.rI[6]; c.r21 = V.rI[7]; c.r22 = V.rI[8];
    c.t0 = V.t32[0]; c.t1 = V.t32[1]; c.t2 = V.t32[2];
    c.fx = V.fx32; c.fy = V.fy32; c.ifx = V.ifxH; c.ify = V.ifyH; c.kx = V.kxH; c.ky = V.kyH;
    c.epsA = V.epsA; c.epsB = V.epsB; c.Wt = V.Wt; c.Ht = V.Ht;
    if (V.Wt < 4 || V.Ht < 4) c.epsB = __int_as_float(0x7f800000);  // no strict interior: always the exact path
    if (first) {
      // ---------------------------------------------------------------- first evaluation: every point
      float sq[6] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};  // sum_k J_ki^2 (fp32, for the fixed-point scales)
      const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int tile = warp; tile < nTiles; tile += NW) {
        const unsigned char * tp = V.tiles + (size_t)tile * LIST_TILE_BYTES;
        const int g0 = tile * LIST_TILE + 4 * lane;
        const uint4 pk4 = reinterpret_cast<const uint4 *>(tp + TILE_OFF_UVT)[lane];
        const float4 dd4 = reinterpret_cast<const float4 *>(tp + TILE_OFF_D)[lane];
        const uint4 gg4 = reinterpret_cast<const uint4 *>(tp + TILE_OFF_G)[lane];
        const uint32_t pk[4] = {pk4.x, pk4.y, pk4.z, pk4.w};
        const float dd[4] = {dd4.x, dd4.y, dd4.z, dd4.w};
        const uint32_t gg[4] = {gg4.x, gg4.y, gg4.z, gg4.w};
        int code[4];
        uint2 lq[4];
        unsigned ulo = 0xffffffffu, uhi = 0u, vlo = 0xffffffffu, vhi = 0u, zlo = 0xffffffffu, zhi = 0u;
#pragma unroll
        for (int q = 0; q < 4; q++) {
          const bool in = g0 + q < n;
          float m[4];
          int pix = warp_point(c, V, pk[q], dd[q], m);  // padding lanes compute on stale data: harmless
          pix = in ? pix : -1;
          const int val = pix >= 0 ? (int)__ldg(I + pix) : 0;
          code[q] = pix >= 0 ? val - (int)(pk[q] >> 24) + 255 : R_CODE_INVISIBLE;
          lq[q] = point_limits(zero4, m);
          if (code[q] != R_CODE_INVISIBLE) atomicAdd(&S.hist[code[q]], 1);
          if (in) {
            const unsigned ui = pk[q] & 0xfff, vi = (pk[q] >> 12) & 0xfff;
            const float u = (float)ui, v = (float)vi;
            const float a = (float)(int)(int16_t)(gg[q] & 0xffff) * V.fxJH, b = (float)(int)(int16_t)(gg[q] >> 16) * V.fyJH;
            const float zi = __fdividef(1.f, dd[q]);
            const float xn = fmaf(c.ifx, u, c.kx), yn = fmaf(c.ify, v, c.ky);
            const float cc = fmaf(a, xn, b * yn);
            const float j0 = a * zi, j1 = b * zi, j2 = cc * zi, j3 = fmaf(cc, yn, b), j4 = fmaf(cc, xn, a), j5 = fmaf(b, xn, -a * yn);
            sq[0] = fmaf(j0, j0, sq[0]);
            sq[1] = fmaf(j1, j1, sq[1]);
            sq[2] = fmaf(j2, j2, sq[2]);
            sq[3] = fmaf(j3, j3, sq[3]);
            sq[4] = fmaf(j4, j4, sq[4]);
            sq[5] = fmaf(j5, j5, sq[5]);
            ulo = min(ulo, ui); uhi = max(uhi, ui); vlo = min(vlo, vi); vhi = max(vhi, vi);
            const unsigned zb = __float_as_uint(zi);  // positive floats order like their bit patterns
            zlo = min(zlo, zb); zhi = max(zhi, zb);
          }
        }
        if (g0 < n) {
          reinterpret_cast<uint2 *>(rcode + tile * LIST_TILE)[lane] =
            make_uint2((uint32_t)code[0] | ((uint32_t)code[1] << 16), (uint32_t)code[2] | ((uint32_t)code[3] << 16));
          uint4 * lp = reinterpret_cast<uint4 *>(lim + tile * LIST_TILE + 4 * lane);
          lp[0] = make_uint4(lq[0].x, lq[0].y, lq[1].x, lq[1].y);
          lp[1] = make_uint4(lq[2].x, lq[2].y, lq[3].x, lq[3].y);
        }
        ulo = __reduce_min_sync(0xffffffffu, ulo); uhi = __reduce_max_sync(0xffffffffu, uhi);
        vlo = __reduce_min_sync(0xffffffffu, vlo); vhi = __reduce_max_sync(0xffffffffu, vhi);
        zlo = __reduce_min_sync(0xffffffffu, zlo); zhi = __reduce_max_sync(0xffffffffu, zhi);
        if (lane == 0) {
          float4 * ts = reinterpret_cast<float4 *>(tileState + tile * 8);
          ts[0] = make_float4(__uint_as_float(ulo | (uhi << 16)), __uint_as_float(vlo | (vhi << 16)),
                              __uint_as_float(zlo) * 0.9999f, __uint_as_float(zhi) * 1.0001f);  // 1/d was approximate
          ts[1] = zero4;
        }
      }
      lap(0);
      // fixed-point scales: 2^e_i with sum_k (J_ki 2^e_i)^2 <= 2^61 (the fp32 sums carry ~1e-5 relative error and
      // the bound keeps two bits of slack below 2^63)
#pragma unroll
      for (int i = 0; i < 6; i++) {
        float v = sq[i];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if (lane == 0) S.redf[warp][i] = v;
      }
      __syncthreads();
      if (tid < 6) {
        double s = 0.0;
        for (int w = 0; w < NW; w++) s += (double)S.redf[w][tid];
        s *= 1.001;
        int e = 0;
        if (s > 0.0 && isfinite(s)) {
          int ex;
          frexp(s, &ex);          // s = f 2^ex, 0.5 <= f < 1  ->  s < 2^ex
          e = (61 - ex) >> 1;     // floor((61 - ex) / 2): s 2^(2e) < 2^61
        }
        e = max(-480, min(480, e));
        S.scaleJ[tid] = ldexp(1.0, e);
      }
      __syncthreads();
      if (tid < MT) {
        int i, j;
        term_pair(tid, i, j);
        S.unscale[tid] = 1.0 / (S.scaleJ[i] * (j < 6 ? S.scaleJ[j] : BSCALE));
      }
      // moments of every point (reads the codes written above: plain loads, ordered by the barriers)
      for (int tile = warp; tile < nTiles; tile += NW) {
#pragma unroll 1
        for (int b = 0; b < LIST_TILE / 32; b++) {
          const int g = tile * LIST_TILE + b * 32;
          if (g >= n) break;
          const int cnt = min(32, n - g);
          const int code = lane < cnt ? (int)(uint16_t)rcode[g + lane] : R_CODE_INVISIBLE;
          move_moments(V, S.scaleJ, Mlo, Mhi, stash, cnt, g + lane, R_CODE_INVISIBLE, code, ti, tj);
        }
      }
      lap(1);
    } else {
      // ---------------------------------------------------------------- later evaluations
      // directional motion bounds of the last step per tile (one lane per tile), accumulated
      for (int t0 = warp; t0 < nTiles; t0 += NW * 32) {
        const int tile = t0 + NW * lane;
        if (tile < nTiles) {
          float4 * ts = reinterpret_cast<float4 *>(tileState + tile * 8);
          const float4 a = ts[0];
          float4 acc = ts[1];
          const unsigned uw = __float_as_uint(a.x), vw = __float_as_uint(a.y);
          const float4 b = tile_motion_bound(S.stp, (float)(uw & 0xffff), (float)(uw >> 16), (float)(vw & 0xffff),
                                             (float)(vw >> 16), a.z, a.w);
          acc.x += b.x; acc.y += b.y; acc.z += b.z; acc.w += b.w;
          ts[1] = acc;
        }
      }
      __syncwarp();
      int qh = 0, qe = 0, qc = 0;  // evaluation queue: head and fill (circular); change queue fill.  Warp-uniform
      // moves the moments of the last `cnt` (<= 32) entries of the change queue: list index | new code << 20
      auto move_batch = [&](int cnt) {
        const uint32_t e = lane < cnt ? chgQ[qc - cnt + lane] : 0u;
        const int idx = (int)(e & 0xfffffu);
        int oldc = R_CODE_INVISIBLE, newc = R_CODE_INVISIBLE;
        if (lane < cnt) {
          newc = (int)(e >> 20);
          oldc = (int)(uint16_t)rcode[idx];
          rcode[idx] = (int16_t)newc;
          if (oldc != R_CODE_INVISIBLE) atomicSub(&S.hist[oldc], 1);
          if (newc != R_CODE_INVISIBLE) atomicAdd(&S.hist[newc], 1);
        }
        move_moments(V, S.scaleJ, Mlo, Mhi, stash, cnt, idx, oldc, newc, ti, tj);
        qc -= cnt;
      };
      // re-evaluates the first `cnt` (<= 32 EB) queued points; lane l owns entries l, l + 32, ...  All loads of a
      // round trip (list entry, residual code and tile bounds; then the target pixels) are in flight together.
      auto eval_batch = [&](int cnt) {
        uint32_t ent[EB], pk[EB];
        float d[EB];
        int oldc[EB], pix[EB], val[EB];
        float4 acc[EB];
        uint2 lq[EB];
#pragma unroll
        for (int j = 0; j < EB; j++) {
          const bool on = j * 32 + lane < cnt;
          ent[j] = on ? evalQ[(qh + j * 32 + lane) & (QE - 1)] : 0xffffffffu;
        }
#pragma unroll
        for (int j = 0; j < EB; j++) {
          pk[j] = 0u; d[j] = 1.f; oldc[j] = R_CODE_INVISIBLE; acc[j] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (ent[j] != 0xffffffffu) {
            const int idx = (int)(ent[j] & ~FLAGGED);
            const int tile = idx / LIST_TILE, k = idx % LIST_TILE;
            const unsigned char * tp = V.tiles + (size_t)tile * LIST_TILE_BYTES;
            pk[j] = reinterpret_cast<const uint32_t *>(tp + TILE_OFF_UVT)[k];
            d[j] = reinterpret_cast<const float *>(tp + TILE_OFF_D)[k];
            oldc[j] = (int)(uint16_t)rcode[idx];
            acc[j] = reinterpret_cast<const float4 *>(tileState + tile * 8)[1];
          }
        }
#pragma unroll
        for (int j = 0; j < EB; j++) {
          float m[4];
          pix[j] = warp_point(c, V, pk[j], d[j], m);
          lq[j] = point_limits(acc[j], m);
          if (ent[j] == 0xffffffffu) pix[j] = -1;
        }
#pragma unroll
        for (int j = 0; j < EB; j++) val[j] = pix[j] >= 0 ? (int)__ldg(I + pix[j]) : 0;
#pragma unroll
        for (int j = 0; j < EB; j++) {
          const bool on = ent[j] != 0xffffffffu;
          const int idx = (int)(ent[j] & ~FLAGGED);
          const int code = pix[j] >= 0 ? val[j] - (int)(pk[j] >> 24) + 255 : R_CODE_INVISIBLE;
          const bool ch = on && code != oldc[j];
          if (on && (ent[j] & FLAGGED)) lim[idx] = lq[j];
          const unsigned mch = __ballot_sync(0xffffffffu, ch);
          if (lane == 0) myChanged += __popc(mch);
          if (checkMode) {
            const unsigned miss = __ballot_sync(0xffffffffu, ch && !(ent[j] & FLAGGED));
            if (lane == 0) myMissed += __popc(miss);
          }
          if (ch) chgQ[qc + __popc(mch & ((1u << lane) - 1u))] = (uint32_t)idx | ((uint32_t)code << 20);
          qc += __popc(mch);
        }
        qh = (qh + cnt) & (QE - 1);
        qe -= cnt;
        __syncwarp();
        while (qc >= 32) move_batch(32);
      };
      // scan: two tiles per step, their limits and bounds loaded together
      for (int tile0 = warp; tile0 < nTiles; tile0 += 2 * NW) {
        uint4 la[2][2];
        float4 accv[2];
#pragma unroll
        for (int t = 0; t < 2; t++) {
          const int tile = tile0 + t * NW;
          if (tile < nTiles) {
            const uint4 * lp = reinterpret_cast<const uint4 *>(lim + tile * LIST_TILE + 4 * lane);
            la[t][0] = lp[0];
            la[t][1] = lp[1];
            accv[t] = reinterpret_cast<const float4 *>(tileState + tile * 8)[1];
          }
        }
#pragma unroll
        for (int t = 0; t < 2; t++) {
          const int tile = tile0 + t * NW;
          if (tile >= nTiles) break;
          const int g0 = tile * LIST_TILE + 4 * lane;
          const uint2 aq = bounds_q(accv[t]);
          const uint2 lm[4] = {make_uint2(la[t][0].x, la[t][0].y), make_uint2(la[t][0].z, la[t][0].w),
                               make_uint2(la[t][1].x, la[t][1].y), make_uint2(la[t][1].z, la[t][1].w)};
#pragma unroll
          for (int q = 0; q < 4; q++) {
            const bool in = g0 + q < n;
            const bool fl = in && limit_reached(aq, lm[q]);
            const bool ev = fl || (checkMode && in);  // check mode looks at every point; only flagged ones may change
            const unsigned m = __ballot_sync(0xffffffffu, ev);
            if (checkMode || m) {
              const unsigned mf = __ballot_sync(0xffffffffu, fl);
              if (lane == 0) myFlagged += __popc(mf);
            }
            if (ev) evalQ[(qh + qe + __popc(m & ((1u << lane) - 1u))) & (QE - 1)] = (uint32_t)(g0 + q) | (fl ? FLAGGED : 0u);
            qe += __popc(m);
          }
          __syncwarp();
          if (qe >= 32 * EB) eval_batch(32 * EB);
        }
      }
      while (qe > 0) eval_batch(min(qe, 32 * EB));
      if (qc > 0) move_batch(qc);
      lap(0);
    }
    __syncthreads();
    // ------------------------------------------------------------------ scale, weights, chi2
    {
      int cvis = 0;
      for (int i = tid; i < HIST_BINS - 1; i += NT) cvis += S.hist[i];
      cvis = __reduce_add_sync(0xffffffffu, cvis);
      if (lane == 0) S.redi[warp] = cvis;
      __syncthreads();
      if (tid < 32) {
        int vis = 0;
        for (int w = 0; w < NW; w++) vis += S.redi[w];
        double sigma = 1.0;
        if (p.scalerKind == VSLAMB200_SCALER_TDIST) sigma = tdist_sigma_warp(S.hist, n - vis, (long)n, p.scalerV);
        if (tid == 0) {
          S.sigma = sigma;
          S.vis = vis;
          V.scale = sigma;
          V.offset = 0.0;
        }
      }
      __syncthreads();
      double c2 = 0.0;
      if (tid < NC) {
        const double r = (double)tid;
        const double w = p.lossKind == VSLAMB200_LOSS_NONE ? 1.0 : loss_weight(p, (r - 0.0) / S.sigma);
        const double wr = w * r;
        S.wA[tid] = w;
        S.wB[tid] = wr;
        const int cnt = tid == 0 ? 0 : S.hist[255 + tid] + S.hist[255 - tid];
        c2 = (double)cnt * (wr * r);
      }
      const double chi2 = block_sum(c2, S.red);  // contains the barriers that publish the table
      if (tid == 0) S.sums[NE_TERMS - 1] = chi2;
    }
    lap(2);
    // ------------------------------------------------------------------ A, b from the moments (fixed order)
    if (tid < MT * NCHUNK) {
      const int t = tid % MT, ch = tid / MT;
      const double * w = t < 21 ? S.wA : S.wB;
      double acc = 0.0;
#pragma unroll 4
      for (int cI = ch * (NC / NCHUNK); cI < (ch + 1) * (NC / NCHUNK); cI++) {
        const long long mv = (long long)(((unsigned long long)Mhi[cI * MT + t] << 32) | (unsigned long long)Mlo[cI * MT + t]);
        acc = fma(w[cI], (double)mv, acc);
      }
      S.part[ch][t] = acc;
    }
    __syncthreads();
    if (tid < MT) {
      double v = 0.0;
#pragma unroll
      for (int ch = 0; ch < NCHUNK; ch++) v += S.part[ch][tid];
      S.sums[tid] = v * S.unscale[tid];
    }
    // the pose before the step, for the next evaluation's motion bounds
    if (tid >= 32 && tid < 41) S.stp.Ro[tid - 32] = (float)V.R[tid - 32];
    if (tid >= 41 && tid < 44) S.stp.to[tid - 41] = (float)V.t[tid - 41];
    __syncthreads();
    lap(3);
    evals++;
    if (tid == 0) {
      S.J.evals[level]++;
      S.J.pointIters += (unsigned long long)S.J.totalPoints;
    }
    solver_step_block(S.J, &S.V, S.sums, S.X, p, level, trace ? trace + (size_t)job * traceJobStride : nullptr, evals >= maxEvals);
    if (!S.J.active) break;
    if (tid == 0) {
      double Rinc[9];
      quat_to_R(S.X.inc.q, Rinc);
      for (int i = 0; i < 9; i++) S.stp.E[i] = (float)(Rinc[i] - ((i % 4 == 0) ? 1.0 : 0.0));
      for (int i = 0; i < 3; i++) S.stp.dt[i] = (float)S.X.inc.t[i];
      S.stp.ifx = (float)V.ifx; S.stp.kx = (float)V.kx; S.stp.ify = (float)V.ify; S.stp.ky = (float)V.ky;
      S.stp.fx = (float)V.fx; S.stp.fy = (float)V.fy;
    }
    __syncthreads();
    lap(6);
  }
  if (lane == 0) {
    atomicAdd(&S.nFlagged, (unsigned long long)myFlagged);
    atomicAdd(&S.nChanged, (unsigned long long)myChanged);
    atomicAdd(&S.nMissed, (unsigned long long)myMissed);
  }
  __syncthreads();
  if (tid == 0) {
    S.J.clk[level][7] += (long long)S.nFlagged;
    S.J.clk[level][8] += (long long)S.nChanged;
    S.J.clk[level][9] += (long long)S.nMissed;
  }
  __syncthreads();
  block_copy(jobs + job, &S.J);
  block_copy(views + S.J.view0, &S.V);
}
}  // namespace inc

cudaError_t launch_gn_inc(
  JobState * jobs, ViewState * views, int job0, int nJobs, SolverParams p, int level, int maxEvals, double * trace,
  size_t traceJobStride, int checkMode, cudaStream_t s)
{
  if (nJobs <= 0) return cudaSuccess;
  cudaError_t e = cudaFuncSetAttribute(inc::gn_inc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)inc::DYN_BYTES);
  if (e != cudaSuccess) return e;
  inc::gn_inc_kernel<<<nJobs, inc::NT, inc::DYN_BYTES, s>>>(jobs, views, p, level, maxEvals, trace, traceJobStride, job0, checkMode);
  return cudaGetLastError();
}
}  // namespace vb
