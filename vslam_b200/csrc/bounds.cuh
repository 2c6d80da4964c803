// Motion bounds for the level kernels that re-evaluate only the interest points whose sampled pixel can have changed.
//
// The inverse-compositional problem of the reference samples the target image at the NEAREST pixel
// (InverseCompositional.cpp:119-120), so the integer residual of a point changes only when its projection crosses a
// rounding boundary or the visibility ring (InverseCompositional.cpp:115-116).  When a point is evaluated, the
// distances of its projection from those boundaries are known per direction; per tile of 128 list entries a rigorous
// bound of the image motion of every Gauss-Newton step is accumulated per direction; a point is looked at again only
// when an accumulated bound reaches the point's limit for that direction.
#pragma once
#include "warp_filter.cuh"

namespace vb
{
constexpr float LIMQ = 2048.f;  // limits and accumulated motion bounds are compared as 16-bit fixed point, 1/2048 px

// what one Gauss-Newton step did, for the motion bounds: the increment exp(-dx) applied on the template side
// (Warp.cpp:153-157: se3 <- se3 * exp(-dx)) and the pose before it
struct StepConsts
{
  float E[9], dt[3];   // rotation of exp(-dx) minus identity, its translation
  float Ro[9], to[3];  // se3 before the step
  float ifx, kx, ify, ky, fx, fy;
};

// Sampled target pixel (or -1) of one interest point and the DIRECTIONAL margins by which its projection may still
// move before the sampled pixel or the visibility can change: m[0] towards +u, m[1] towards -u, m[2] towards +v,
// m[3] towards -v (pixels).  Same decisions as the dense kernel's warp_pick_ring / warp_exact pair (warp_filter.cuh).
// With u' = iu + fu, fu in [-0.5, 0.5], the rounded pixel changes at fu = +-0.5; on the visibility ring (iu == 1 or
// iu == W-1: visible iff u' > 1 resp. u' < W-1, InverseCompositional.cpp:115-116) the nearer boundary is fu = 0.
__device__ __forceinline__ bool directional_margins(const WarpConsts & c, const WarpEval & e, float (&m)[4])
{
  const int iu = e.iu, iv = e.iv;
  float pu = 0.5f - e.fu, nu = 0.5f + e.fu, pv = 0.5f - e.fv, nv = 0.5f + e.fv;
  const bool ringu = iu == 1 || iu == c.Wt - 1, ringv = iv == 1 || iv == c.Ht - 1;
  if (ringu) {
    if (e.fu > 0.f) nu = e.fu; else pu = -e.fu;
  }
  if (ringv) {
    if (e.fv > 0.f) nv = e.fv; else pv = -e.fv;
  }
  const float dist = fminf(fminf(pu, nu), fminf(pv, nv));
  const bool ok = e.okb && dist > e.eps;  // the same criterion as warp_pick_interior / warp_pick_ring
  const float slack = e.eps + 2e-5f;  // the fp32 evaluation's own error, the rounding of the limit arithmetic
  m[0] = ok ? pu - slack : 0.f;
  m[1] = ok ? nu - slack : 0.f;
  m[2] = ok ? pv - slack : 0.f;
  m[3] = ok ? nv - slack : 0.f;
  return ok;
}
__device__ __forceinline__ int warp_point(const WarpConsts & c, const ViewState & V, uint32_t pk, float d, float (&m)[4])
{
  const WarpEval e = warp_eval(c, pk, d);
  const int iu = e.iu, iv = e.iv;
  const bool ok = directional_margins(c, e, m);
  bool vis = (unsigned)(iu - 1) <= (unsigned)(c.Wt - 2) && (unsigned)(iv - 1) <= (unsigned)(c.Ht - 2);
  vis = vis && !(iu == 1 && e.fu < 0.f) && !(iu == c.Wt - 1 && e.fu > 0.f);
  vis = vis && !(iv == 1 && e.fv < 0.f) && !(iv == c.Ht - 1 && e.fv > 0.f);
  if (!ok) return warp_exact(V, pk & 0xfff, (pk >> 12) & 0xfff, d);
  return vis ? iv * c.Wt + iu : -1;
}

// 16-bit fixed-point forms (1/2048 px): a limit is rounded DOWN, an accumulated bound UP, so the comparison
// accumulated >= limit never flags later than the real numbers would.  NaN limit -> 0 (always flagged), NaN or
// infinite bound -> 65535 (flags everything; limits saturate at 65535 too).
__device__ __forceinline__ uint32_t limit_q(float x)
{
  const float y = x * LIMQ;
  return y > 0.f ? (uint32_t)__float2int_rd(fminf(y, 65535.f)) : 0u;
}
__device__ __forceinline__ uint32_t bound_q(float x)
{
  const float y = x * LIMQ;
  return y < 65535.f ? (uint32_t)__float2int_ru(fmaxf(y, 0.f)) : 65535u;
}
// the point's limits {+u | -u << 16, +v | -v << 16} given the tile's accumulated bounds at evaluation time
__device__ __forceinline__ uint2 point_limits(const float4 acc, const float (&m)[4])
{
  return make_uint2(limit_q(acc.x + m[0]) | (limit_q(acc.y + m[1]) << 16), limit_q(acc.z + m[2]) | (limit_q(acc.w + m[3]) << 16));
}
__device__ __forceinline__ uint2 bounds_q(const float4 acc)
{
  return make_uint2(bound_q(acc.x) | (bound_q(acc.y) << 16), bound_q(acc.z) | (bound_q(acc.w) << 16));
}
__device__ __forceinline__ bool limit_reached(const uint2 accq, const uint2 lim)
{
  return (__vcmpgeu2(accq.x, lim.x) | __vcmpgeu2(accq.y, lim.y)) != 0u;
}

// ---- all-integer form used by the bounded cluster kernel (gn_level.cu): accumulated bounds and limits are 15-bit
// fixed point (1/2048 px, saturating at 32767 = 16 px), two per 32-bit word {+ | - << 16}.  Bound increments are rounded
// UP before they are accumulated, a point's limit is accumulated bound + margin rounded DOWN, so "accumulated >= limit"
// never fires later than the real numbers would; a saturated accumulator flags every point of its tile.
__device__ __forceinline__ uint32_t margin_q15(float m)  // NaN -> 0
{
  return (uint32_t)min(max(__float2int_rd(m * LIMQ), 0), 32767);
}
__device__ __forceinline__ uint32_t bound_q15(float b)  // NaN, inf -> 32767
{
  const float y = b * LIMQ;
  return y < 32767.f ? (uint32_t)__float2int_ru(fmaxf(y, 0.f)) : 32767u;
}
__device__ __forceinline__ uint32_t sat_add_q15x2(uint32_t a, uint32_t lo, uint32_t hi)  // halfwords of a plus (lo, hi)
{
  return min((a & 0xffffu) + lo, 32767u) | (min((a >> 16) + hi, 32767u) << 16);
}
__device__ __forceinline__ uint2 point_limits_q15(const uint2 acc, const float (&m)[4])
{
  return make_uint2(sat_add_q15x2(acc.x, margin_q15(m[0]), margin_q15(m[1])), sat_add_q15x2(acc.y, margin_q15(m[2]), margin_q15(m[3])));
}
// Scan test for one point: with A = accumulated + 0x00010001 per word, (L | 0x80008000) - A keeps bit 15 of a halfword
// iff L >= A, i.e. iff the accumulated bound is still below the limit (no borrow crosses the halfwords: A <= 0x8000).
// Returns the AND over the point's four directions in bits 15 and 31 (both set <=> nothing reached).
__device__ __forceinline__ uint32_t limits_clear(const uint2 A, uint32_t Lu, uint32_t Lv)
{
  return ((Lu | 0x80008000u) - A.x) & ((Lv | 0x80008000u) - A.y);
}

// Upper bounds (pixels) of the image motion that one Gauss-Newton step causes, per direction (+u, -u, +v, -v), for
// ANY point whose template pixel lies in [ulo,uhi] x [vlo,vhi] and whose inverse depth lies in [zlo,zhi].
// With p = d (xn, yn, 1), zi = 1/d:  q_a = T_a p,  q_b = T_a (dT p)  =>  with s = q_a / d = R_a (xn,yn,1) + t_a zi,
// m = ((dR - I)(xn,yn,1) + dt zi),  g = R_a m = (q_b - q_a) / d, the projected x' = q_x / q_z moves by EXACTLY
//     dx' = (g_x - x'_a g_z) / (s_z + g_z),      x'_a = s_x / s_z,
// s, m, g are affine in (xn, yn, zi), so centre +- radius intervals bound them without slack; the products are
// bounded term by term; the denominator is positive.  The numerator's interval [c - r, c + r] gives the two signed
// bounds.  Each is inflated by 1 % + 1e-5 px, which covers the fp32 evaluation of the bound.
__device__ __forceinline__ float4 tile_motion_bound(const StepConsts & k, float ulo, float uhi, float vlo, float vhi, float zlo, float zhi)
{
  const float xlo = fmaf(k.ifx, ulo, k.kx), xhi = fmaf(k.ifx, uhi, k.kx);
  const float ylo = fmaf(k.ify, vlo, k.ky), yhi = fmaf(k.ify, vhi, k.ky);
  const float xc = 0.5f * (xlo + xhi), xr = 0.5f * fabsf(xhi - xlo) + 1e-6f;
  const float yc = 0.5f * (ylo + yhi), yr = 0.5f * fabsf(yhi - ylo) + 1e-6f;
  const float zc = 0.5f * (zlo + zhi), zr = 0.5f * (zhi - zlo) + 1e-5f * zhi;
  float mc[3], mr[3], gc[3], gr[3], sc[3], sr[3];
#pragma unroll
  for (int i = 0; i < 3; i++) {
    mc[i] = fmaf(k.E[3 * i], xc, fmaf(k.E[3 * i + 1], yc, fmaf(k.dt[i], zc, k.E[3 * i + 2])));
    mr[i] = fabsf(k.E[3 * i]) * xr + fabsf(k.E[3 * i + 1]) * yr + fabsf(k.dt[i]) * zr;
    sc[i] = fmaf(k.Ro[3 * i], xc, fmaf(k.Ro[3 * i + 1], yc, fmaf(k.to[i], zc, k.Ro[3 * i + 2])));
    sr[i] = fabsf(k.Ro[3 * i]) * xr + fabsf(k.Ro[3 * i + 1]) * yr + fabsf(k.to[i]) * zr;
  }
#pragma unroll
  for (int i = 0; i < 3; i++) {
    gc[i] = k.Ro[3 * i] * mc[0] + k.Ro[3 * i + 1] * mc[1] + k.Ro[3 * i + 2] * mc[2];
    gr[i] = fabsf(k.Ro[3 * i]) * mr[0] + fabsf(k.Ro[3 * i + 1]) * mr[1] + fabsf(k.Ro[3 * i + 2]) * mr[2];
  }
  const float INF = __int_as_float(0x7f800000);
  const float4 ALL = make_float4(INF, INF, INF, INF);
  const float szmin = sc[2] - sr[2];
  if (!(szmin > 0.25f)) return ALL;
  const float gz = fabsf(gc[2]) + gr[2];
  const float den = szmin - gz;
  if (!(den > 0.125f)) return ALL;
  const float isz = 1.f / sc[2], iszmin = 1.f / szmin;
  const float xpc = sc[0] * isz, xpr = (sr[0] + fabsf(xpc) * sr[2]) * iszmin;
  const float ypc = sc[1] * isz, ypr = (sr[1] + fabsf(ypc) * sr[2]) * iszmin;
  const float cu = fmaf(-xpc, gc[2], gc[0]), ru = gr[0] + fabsf(xpc) * gr[2] + xpr * gz;
  const float cv = fmaf(-ypc, gc[2], gc[1]), rv = gr[1] + fabsf(ypc) * gr[2] + ypr * gz;
  const float su = k.fx / den, sv = k.fy / den;
  // NaN propagates (fmaxf would drop it: keep the sums explicit) and flags everything through bound_q
  float4 b;
  b.x = fmaf(su * (ru + cu > 0.f ? ru + cu : (ru + cu == ru + cu ? 0.f : ru + cu)), 1.01f, 1e-5f);
  b.y = fmaf(su * (ru - cu > 0.f ? ru - cu : (ru - cu == ru - cu ? 0.f : ru - cu)), 1.01f, 1e-5f);
  b.z = fmaf(sv * (rv + cv > 0.f ? rv + cv : (rv + cv == rv + cv ? 0.f : rv + cv)), 1.01f, 1e-5f);
  b.w = fmaf(sv * (rv - cv > 0.f ? rv - cv : (rv - cv == rv - cv ? 0.f : rv - cv)), 1.01f, 1e-5f);
  return b;
}

}  // namespace vb
