// Filtered SE3 warp of one interest point: fp32 evaluation with a rigorous error bound, exact fp64 fallback.
// Shared by the Gauss-Newton level kernels (Warp.cpp:158-165, Camera.cpp:23-35, InverseCompositional.cpp:115-120).
#pragma once
#include "common.cuh"

namespace vb
{
// Exact warp of one interest point, operation for operation the oracle's (fp64, explicit fma):
// Camera::image2camera (Camera.cpp:32-35), se3 * p (Warp.cpp:160-161), Camera::camera2image
// (Camera.cpp:23-30), visibility (InverseCompositional.cpp:115-116), std::round.
static __device__ __noinline__ int warp_exact(const ViewState & V, int u, int v, float df)
{
  const double d = (double)df;
  const double X = __fma_rn(V.kx, d, __dmul_rn(V.ifx, __dmul_rn((double)u, d)));
  const double Y = __fma_rn(V.ky, d, __dmul_rn(V.ify, __dmul_rn((double)v, d)));
  const double qx = __fma_rn(V.R[2], d, __fma_rn(V.R[1], Y, __fma_rn(V.R[0], X, V.t[0])));
  const double qy = __fma_rn(V.R[5], d, __fma_rn(V.R[4], Y, __fma_rn(V.R[3], X, V.t[1])));
  const double qz = __fma_rn(V.R[8], d, __fma_rn(V.R[7], Y, __fma_rn(V.R[6], X, V.t[2])));
  if (!(qz > 0.0)) return -1;
  const double up = __ddiv_rn(__fma_rn(V.fx, qx, __dmul_rn(V.cx, qz)), qz);
  const double vp = __ddiv_rn(__fma_rn(V.fy, qy, __dmul_rn(V.cy, qz)), qz);
  if (!(1.0 < up && up < (double)(V.Wt - 1) && 1.0 < vp && vp < (double)(V.Ht - 1) && isfinite(up))) return -1;
  // std::round (half away from zero) of a positive value == floor(x + 0.5)
  return __double2int_rd(vp + 0.5) * V.Wt + __double2int_rd(up + 0.5);
}

struct WarpConsts
{
  float r00, r01, r02, r10, r11, r12, r20, r21, r22, t0, t1, t2, fx, fy, ifx, ify, kx, ky, epsA, epsB;
  int Wt, Ht;
};

// Sampled target pixel index of one interest point, or -1 when it is not visible.  The index and the
// visibility flag are integer decisions and must be the fp64 oracle's.  A filtered evaluation gets them
// without fp64 in the common case: in exact arithmetic u' = u + fx*(Dx - xn*Dz)/(z + Dz) with
// D = (R - I) p + t, because p was back-projected from the integer pixel (u, v).  The displacement du is
// small, so fp32 carries it with an absolute error far below a pixel.  With iu = u + rint(du) and the exact
// fp32 remainder fu = du - rint(du) in [-0.5, 0.5], u' = iu + fu, so
//   - the rounded target round(u') is iu unless |fu| is within eps of 0.5,
//   - visibility 1 < u' < W-1 is decided by iu alone, except on the ring iu == 1 (visible iff fu > 0) and
//     iu == W-1 (visible iff fu < 0), where it is safe unless |fu| <= eps.
// warp_eval() is branch-free (the four points of a thread interleave).  When every rounded target of the warp lies
// strictly inside the ring (the common case, decided by two unsigned compares per point) warp_pick_interior()
// finishes with the rounding distance alone; otherwise warp_pick_ring() adds the visibility logic.  Both report
// `ok`; the few points within eps of a boundary are redone by warp_exact().
struct WarpEval
{
  float fu, fv, eps;
  int iu, iv;
  bool okb;       // depth / magnitude guards of the fp32 evaluation
  bool interior;  // 2 <= iu <= W-2 and 2 <= iv <= H-2
};
__device__ __forceinline__ WarpEval warp_eval(const WarpConsts & c, uint32_t pk, float d)
{
  WarpEval e;
  const int u = pk & 0xfff, v = (pk >> 12) & 0xfff;
  const float xn = fmaf((float)u, c.ifx, c.kx), yn = fmaf((float)v, c.ify, c.ky);
  const float px = xn * d, py = yn * d;
  const float Dx = fmaf(c.r02, d, fmaf(c.r01, py, fmaf(c.r00, px, c.t0)));
  const float Dy = fmaf(c.r12, d, fmaf(c.r11, py, fmaf(c.r10, px, c.t1)));
  const float Dz = fmaf(c.r22, d, fmaf(c.r21, py, fmaf(c.r20, px, c.t2)));
  const float qz = d + Dz;
  float iz;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(iz) : "f"(qz));  // 1 ulp; covered by the 2^-18 (|du|+|dv|) term of eps
  const float du = (c.fx * iz) * fmaf(-xn, Dz, Dx);
  const float dv = (c.fy * iz) * fmaf(-yn, Dz, Dy);
  // nearest integer of the displacement through the 1.5 * 2^23 trick (|du| < 4096 is part of okb)
  const float MAGIC = 12582912.f;
  const float mu = du + MAGIC, mv = dv + MAGIC;
  e.fu = du - (mu - MAGIC);
  e.fv = dv - (mv - MAGIC);
  e.iu = u + (__float_as_int(mu) - 0x4B400000);
  e.iv = v + (__float_as_int(mv) - 0x4B400000);
  const float s = fabsf(du) + fabsf(dv);
  e.eps = fmaf(fmaf(c.epsA, d, c.epsB), fabsf(iz), fmaf(3.8e-6f, s, 6.1e-5f));
  e.okb = qz > 0.25f * d && s < 4096.f;
  e.interior = (unsigned)(e.iu - 2) <= (unsigned)(c.Wt - 4) && (unsigned)(e.iv - 2) <= (unsigned)(c.Ht - 4);
  return e;
}
__device__ __forceinline__ int warp_pick_interior(const WarpConsts & c, const WarpEval & e, bool & ok)
{
  ok = e.okb && 0.5f - fmaxf(fabsf(e.fu), fabsf(e.fv)) > e.eps;  // distance to the nearest rounding boundary
  return e.iv * c.Wt + e.iu;
}
__device__ __forceinline__ int warp_pick_ring(const WarpConsts & c, const WarpEval & e, bool & ok)
{
  const int iu = e.iu, iv = e.iv;
  const float au = fabsf(e.fu), av = fabsf(e.fv);
  const bool ringu = iu == 1 || iu == c.Wt - 1, ringv = iv == 1 || iv == c.Ht - 1;
  float dist = 0.5f - fmaxf(au, av);
  dist = ringu ? fminf(dist, au) : dist;  // on the ring: distance to the visibility boundary as well
  dist = ringv ? fminf(dist, av) : dist;
  ok = e.okb && dist > e.eps;
  bool vis = (unsigned)(iu - 1) <= (unsigned)(c.Wt - 2) && (unsigned)(iv - 1) <= (unsigned)(c.Ht - 2);
  vis = vis && !(iu == 1 && e.fu < 0.f) && !(iu == c.Wt - 1 && e.fu > 0.f);
  vis = vis && !(iv == 1 && e.fv < 0.f) && !(iv == c.Ht - 1 && e.fv > 0.f);
  return vis ? iv * c.Wt + iu : -1;
}
}  // namespace vb
