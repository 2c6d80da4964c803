// Gauss-Newton iteration kernels of the inverse-compositional SE3 alignment.
//   residual kernel   : warp + nearest sample + residual + residual histogram; the last block of a view
//                       turns the histogram into the robust scale and a 511-entry weight table
//                       (InverseCompositional.cpp:113-136, Scaler.cpp:21-67, Loss.cpp:51-92)
//   accumulate kernel : 1x6 Jacobian, weights, 21+6+1 products, block partials; the last block of a job
//                       reduces all partials in fixed order and runs the 6x6 solve / pose update /
//                       termination logic (NormalEquations.cpp:53-61, GaussNewton.cpp:44-99,
//                       SE3Alignment.cpp:37-47,101-102)
// Residuals are u8 - u8, i.e. integers in [-255,255]: every scaler of the reference is a function of the
// residual histogram, so the "up to 30 passes over r" of ScalerTDistribution collapse into one pass plus
// a 30-step loop over 511 bins.
#include "common.cuh"

namespace vb
{
constexpr int THREADS = 256;

__device__ __forceinline__ double block_sum(double v, double * sm /* >= 8 doubles */)
{
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) sm[threadIdx.x >> 5] = v;
  __syncthreads();
  double s = 0;
#pragma unroll
  for (int w = 0; w < THREADS / 32; w++) s += sm[w];
  return s;
}

__device__ __forceinline__ double loss_weight(const SolverParams & p, double r)
{
  switch (p.lossKind) {
    case VSLAMB200_LOSS_TUKEY: {
      const double C = 4.6851;
      if (fabs(r) < C) {
        const double rc = r / C;
        const double q = 1.0 - rc * rc;
        return q * q;
      }
      return 0.0;
    }
    case VSLAMB200_LOSS_HUBER:
      if (fabs(r) < p.huberC) return 1.0;
      return (p.huberC * r > 0.0 ? 1.0 : -1.0) / r;
    case VSLAMB200_LOSS_TDIST:
      return (p.lossV + 1.0) / (p.lossV + r * r);
    default:
      return 1.0;
  }
}

// Scale / offset from the residual histogram (h[b] = #visible points with r = b-255; invisible points
// count as r = 0, InverseCompositional.cpp:111,130) and the weight table.  Runs in one block.
__device__ void scale_and_weights(ViewState & V, const SolverParams & p, int * h /* smem 512 */, double * red)
{
  const int tid = threadIdx.x;
  __shared__ double s_scale, s_offset;
  // total visible
  int c0 = h[tid] + h[tid + 256];
  double vis = block_sum((double)c0, red);
  const long n = V.nIp;
  __syncthreads();
  if (tid == 0) h[255] += (int)(n - (long)vis);  // zeros of the invisible points
  __syncthreads();
  double scale = 1.0, offset = 0.0;
  const double r0 = (double)(tid - 255), r1 = (double)(tid + 1);
  const double n0 = (double)h[tid], n1 = (double)h[tid + 256];
  if (p.scalerKind == VSLAMB200_SCALER_TDIST) {
    double step = 1.79769313486231570e308, sigma = 1.0;
    for (int it = 0; it < 30 && step > 1e-5; it++) {
      double term = 0.0;
      if (n0 > 0) {
        const double q = r0 / sigma;
        term += n0 * (r0 * r0 * (p.scalerV + 1) / (p.scalerV + q * q));
      }
      if (n1 > 0) {
        const double q = r1 / sigma;
        term += n1 * (r1 * r1 * (p.scalerV + 1) / (p.scalerV + q * q));
      }
      const double sum = block_sum(term, red);
      const double s = sqrt(sum / (double)n);
      step = fabs(sigma - s);
      sigma = s;
    }
    scale = sigma;
  } else if (p.scalerKind == VSLAMB200_SCALER_MEAN || p.scalerKind == VSLAMB200_SCALER_MEDIAN) {
    if (n > 0) {
      if (p.scalerKind == VSLAMB200_SCALER_MEAN) {
        const double sum = block_sum(n0 * r0 + n1 * r1, red);  // exact: integers < 2^53
        offset = sum / (double)n;
      } else {
        if (tid == 0) {
          // order statistics of the sorted residual vector (algorithm.cpp:109-120)
          long ka, kb;
          if (n % 2 == 0) {
            ka = n / 2 - 1;
            kb = n == 2 ? ka : n / 2 + 1;
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
      const double a = block_sum(n0 * fabs(r0 - offset) + n1 * fabs(r1 - offset), red);
      scale = sqrt(a / (double)(n - 1));
    }
  }
  if (tid == 0) {
    s_scale = scale;
    V.scale = scale;
    V.offset = offset;
  }
  __syncthreads();
  scale = s_scale;
  V.wlut[tid] = (float)loss_weight(p, (r0 - offset) / scale);
  if (tid < 255) V.wlut[tid + 256] = (float)loss_weight(p, (r1 - offset) / scale);
}

// Exact warp of one interest point, operation for operation the oracle's (fp64, explicit fma):
// Camera::image2camera (Camera.cpp:32-35), se3 * p (Warp.cpp:160-161), Camera::camera2image
// (Camera.cpp:23-30), visibility (InverseCompositional.cpp:115-116), std::round.
__device__ __noinline__ int warp_exact(const ViewState & V, int u, int v, float df)
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

// Residual kernel.  The sampled pixel index and the visibility flag are integer decisions, and they must
// be the fp64 oracle's.  A filtered evaluation gets them without fp64 in the common case: in exact
// arithmetic u' = u + fx*(Dx - xn*Dz)/(z + Dz) with D = (R - I) p + t, because p was back-projected from
// the integer pixel (u, v).  The displacement is small, so fp32 carries it with an absolute error far
// below a pixel; only when it lands within the error bound eps of a rounding (k + 0.5) or visibility
// (integer) boundary is the point re-evaluated by warp_exact().
template <int PPT>
__global__ void __launch_bounds__(THREADS, 4) residual_kernel(
  ViewState * __restrict__ views, const int * __restrict__ activeViews, SolverParams p)
{
  ViewState & V = views[activeViews[blockIdx.y]];
  if ((int)blockIdx.x >= V.nChunks) return;
  __shared__ int hist[HIST_BINS];
  __shared__ double red[THREADS / 32];
  __shared__ bool isLast;
  const int tid = threadIdx.x;
  hist[tid] = 0;
  hist[tid + 256] = 0;
  __syncthreads();

  const float r00 = V.rI[0], r01 = V.rI[1], r02 = V.rI[2], r10 = V.rI[3], r11 = V.rI[4], r12 = V.rI[5],
              r20 = V.rI[6], r21 = V.rI[7], r22 = V.rI[8];
  const float t0 = V.t32[0], t1 = V.t32[1], t2 = V.t32[2];
  const float fx = V.fx32, fy = V.fy32, ifx = V.ifxH, ify = V.ifyH, kx = V.kxH, ky = V.kyH;
  const float epsA = V.epsA, epsB = V.epsB;
  const int Wt = V.Wt, Ht = V.Ht;
  const uint8_t * __restrict__ I = V.I;
  const uint32_t * __restrict__ uvt = V.uvt;
  const float * __restrict__ dep = V.d;
  int16_t * __restrict__ rout = V.r;
  const int nIp = V.nIp;
  const int base = blockIdx.x * (THREADS * PPT);
  const bool needHist = p.lossKind != VSLAMB200_LOSS_NONE;

#pragma unroll 4
  for (int k = 0; k < PPT; k++) {
    const int idx = base + k * THREADS + tid;
    if (idx < nIp) {
      const uint32_t pk = __ldg(uvt + idx);
      const float d = __ldg(dep + idx);
      const int u = pk & 0xfff, v = (pk >> 12) & 0xfff, T = pk >> 24;
      const float xn = fmaf((float)u, ifx, kx), yn = fmaf((float)v, ify, ky);
      const float px = xn * d, py = yn * d;
      const float Dx = fmaf(r02, d, fmaf(r01, py, fmaf(r00, px, t0)));
      const float Dy = fmaf(r12, d, fmaf(r11, py, fmaf(r10, px, t1)));
      const float Dz = fmaf(r22, d, fmaf(r21, py, fmaf(r20, px, t2)));
      const float qz = d + Dz;
      const float iz = __frcp_rn(qz);
      const float du = fx * fmaf(-xn, Dz, Dx) * iz;
      const float dv = fy * fmaf(-yn, Dz, Dy) * iz;
      // distance of the displacement to the nearest multiple of 0.5 (rounding and visibility boundaries)
      const float fu = du + du, fv = dv + dv;
      const float distu = 0.5f * fabsf(fu - rintf(fu)), distv = 0.5f * fabsf(fv - rintf(fv));
      const float eps = fmaf(epsA, d, epsB) * fabsf(iz) + 3.8e-6f * (fabsf(du) + fabsf(dv)) + 6.1e-5f;
      int pix;
      if (!(qz > 0.25f * d) || !(fminf(distu, distv) > eps) || !(fabsf(du) + fabsf(dv) < 4096.f)) {
        pix = warp_exact(V, u, v, d);
      } else {
        // boundaries are integers: compare the displacement with exact integer offsets
        const bool vis = du > (float)(1 - u) && du < (float)(Wt - 1 - u) && dv > (float)(1 - v) && dv < (float)(Ht - 1 - v);
        const int iu = u + __float2int_rd(du + 0.5f), iv = v + __float2int_rd(dv + 0.5f);
        pix = vis ? iv * Wt + iu : -1;
      }
      int16_t rr = R_INVISIBLE;
      if (pix >= 0) {
        const int r = (int)__ldg(I + pix) - T;
        rr = (int16_t)r;
        if (needHist) atomicAdd(&hist[r + 255], 1);
      }
      rout[idx] = rr;
    }
  }
  if (!needHist) return;
  __syncthreads();
  {
    const int a = hist[tid], b = hist[tid + 256];
    if (a) atomicAdd(&V.hist[tid], a);
    if (b) atomicAdd(&V.hist[tid + 256], b);
  }
  __threadfence();
  __syncthreads();
  if (tid == 0) isLast = atomicAdd(&V.ticketA, 1u) == (unsigned)(V.nChunks - 1);
  __syncthreads();
  if (!isLast) return;
  __threadfence();
  hist[tid] = __ldcg(&V.hist[tid]);
  hist[tid + 256] = __ldcg(&V.hist[tid + 256]);
  V.hist[tid] = 0;
  V.hist[tid + 256] = 0;
  if (tid == 0) V.ticketA = 0;
  __syncthreads();
  scale_and_weights(V, p, hist, red);
}

void launch_residual(ViewState * views, const int * activeViews, int nActiveViews, int maxChunks, int chunk,
                     SolverParams p, cudaStream_t s)
{
  if (nActiveViews <= 0 || maxChunks <= 0) return;
  dim3 grid(maxChunks, nActiveViews);
  switch (chunk / THREADS) {
    case 2: residual_kernel<2><<<grid, THREADS, 0, s>>>(views, activeViews, p); break;
    case 4: residual_kernel<4><<<grid, THREADS, 0, s>>>(views, activeViews, p); break;
    case 8: residual_kernel<8><<<grid, THREADS, 0, s>>>(views, activeViews, p); break;
    default: residual_kernel<16><<<grid, THREADS, 0, s>>>(views, activeViews, p); break;
  }
}
}  // namespace vb
