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

template <int PPT>
__global__ void __launch_bounds__(THREADS) residual_kernel(
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

  const double R0 = V.R[0], R1 = V.R[1], R2 = V.R[2], R3 = V.R[3], R4 = V.R[4], R5 = V.R[5], R6 = V.R[6],
               R7 = V.R[7], R8 = V.R[8];
  const double t0 = V.t[0], t1 = V.t[1], t2 = V.t[2];
  const double fx = V.fx, fy = V.fy, cx = V.cx, cy = V.cy;
  const double ifx = V.ifx, ify = V.ify, kx = V.kx, ky = V.ky;
  const double umax = (double)(V.Wt - 1), vmax = (double)(V.Ht - 1);
  const int Wt = V.Wt;
  const uint8_t * __restrict__ I = V.I;
  const uint32_t * __restrict__ uvt = V.uvt;
  const float * __restrict__ dep = V.d;
  int16_t * __restrict__ rout = V.r;
  const int nIp = V.nIp;
  const int base = blockIdx.x * (THREADS * PPT);
  const bool needHist = p.lossKind != VSLAMB200_LOSS_NONE;

#pragma unroll
  for (int k = 0; k < PPT; k++) {
    const int idx = base + k * THREADS + tid;
    if (idx < nIp) {
      const uint32_t pk = __ldg(uvt + idx);
      const double d = (double)__ldg(dep + idx);
      const int u = pk & 0xfff, v = (pk >> 12) & 0xfff, T = pk >> 24;
      // Camera::image2camera (Camera.cpp:32-35), same operation order as the oracle
      const double X = __fma_rn(kx, d, __dmul_rn(ifx, __dmul_rn((double)u, d)));
      const double Y = __fma_rn(ky, d, __dmul_rn(ify, __dmul_rn((double)v, d)));
      // se3 * p (Warp.cpp:160-161)
      const double qx = __fma_rn(R2, d, __fma_rn(R1, Y, __fma_rn(R0, X, t0)));
      const double qy = __fma_rn(R5, d, __fma_rn(R4, Y, __fma_rn(R3, X, t1)));
      const double qz = __fma_rn(R8, d, __fma_rn(R7, Y, __fma_rn(R6, X, t2)));
      int16_t rr = R_INVISIBLE;
      if (qz > 0.0) {
        // Camera::camera2image (Camera.cpp:23-30)
        const double up = __ddiv_rn(__fma_rn(fx, qx, __dmul_rn(cx, qz)), qz);
        const double vp = __ddiv_rn(__fma_rn(fy, qy, __dmul_rn(cy, qz)), qz);
        // InverseCompositional.cpp:115-116
        if (1.0 < up && up < umax && 1.0 < vp && vp < vmax && isfinite(up)) {
          // std::round (half away from zero) of a positive value == floor(x + 0.5)
          const int iu = __double2int_rd(up + 0.5), iv = __double2int_rd(vp + 0.5);
          const int r = (int)__ldg(I + (size_t)iv * Wt + iu) - T;
          rr = (int16_t)r;
          if (needHist) atomicAdd(&hist[r + 255], 1);
        }
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
