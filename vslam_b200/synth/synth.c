/*
 * Deterministic synthetic RGB-D generator.
 *
 * Scene: a slanted textured plane n.X = d in the world frame (= the reference
 * camera frame).  A frame with pose T (P_frame = T * P_world, the convention of
 * the reference's Frame, core/src/Frame.h:94) is ray-cast analytically: pixel
 * ray K^-1 (u,v,1), moved to the world by T^-1, intersected with the plane,
 * texture sampled at the hit; depth = camera-z of the hit, rounded to float so
 * that host doubles and device floats hold the same value.
 *
 * This is shared test/bench input machinery, not part of the alignment path
 * and not part of the oracle.  Plain C, OpenMP over rows.
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

static uint64_t splitmix64(uint64_t *s)
{
  uint64_t z = (*s += 0x9E3779B97F4A7C15ull);
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
static double u01(uint64_t *s) { return (double)(splitmix64(s) >> 11) * (1.0 / 9007199254740992.0); }

static uint64_t hash2(int64_t a, int64_t b, uint64_t seed)
{
  uint64_t s = seed ^ ((uint64_t)a * 0x9E3779B97F4A7C15ull) ^ ((uint64_t)b * 0xC2B2AE3D27D4EB4Full);
  return splitmix64(&s);
}

/* tangent (v, w) -> R (row major 3x3), t;  standard SE3 exponential */
void vsynth_se3_exp(const double xi[6], double R[9], double t[3])
{
  const double wx = xi[3], wy = xi[4], wz = xi[5];
  const double th2 = wx * wx + wy * wy + wz * wz, th = sqrt(th2);
  double A, B, C;
  if (th < 1e-8) {
    A = 1.0 - th2 / 6.0;
    B = 0.5 - th2 / 24.0;
    C = 1.0 / 6.0 - th2 / 120.0;
  } else {
    A = sin(th) / th;
    B = (1.0 - cos(th)) / th2;
    C = (th - sin(th)) / (th2 * th);
  }
  const double W[9] = {0, -wz, wy, wz, 0, -wx, -wy, wx, 0};
  double W2[9];
  for (int i = 0; i < 3; i++)
    for (int j = 0; j < 3; j++) {
      double s = 0;
      for (int k = 0; k < 3; k++) s += W[i * 3 + k] * W[k * 3 + j];
      W2[i * 3 + j] = s;
    }
  double V[9];
  for (int i = 0; i < 9; i++) {
    const double I = (i % 4 == 0) ? 1.0 : 0.0;
    R[i] = I + A * W[i] + B * W2[i];
    V[i] = I + B * W[i] + C * W2[i];
  }
  for (int i = 0; i < 3; i++) t[i] = V[i * 3] * xi[0] + V[i * 3 + 1] * xi[1] + V[i * 3 + 2] * xi[2];
}

/* pair motion of SURVEY §8(d): v ~ U[-0.02,0.02]^3 m, w ~ U[-0.01,0.01]^3 rad */
void vsynth_pair_twist(uint64_t seed, double xi[6])
{
  uint64_t s = seed * 0xD1342543DE82EF95ull + 0x1234;
  for (int i = 0; i < 3; i++) xi[i] = (2.0 * u01(&s) - 1.0) * 0.02;
  for (int i = 3; i < 6; i++) xi[i] = (2.0 * u01(&s) - 1.0) * 0.01;
}

/*
 * Render one frame.
 *   R, t      pose of the frame (P_frame = R * P_world + t)
 *   gray      H*W u8 row-major out
 *   depth     H*W f32 row-major out
 *   holes     0: none; otherwise fraction*1000 of 8x8 blocks whose depth is zeroed
 */
void vsynth_render(
  int W, int H, double fx, double fy, double cx, double cy, const double R[9], const double t[3],
  uint64_t scene_seed, int holes_permille, uint64_t hole_seed, uint8_t *gray, float *depth)
{
  double nrm[3] = {0.10, -0.05, 1.0};
  const double nl = sqrt(nrm[0] * nrm[0] + nrm[1] * nrm[1] + nrm[2] * nrm[2]);
  for (int i = 0; i < 3; i++) nrm[i] /= nl;
  const double dpl = 2.0;

  double amp[5], freq[5], cs[5], sn[5], ph[5];
  uint64_t s = scene_seed;
  for (int k = 0; k < 5; k++) {
    const double ang = 2.0 * M_PI * u01(&s);
    cs[k] = cos(ang);
    sn[k] = sin(ang);
    ph[k] = 2.0 * M_PI * u01(&s);
    freq[k] = 1.5 * pow(2.0, k);
    amp[k] = 48.0 / pow(1.5, k);
  }
  /* world-frame quantities: camera centre c = -R^T t, n' = R n (plane normal in camera frame) */
  double Rt_t[3], nc[3];
  for (int i = 0; i < 3; i++) {
    Rt_t[i] = R[0 * 3 + i] * t[0] + R[1 * 3 + i] * t[1] + R[2 * 3 + i] * t[2];
    nc[i] = R[i * 3 + 0] * nrm[0] + R[i * 3 + 1] * nrm[1] + R[i * 3 + 2] * nrm[2];
  }
  const double num = dpl + nrm[0] * Rt_t[0] + nrm[1] * Rt_t[1] + nrm[2] * Rt_t[2];

#pragma omp parallel for schedule(static)
  for (int v = 0; v < H; v++) {
    for (int u = 0; u < W; u++) {
      const double dx = (u - cx) / fx, dy = (v - cy) / fy;
      const double den = nc[0] * dx + nc[1] * dy + nc[2];
      double lam = num / den;
      uint8_t g = 0;
      float z = 0.f;
      if (den > 1e-9 && lam > 0) {
        const double pc[3] = {lam * dx - t[0], lam * dy - t[1], lam - t[2]};
        const double X = R[0] * pc[0] + R[3] * pc[1] + R[6] * pc[2];
        const double Y = R[1] * pc[0] + R[4] * pc[1] + R[7] * pc[2];
        double val = 127.5;
        for (int k = 0; k < 5; k++) val += amp[k] * sin(2.0 * M_PI * freq[k] * (cs[k] * X + sn[k] * Y) + ph[k]);
        const int64_t gx = (int64_t)floor(X / 0.004), gy = (int64_t)floor(Y / 0.004);
        const uint64_t hh = hash2(gx, gy, scene_seed ^ 0xABCDEFull);
        val += ((double)(hh >> 11) * (1.0 / 9007199254740992.0) * 2.0 - 1.0) * 6.0;
        if (val < 0) val = 0;
        if (val > 255) val = 255;
        g = (uint8_t)floor(val + 0.5);
        z = (float)lam;
      }
      if (holes_permille > 0) {
        const uint64_t hb = hash2(u / 8, v / 8, hole_seed);
        if ((int)(hb % 1000) < holes_permille) z = 0.f;
      }
      gray[(size_t)v * W + u] = g;
      depth[(size_t)v * W + u] = z;
    }
  }
}

/* smooth sequence twist of SURVEY §8(d) config 2: xi_t = 0.5*xi_max*sin(0.05 t + phi_j), absolute pose = exp(sum) */
void vsynth_sequence_pose(uint64_t seed, int tframe, double R[9], double t[3])
{
  /* absolute pose built by composing per-frame relative twists so that inter-frame motion stays in the TUM envelope */
  uint64_t s = seed * 0x9E3779B97F4A7C15ull + 77;
  double phi[6];
  for (int j = 0; j < 6; j++) phi[j] = 2.0 * M_PI * u01(&s);
  double Rc[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, tc[3] = {0, 0, 0};
  for (int k = 1; k <= tframe; k++) {
    double xi[6], Rk[9], tk[3];
    for (int j = 0; j < 6; j++) xi[j] = 0.5 * (j < 3 ? 0.02 : 0.01) * sin(0.05 * k + phi[j]);
    vsynth_se3_exp(xi, Rk, tk);
    /* T_k = exp(xi_k) * T_{k-1} */
    double Rn[9], tn[3];
    for (int i = 0; i < 3; i++) {
      for (int j = 0; j < 3; j++) {
        double a = 0;
        for (int m = 0; m < 3; m++) a += Rk[i * 3 + m] * Rc[m * 3 + j];
        Rn[i * 3 + j] = a;
      }
      tn[i] = Rk[i * 3] * tc[0] + Rk[i * 3 + 1] * tc[1] + Rk[i * 3 + 2] * tc[2] + tk[i];
    }
    memcpy(Rc, Rn, sizeof(Rn));
    memcpy(tc, tn, sizeof(tn));
  }
  memcpy(R, Rc, sizeof(Rc));
  memcpy(t, tc, sizeof(tc));
}
