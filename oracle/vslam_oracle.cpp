/*
 * ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path.
 *
 * Plain C++17 CPU restatement (no Eigen / Sophus / OpenCV) of the dense RGB-D
 * direct-alignment hot path of phildue/vslam.  Only tests/, __graft_entry__.smoke()
 * and bench.py's cpu_baseline / --impl reference legs may build, load or call this.
 *
 * The reference itself cannot be compiled in this image (Eigen, Sophus, OpenCV C++,
 * easylogging++, matplotlib-cpp are absent), so every function below restates the
 * reference source it cites (paths relative to /root/reference/src/vslam/src/).
 *
 * PARITY PINNING
 *  - pyr_down_u8 / blur3 / sobel3 restate OpenCV (cv::buildPyramid, cv::GaussianBlur 3x3,
 *    cv::Sobel CV_16S, called at core/src/Frame.cpp:264,225,227-228).  Pinned bit-exactly
 *    against cv2 4.13 outputs committed under tests/golden/ (tests/golden/make_golden.py).
 *  - bilinear / resize / median / camera projection / warp Jacobian are pinned by the
 *    known-answer structures of the reference's own tests (core/test/test_algorithm.cpp:27-33,
 *    57-68; core/test/test_camera.cpp:59-111; lukas_kanade/test/test_warp.cpp:29-41).
 *  - The solver path (InverseCompositional / GaussNewton / SE3Alignment) has NO golden
 *    vectors in the reference (all its hot-path tests are DISABLED_ and need a dataset that
 *    is not shipped): parity unpinned for that part; Sophus/Eigen internals are restated from
 *    their published closed forms, so agreement with a real reference build is
 *    tolerance-level, not bitwise.
 *
 * Floating point: built with -ffp-contract=off; fused operations are written as fma()
 * explicitly so that the device path can match the warp arithmetic bit for bit.
 */
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <memory>
#include <vector>

namespace vo
{
using std::size_t;

// ------------------------------------------------------------------------------------------
// small fixed-size linear algebra
// ------------------------------------------------------------------------------------------
struct Mat6
{
  double a[36];
  double & operator()(int i, int j) { return a[i * 6 + j]; }
  double operator()(int i, int j) const { return a[i * 6 + j]; }
};
struct Vec6
{
  double v[6];
  double & operator[](int i) { return v[i]; }
  double operator[](int i) const { return v[i]; }
};

// determinant by Gaussian elimination with partial pivoting (Eigen uses PartialPivLU for 6x6,
// least_squares/src/GaussNewton.cpp:52)
static double det6(const Mat6 & A)
{
  double m[36];
  std::memcpy(m, A.a, sizeof(m));
  double det = 1.0;
  for (int c = 0; c < 6; c++) {
    int p = c;
    for (int r = c + 1; r < 6; r++)
      if (std::fabs(m[r * 6 + c]) > std::fabs(m[p * 6 + c])) p = r;
    if (m[p * 6 + c] == 0.0) return 0.0;
    if (p != c) {
      for (int k = 0; k < 6; k++) std::swap(m[p * 6 + k], m[c * 6 + k]);
      det = -det;
    }
    det *= m[c * 6 + c];
    for (int r = c + 1; r < 6; r++) {
      const double f = m[r * 6 + c] / m[c * 6 + c];
      for (int k = c; k < 6; k++) m[r * 6 + k] -= f * m[c * 6 + k];
    }
  }
  return det;
}

// A x = b via LDL^T (least_squares/src/GaussNewton.cpp:76, Eigen ldlt(); Eigen pivots on the
// diagonal, which changes only rounding for the SPD systems met here)
static Vec6 ldlt_solve(const Mat6 & A, const Vec6 & b)
{
  double L[36] = {0}, D[6];
  for (int j = 0; j < 6; j++) {
    double d = A(j, j);
    for (int k = 0; k < j; k++) d -= L[j * 6 + k] * L[j * 6 + k] * D[k];
    D[j] = d;
    L[j * 6 + j] = 1.0;
    for (int i = j + 1; i < 6; i++) {
      double s = A(i, j);
      for (int k = 0; k < j; k++) s -= L[i * 6 + k] * L[j * 6 + k] * D[k];
      L[i * 6 + j] = s / d;
    }
  }
  double y[6];
  for (int i = 0; i < 6; i++) {
    double s = b[i];
    for (int k = 0; k < i; k++) s -= L[i * 6 + k] * y[k];
    y[i] = s;
  }
  Vec6 x;
  for (int i = 5; i >= 0; i--) {
    double s = y[i] / D[i];
    for (int k = i + 1; k < 6; k++) s -= L[k * 6 + i] * x[k];
    x[i] = s;
  }
  return x;
}

// inverse by Gauss-Jordan with partial pivoting (SE3Alignment.cpp:101 A().inverse())
static Mat6 inverse6(const Mat6 & A)
{
  double m[6][12];
  for (int i = 0; i < 6; i++)
    for (int j = 0; j < 6; j++) {
      m[i][j] = A(i, j);
      m[i][6 + j] = i == j ? 1.0 : 0.0;
    }
  for (int c = 0; c < 6; c++) {
    int p = c;
    for (int r = c + 1; r < 6; r++)
      if (std::fabs(m[r][c]) > std::fabs(m[p][c])) p = r;
    if (p != c)
      for (int k = 0; k < 12; k++) std::swap(m[p][k], m[c][k]);
    const double inv = 1.0 / m[c][c];
    for (int k = 0; k < 12; k++) m[c][k] *= inv;
    for (int r = 0; r < 6; r++) {
      if (r == c) continue;
      const double f = m[r][c];
      if (f == 0.0) continue;
      for (int k = 0; k < 12; k++) m[r][k] -= f * m[c][k];
    }
  }
  Mat6 R;
  for (int i = 0; i < 6; i++)
    for (int j = 0; j < 6; j++) R(i, j) = m[i][6 + j];
  return R;
}

// ------------------------------------------------------------------------------------------
// SE3 = unit quaternion (x,y,z,w) + translation, Sophus conventions (core/src/types.h:39;
// tangent order [upsilon; omega]).  Closed forms restated from Sophus' so3.hpp / se3.hpp.
// ------------------------------------------------------------------------------------------
struct SE3
{
  double q[4] = {0, 0, 0, 1};  // x y z w
  double t[3] = {0, 0, 0};
};

static void quat_to_R(const double q[4], double R[9])
{
  const double x = q[0], y = q[1], z = q[2], w = q[3];
  const double tx = 2 * x, ty = 2 * y, tz = 2 * z;
  const double twx = tx * w, twy = ty * w, twz = tz * w;
  const double txx = tx * x, txy = ty * x, txz = tz * x;
  const double tyy = ty * y, tyz = tz * y, tzz = tz * z;
  R[0] = 1 - (tyy + tzz);
  R[1] = txy - twz;
  R[2] = txz + twy;
  R[3] = txy + twz;
  R[4] = 1 - (txx + tzz);
  R[5] = tyz - twx;
  R[6] = txz - twy;
  R[7] = tyz + twx;
  R[8] = 1 - (txx + tyy);
}
static void quat_mul(const double a[4], const double b[4], double o[4])
{
  const double ax = a[0], ay = a[1], az = a[2], aw = a[3];
  const double bx = b[0], by = b[1], bz = b[2], bw = b[3];
  o[0] = aw * bx + ax * bw + ay * bz - az * by;
  o[1] = aw * by + ay * bw + az * bx - ax * bz;
  o[2] = aw * bz + az * bw + ax * by - ay * bx;
  o[3] = aw * bw - ax * bx - ay * by - az * bz;
}
static void quat_normalize(double q[4])
{
  const double n = std::sqrt(q[0] * q[0] + q[1] * q[1] + q[2] * q[2] + q[3] * q[3]);
  for (int i = 0; i < 4; i++) q[i] /= n;
}
static void rotate(const double q[4], const double p[3], double o[3])
{
  double R[9];
  quat_to_R(q, R);
  for (int i = 0; i < 3; i++) o[i] = R[i * 3] * p[0] + R[i * 3 + 1] * p[1] + R[i * 3 + 2] * p[2];
}
static SE3 se3_mul(const SE3 & a, const SE3 & b)
{
  SE3 o;
  quat_mul(a.q, b.q, o.q);
  quat_normalize(o.q);
  double rt[3];
  rotate(a.q, b.t, rt);
  for (int i = 0; i < 3; i++) o.t[i] = rt[i] + a.t[i];
  return o;
}
static SE3 se3_inv(const SE3 & a)
{
  SE3 o;
  o.q[0] = -a.q[0];
  o.q[1] = -a.q[1];
  o.q[2] = -a.q[2];
  o.q[3] = a.q[3];
  double rt[3];
  rotate(o.q, a.t, rt);
  for (int i = 0; i < 3; i++) o.t[i] = -rt[i];
  return o;
}
static SE3 se3_exp(const Vec6 & xi)
{
  const double wx = xi[3], wy = xi[4], wz = xi[5];
  const double th2 = wx * wx + wy * wy + wz * wz;
  const double th = std::sqrt(th2);
  double imag, real;
  if (th2 < 1e-20) {
    const double th4 = th2 * th2;
    imag = 0.5 - th2 / 48.0 + th4 / 3840.0;
    real = 1.0 - th2 / 8.0 + th4 / 384.0;
  } else {
    imag = std::sin(0.5 * th) / th;
    real = std::cos(0.5 * th);
  }
  SE3 o;
  o.q[0] = imag * wx;
  o.q[1] = imag * wy;
  o.q[2] = imag * wz;
  o.q[3] = real;
  double B, C;
  if (th2 < 1e-20) {
    B = 0.5;
    C = 1.0 / 6.0;
  } else {
    B = (1.0 - std::cos(th)) / th2;
    C = (th - std::sin(th)) / (th2 * th);
  }
  // V = I + B*W + C*W^2,  t = V * upsilon
  const double ux = xi[0], uy = xi[1], uz = xi[2];
  const double c1[3] = {wy * uz - wz * uy, wz * ux - wx * uz, wx * uy - wy * ux};           // w x u
  const double c2[3] = {wy * c1[2] - wz * c1[1], wz * c1[0] - wx * c1[2], wx * c1[1] - wy * c1[0]};  // w x (w x u)
  o.t[0] = ux + B * c1[0] + C * c2[0];
  o.t[1] = uy + B * c1[1] + C * c2[1];
  o.t[2] = uz + B * c1[2] + C * c2[2];
  return o;
}
static Vec6 se3_log(const SE3 & T)
{
  const double n2 = T.q[0] * T.q[0] + T.q[1] * T.q[1] + T.q[2] * T.q[2];
  const double n = std::sqrt(n2), w = T.q[3];
  double f;
  if (n2 < 1e-20) {
    f = 2.0 / w - 2.0 / 3.0 * n2 / (w * w * w);
  } else if (std::fabs(w) < 1e-10) {
    f = (w > 0 ? M_PI : -M_PI) / n;
  } else {
    f = 2.0 * std::atan(n / w) / n;
  }
  const double wx = f * T.q[0], wy = f * T.q[1], wz = f * T.q[2];
  const double th2 = wx * wx + wy * wy + wz * wz, th = std::sqrt(th2);
  double c;
  if (th2 < 1e-12) {
    c = 1.0 / 12.0;
  } else {
    const double hs = std::sin(0.5 * th), hc = std::cos(0.5 * th);
    c = (1.0 - th * hc / (2.0 * hs)) / th2;
  }
  // upsilon = V^-1 t = t - 0.5 w x t + c w x (w x t)
  const double * t = T.t;
  const double c1[3] = {wy * t[2] - wz * t[1], wz * t[0] - wx * t[2], wx * t[1] - wy * t[0]};
  const double c2[3] = {wy * c1[2] - wz * c1[1], wz * c1[0] - wx * c1[2], wx * c1[1] - wy * c1[0]};
  Vec6 xi;
  xi[0] = t[0] - 0.5 * c1[0] + c * c2[0];
  xi[1] = t[1] - 0.5 * c1[1] + c * c2[1];
  xi[2] = t[2] - 0.5 * c1[2] + c * c2[2];
  xi[3] = wx;
  xi[4] = wy;
  xi[5] = wz;
  return xi;
}

// ------------------------------------------------------------------------------------------
// image operations
// ------------------------------------------------------------------------------------------
static inline int reflect101(int i, int n)
{
  if (n == 1) return 0;
  while (i < 0 || i >= n) {
    if (i < 0) i = -i;
    if (i >= n) i = 2 * n - 2 - i;
  }
  return i;
}

// Frame ingestion of the node (src/ros/nodes/NodeRgbdAlignment.cpp:177-199): cv::cvtColor(mat, mat, COLOR_RGB2GRAY)
// on an 8UC3 image (:182).  OpenCV's 8-bit path is the integer formula below (15-bit coefficients 9798 / 19235 /
// 3735, round to nearest); pinned against cv2 4.13.0 (tests/golden/cv_ops.npz).
static void rgb_to_gray_u8(const uint8_t * rgb, size_t n, int bgr, uint8_t * gray)
{
  for (size_t i = 0; i < n; i++) {
    const int c0 = rgb[3 * i], c1 = rgb[3 * i + 1], c2 = rgb[3 * i + 2];
    const int r = bgr ? c2 : c0, b = bgr ? c0 : c2;
    gray[i] = (uint8_t)((r * 9798 + c1 * 19235 + b * 3735 + (1 << 14)) >> 15);
  }
}
// depth = depth.array().isNaN().select(0, depth) (NodeRgbdAlignment.cpp:190, utils.cpp:74); 16-bit sensor depth is
// divided by the dataset's factor in double (loadDepth(...) / 5000.0, odometry/test/test_alignment_se3.cpp:99)
static void depth_ingest(const double * in, size_t n, double divisor, double * out)
{
  for (size_t i = 0; i < n; i++) {
    const double v = in[i] / divisor;
    out[i] = std::isnan(v) ? 0.0 : v;
  }
}

// cv::pyrDown as used by cv::buildPyramid (core/src/Frame.cpp:264):
// 5x5 binomial [1 4 6 4 1]^2, reflect-101, (s+128)>>8, even rows/cols, size ((W+1)/2,(H+1)/2)
static void pyr_down_u8(const uint8_t * src, int W, int H, uint8_t * dst)
{
  const int Wd = (W + 1) / 2, Hd = (H + 1) / 2;
  static const int k[5] = {1, 4, 6, 4, 1};
  std::vector<int> row((size_t)5 * Wd);
  for (int y = 0; y < Hd; y++) {
    for (int i = 0; i < 5; i++) {
      const uint8_t * s = src + (size_t)reflect101(2 * y + i - 2, H) * W;
      int * r = row.data() + (size_t)i * Wd;
      for (int x = 0; x < Wd; x++) {
        int acc = 0;
        for (int j = 0; j < 5; j++) acc += k[j] * s[reflect101(2 * x + j - 2, W)];
        r[x] = acc;
      }
    }
    for (int x = 0; x < Wd; x++) {
      int acc = 0;
      for (int i = 0; i < 5; i++) acc += k[i] * row[(size_t)i * Wd + x];
      dst[(size_t)y * Wd + x] = (uint8_t)((acc + 128) >> 8);
    }
  }
}

// cv::GaussianBlur(3x3, sigma 0, BORDER_DEFAULT) on u8: ([1 2 1]^2 + 8) >> 4 (core/src/Frame.cpp:225)
static void blur3_u8(const uint8_t * src, int W, int H, uint8_t * dst)
{
  for (int y = 0; y < H; y++) {
    const uint8_t * r0 = src + (size_t)reflect101(y - 1, H) * W;
    const uint8_t * r1 = src + (size_t)y * W;
    const uint8_t * r2 = src + (size_t)reflect101(y + 1, H) * W;
    for (int x = 0; x < W; x++) {
      const int xm = reflect101(x - 1, W), xp = reflect101(x + 1, W);
      const int s = (r0[xm] + 2 * r0[x] + r0[xp]) + 2 * (r1[xm] + 2 * r1[x] + r1[xp]) +
                    (r2[xm] + 2 * r2[x] + r2[xp]);
      dst[(size_t)y * W + x] = (uint8_t)((s + 8) >> 4);
    }
  }
}

// cv::Sobel(CV_16S, ksize 3) dx and dy, reflect-101, unscaled (core/src/Frame.cpp:227-228)
static void sobel3_s16(const uint8_t * g, int W, int H, int16_t * dx, int16_t * dy)
{
  for (int y = 0; y < H; y++) {
    const uint8_t * r0 = g + (size_t)reflect101(y - 1, H) * W;
    const uint8_t * r1 = g + (size_t)y * W;
    const uint8_t * r2 = g + (size_t)reflect101(y + 1, H) * W;
    for (int x = 0; x < W; x++) {
      const int xm = reflect101(x - 1, W), xp = reflect101(x + 1, W);
      dx[(size_t)y * W + x] =
        (int16_t)((r0[xp] - r0[xm]) + 2 * (r1[xp] - r1[xm]) + (r2[xp] - r2[xm]));
      dy[(size_t)y * W + x] =
        (int16_t)((r2[xm] + 2 * r2[x] + r2[xp]) - (r0[xm] + 2 * r0[x] + r0[xp]));
    }
  }
}

// algorithm::bilinearInterpolation (core/src/algorithm.h:37-77) on a row-major plane;
// keeps the reference's Q12/Q21 transposition
template <typename T>
static double bilinear_ref(const T * m, int W, double x, double y)
{
  const double x1 = std::floor(x), x2 = std::ceil(x), y1 = std::floor(y), y2 = std::ceil(y);
  const double Q11 = m[(size_t)(int)y1 * W + (int)x1];
  const double Q12 = m[(size_t)(int)y1 * W + (int)x2];
  const double Q21 = m[(size_t)(int)y2 * W + (int)x1];
  const double Q22 = m[(size_t)(int)y2 * W + (int)x2];
  double R1, R2;
  if (x2 == x1) {
    R1 = Q11;
    R2 = Q12;
  } else {
    R1 = ((x2 - x) / (x2 - x1)) * Q11 + ((x - x1) / (x2 - x1)) * Q21;
    R2 = ((x2 - x) / (x2 - x1)) * Q12 + ((x - x1) / (x2 - x1)) * Q22;
  }
  if (y2 == y1) return R1;
  return ((y2 - y) / (y2 - y1)) * R1 + ((y - y1) / (y2 - y1)) * R2;
}

// algorithm::median (core/src/algorithm.cpp:109-120): even n -> (v[n/2-1] + v[n/2+1])/2.
// n == 2 reads v[2] out of bounds in the reference (undefined): the build's documented rule is v[0].
static double median_ref(std::vector<double> & v, long * n2_hits)
{
  std::sort(v.begin(), v.end());
  const int n = (int)v.size();
  if (n % 2 == 0) {
    if (n == 2) {
      if (n2_hits) (*n2_hits)++;
      return v[0];
    }
    return (v[n / 2 - 1] + v[n / 2 + 1]) / 2;
  }
  return v[n / 2];
}

// one depth pyramid step: algorithm::medianBlur(d,3,3, v<=0 masked) (core/src/algorithm.h:157-184)
// then algorithm::resize(.,0.5) (algorithm.h:84-101) == blur(2i,2j); size (int(H*0.5), int(W*0.5)).
// The full-resolution blur is materialised like the reference does (needed for the CPU baseline to
// be a fair restatement); f32_storage rounds the output to float (device storage format).
static void depth_down(
  const double * src, int W, int H, double * dst, int f32_storage, long * n2_hits)
{
  std::vector<double> blur((size_t)W * H, 0.0);
  std::vector<double> vals;
  vals.reserve(9);
  for (int i = 1; i < H - 1; i++) {
    for (int j = 1; j < W - 1; j++) {
      vals.clear();
      for (int ki = i - 1; ki <= i + 1; ki++)
        for (int kj = j - 1; kj <= j + 1; kj++) {
          const double v = src[(size_t)ki * W + kj];
          if (!(v <= 0.0)) vals.push_back(v);
        }
      // the n==2 counter only counts sites that survive the decimation
      long * cnt = (i % 2 == 0 && j % 2 == 0) ? n2_hits : nullptr;
      blur[(size_t)i * W + j] = vals.empty() ? 0.0 : median_ref(vals, cnt);
    }
  }
  const int Hd = (int)(H * 0.5), Wd = (int)(W * 0.5);
  for (int i = 0; i < Hd; i++)
    for (int j = 0; j < Wd; j++) {
      double v = bilinear_ref(blur.data(), W, j * 2.0, i * 2.0);
      if (f32_storage) v = (double)(float)v;
      dst[(size_t)i * Wd + j] = v;
    }
}

// ------------------------------------------------------------------------------------------
// Camera (core/src/Camera.cpp:23-57) and Frame (core/src/Frame.cpp:215-275)
// ------------------------------------------------------------------------------------------
struct Camera
{
  double fx, fy, cx, cy;
  // analytic K^-1 (Eigen computes it numerically; equal up to rounding)
  double ifx() const { return 1.0 / fx; }
  double ify() const { return 1.0 / fy; }
  double kx() const { return -cx / fx; }
  double ky() const { return -cy / fy; }
  Camera resized(double s) const { return {fx * s, fy * s, cx * s, cy * s}; }  // Camera.cpp:53-57
};

// Camera::image2camera (Camera.cpp:32-35): Kinv * ((u,v,1)*d).  Operation order fixed here and
// mirrored on the device: X = fma(kx, d, ifx*(u*d)), Y = fma(ky, d, ify*(v*d)), Z = d.
static inline void image2camera(const Camera & c, int u, int v, double d, double p[3])
{
  p[0] = std::fma(c.kx(), d, c.ifx() * ((double)u * d));
  p[1] = std::fma(c.ky(), d, c.ify() * ((double)v * d));
  p[2] = d;
}

struct Frame
{
  int nLevels = 0;
  std::vector<int> W, H;     // intensity sizes per level ((W+1)/2 rule)
  std::vector<int> Wd, Hd;   // depth sizes per level (int(W*0.5) rule) — differ at odd sizes
  std::vector<std::vector<uint8_t>> I;
  std::vector<std::vector<double>> Z;
  std::vector<std::vector<int16_t>> dIx, dIy;
  std::vector<std::vector<double>> pcl;  // 3 doubles per pixel, index v*Wd+u (Frame.cpp:233-253)
  std::vector<Camera> cam;
  long n2_hits = 0;
};

// Frame ctor + computePyramid + computeDerivatives + computePcl in the node's order
// (src/ros/nodes/NodeRgbdAlignment.cpp:194-197)
static Frame * frame_create(
  int W, int H, const uint8_t * gray, const double * depth, double fx, double fy, double cx,
  double cy, int nLevels, int f32_storage)
{
  // Frame.cpp:150-155
  if (std::abs(W / 2 - cx) > 10) return nullptr;
  auto f = new Frame();
  f->nLevels = nLevels;
  f->W.resize(nLevels);
  f->H.resize(nLevels);
  f->Wd.resize(nLevels);
  f->Hd.resize(nLevels);
  f->I.resize(nLevels);
  f->Z.resize(nLevels);
  f->dIx.resize(nLevels);
  f->dIy.resize(nLevels);
  f->pcl.resize(nLevels);
  f->cam.resize(nLevels);
  f->W[0] = f->Wd[0] = W;
  f->H[0] = f->Hd[0] = H;
  f->I[0].assign(gray, gray + (size_t)W * H);
  f->Z[0].assign(depth, depth + (size_t)W * H);
  if (f32_storage)
    for (auto & z : f->Z[0]) z = (double)(float)z;
  const Camera c0{fx, fy, cx, cy};
  // computePyramid (Frame.cpp:255-275)
  for (int l = 0; l < nLevels; l++) {
    if (l > 0) {
      f->W[l] = (f->W[l - 1] + 1) / 2;
      f->H[l] = (f->H[l - 1] + 1) / 2;
      f->I[l].resize((size_t)f->W[l] * f->H[l]);
      pyr_down_u8(f->I[l - 1].data(), f->W[l - 1], f->H[l - 1], f->I[l].data());
    }
    f->cam[l] = c0.resized(std::pow(0.5, l));
  }
  for (int l = 1; l < nLevels; l++) {
    f->Wd[l] = (int)(f->Wd[l - 1] * 0.5);
    f->Hd[l] = (int)(f->Hd[l - 1] * 0.5);
    f->Z[l].resize((size_t)f->Wd[l] * f->Hd[l]);
    depth_down(f->Z[l - 1].data(), f->Wd[l - 1], f->Hd[l - 1], f->Z[l].data(), f32_storage, &f->n2_hits);
  }
  // computeDerivatives (Frame.cpp:215-232)
  for (int l = 0; l < nLevels; l++) {
    const size_t n = (size_t)f->W[l] * f->H[l];
    std::vector<uint8_t> blur(n);
    blur3_u8(f->I[l].data(), f->W[l], f->H[l], blur.data());
    f->dIx[l].resize(n);
    f->dIy[l].resize(n);
    sobel3_s16(blur.data(), f->W[l], f->H[l], f->dIx[l].data(), f->dIy[l].data());
  }
  // computePcl (Frame.cpp:233-253)
  for (int l = 0; l < nLevels; l++) {
    const int w = f->Wd[l], h = f->Hd[l];
    f->pcl[l].assign((size_t)3 * w * h, 0.0);
    for (int v = 0; v < h; v++)
      for (int u = 0; u < w; u++) {
        const double d = f->Z[l][(size_t)v * w + u];
        if (std::isfinite(d) && d > 0.0) image2camera(f->cam[l], u, v, d, &f->pcl[l][3 * ((size_t)v * w + u)]);
      }
  }
  return f;
}

// ------------------------------------------------------------------------------------------
// least_squares: scalers (Scaler.cpp:21-67), losses (Loss.cpp:51-92), normal equations
// ------------------------------------------------------------------------------------------
enum LossKind { LOSS_NONE = 0, LOSS_QUADRATIC = 1, LOSS_TUKEY = 2, LOSS_HUBER = 3, LOSS_TDIST = 4 };
enum ScalerKind { SCALER_IDENTITY = 0, SCALER_MEDIAN = 1, SCALER_MEAN = 2, SCALER_TDIST = 3 };

struct Config
{
  double minGradient;
  int lossKind;
  double huberC;
  double lossV;      // nu of LossTDistribution (Loss.h:103)
  int scalerKind;
  double scalerV;    // nu of ScalerTDistribution (Scaler.h:61-64)
  double minStepSize;
  int maxIterations;
  int includePrior;
  int f32Storage;
  // EXTENSION (no counterpart in the reference): solver 0 = Gauss-Newton (GaussNewton.cpp),
  // 1 = Levenberg-Marquardt damping on the same normal equations
  int solverKind;
  double lmLambda0, lmUp, lmDown;
  // EXTENSION: sampling 0 = nearest pixel, inverse compositional (the reference's aligner); 1 = bilinear sampling with
  // per-iteration Jacobians at the warped point, optionally with a depth residual weighted by depthWeight
  int sampling;
  double depthWeight;
  // chi2 of one view's normal equations.  1: sum_k w_k r_k^2, the r^T W r that NormalEquations.h documents.
  // 0: w_0 * sum_k r_k^2, what NormalEquations.cpp:58 `(r * w).transpose() * r` evaluates to in a Release build: r and w
  // are both column vectors, Eigen only ASSERTS the inner dimension of a product (compiled out by NDEBUG) and takes it
  // from the left factor, whose column count is 1, so (r * w)(k) = r(k) * w(0) - the weight of the FIRST interest point
  // (0 when it is invisible) times the unweighted sum of squares.  Derived from Eigen's product rules, and what
  // oracle/_ref (the reference's sources against oracle/ref_shim) computes; it cannot be measured against real Eigen here.
  int chi2Mode;
};

struct Scale
{
  double offset, scale;
};

static Scale compute_scale(const Config & c, const std::vector<double> & r)
{
  const size_t n = r.size();
  switch (c.scalerKind) {
    case SCALER_TDIST: {
      // Scaler.cpp:49-67; defaults 30 iterations, 1e-5 (Scaler.h:61-64)
      double step = std::numeric_limits<double>::max(), sigma = 1.0;
      for (int it = 0; it < 30 && step > 1e-5; it++) {
        double sum = 0.0;
        for (size_t i = 0; i < n; i++) {
          const double q = r[i] / sigma;
          sum += r[i] * r[i] * (c.scalerV + 1) / (c.scalerV + q * q);
        }
        const double s = std::sqrt(sum / (double)n);
        step = std::fabs(sigma - s);
        sigma = s;
      }
      return {0.0, sigma};
    }
    case SCALER_MEAN: {
      if (n == 0) return {0.0, 1.0};
      double s = 0;
      for (double v : r) s += v;
      const double mean = s / (double)n;
      double a = 0;
      for (double v : r) a += std::fabs(v - mean);
      return {mean, std::sqrt(a / (double)(n - 1))};
    }
    case SCALER_MEDIAN: {
      std::vector<double> v(r);
      std::sort(v.begin(), v.end());
      double med;
      const long m = (long)n;
      if (m == 0) return {0.0, 1.0};  // reference: UB on empty input
      if (m % 2 == 0)
        med = m == 2 ? v[0] : (v[m / 2 - 1] + v[m / 2 + 1]) / 2;
      else
        med = v[m / 2];
      double a = 0;
      for (double x : r) a += std::fabs(x - med);
      return {med, std::sqrt(a / (double)(m - 1))};
    }
    default:
      return {0.0, 1.0};
  }
}

static inline double compute_weight(const Config & c, double r)
{
  switch (c.lossKind) {
    case LOSS_TUKEY: {
      const double C = 4.6851;
      if (std::fabs(r) < C) {
        const double rc = r / C;
        const double q = 1.0 - rc * rc;
        return q * q;
      }
      return 0.0;
    }
    case LOSS_HUBER:
      if (std::fabs(r) < c.huberC) return 1.0;
      return (c.huberC * r > 0.0 ? 1.0 : -1.0) / r;
    case LOSS_TDIST:
      return (c.lossV + 1.0) / (c.lossV + r * r);
    default:
      return 1.0;
  }
}

struct NE
{
  Mat6 A;
  Vec6 b;
  double chi2 = 0;
  long n = 0;
  double scale = 1, offset = 0;
  NE()
  {
    std::memset(A.a, 0, sizeof(A.a));
    std::memset(b.v, 0, sizeof(b.v));
  }
};

// ------------------------------------------------------------------------------------------
// WarpSE3 + InverseCompositional for one (keyframe -> target) view
// (lukas_kanade/src/Warp.cpp:135-201,233-239; InverseCompositional.cpp:63-93,104-154)
// ------------------------------------------------------------------------------------------
static const int16_t R_NOT_INTEREST = -32768;
static const int16_t R_INVISIBLE = -32767;

struct View
{
  const Frame * from;
  const Frame * to;
  int level;
  int W, H;          // from->width(level), height(level) (intensity dims)
  Camera camProj;    // _camCur  = from->camera(level): used by apply()  (argument swap)
  Camera camJac;     // _camRef  = to->camera(level):   scales J
  SE3 se3;           // poseCur * poseRef^-1
  SE3 poseRef;
  std::vector<int> ipU, ipV;   // retained interest points, row-major scan order
  std::vector<double> J;       // N x 6 row-major
  bool hasPrior = false;
  Vec6 xPred;
  // scratch of the last computeNormalEquations
  std::vector<double> r, w;
};

// Warp::apply(u,v) (Warp.cpp:158-165) + Camera::camera2image (Camera.cpp:23-30).
// q = R p + t evaluated as nested fma (innermost first: column 0, then 1, then 2);
// u' = fma(fx, qx, cx*qz) / qz.  Mirrored bit for bit on the device.
static inline bool warp_apply(
  const double R[9], const double t[3], const Camera & cam, const double p[3], double & up, double & vp)
{
  if (!(p[2] > 0.0)) return false;
  const double qx = std::fma(R[2], p[2], std::fma(R[1], p[1], std::fma(R[0], p[0], t[0])));
  const double qy = std::fma(R[5], p[2], std::fma(R[4], p[1], std::fma(R[3], p[0], t[1])));
  const double qz = std::fma(R[8], p[2], std::fma(R[7], p[1], std::fma(R[6], p[0], t[2])));
  if (qz <= 0.0) return false;
  up = std::fma(cam.fx, qx, cam.cx * qz) / qz;
  vp = std::fma(cam.fy, qy, cam.cy * qz) / qz;
  return true;
}

static void view_init(
  View & V, const Config & cfg, const Frame * from, const Frame * to, int level, const SE3 & poseCur,
  const SE3 & poseRef, bool multiFrameMask)
{
  V.from = from;
  V.to = to;
  V.level = level;
  V.W = from->W[level];
  V.H = from->H[level];
  V.camProj = from->cam[level];
  V.camJac = to->cam[level];
  V.poseRef = poseRef;
  V.se3 = se3_mul(poseCur, se3_inv(poseRef));
  const int W = V.W, H = V.H;
  const double minGrad2 = cfg.minGradient * cfg.minGradient;
  const double s = 1.0 / std::pow(0.5, level);
  const int16_t * gx = from->dIx[level].data();
  const int16_t * gy = from->dIy[level].data();
  const double * pcl = from->pcl[level].data();
  V.ipU.clear();
  V.ipV.clear();
  V.J.clear();
  V.ipU.reserve((size_t)W * H);
  V.ipV.reserve((size_t)W * H);
  V.J.reserve((size_t)W * H * 6);
  const double fxJ = V.camJac.fx, fyJ = V.camJac.fy;
  for (int v = 0; v < H; v++) {
    for (int u = 0; u < W; u++) {
      const size_t i = (size_t)v * W + u;
      const double dx = gx[i], dy = gy[i];
      const double g2 = dx * dx + dy * dy;
      if (!(g2 >= minGrad2)) continue;
      // SE3Alignment.cpp:87-94 (single) / :127 (multi)
      if (multiFrameMask) {
        if (!(from->Z[level][i] > 0.0)) continue;
      } else {
        if (!(from->Z[0][(size_t)(int)(v * s) * from->Wd[0] + (int)(u * s)] > 0.0)) continue;
      }
      // WarpSE3::J (Warp.cpp:166-201): NaN row if z <= 0 -> dropped (InverseCompositional.cpp:82)
      const double * p = pcl + 3 * i;
      if (p[2] <= 0.0) continue;
      const double x = p[0], y = p[1], zi = 1. / p[2], zi2 = zi * zi;
      double j0[6], j1[6];
      j0[0] = zi;
      j0[1] = 0.0;
      j0[2] = -x * zi2;
      j0[3] = y * j0[2];
      j0[4] = 1.0 - x * j0[2];
      j0[5] = -y * zi;
      j1[0] = 0.0;
      j1[1] = zi;
      j1[2] = -y * zi2;
      j1[3] = -1.0 + y * j1[2];
      j1[4] = -j0[3];
      j1[5] = x * zi;
      double Ji[6], nrm = 0;
      for (int k = 0; k < 6; k++) {
        Ji[k] = (j0[k] * fxJ) * dx + (j1[k] * fyJ) * dy;
        nrm += Ji[k] * Ji[k];
      }
      if (!std::isfinite(std::sqrt(nrm))) continue;
      V.ipU.push_back(u);
      V.ipV.push_back(v);
      V.J.insert(V.J.end(), Ji, Ji + 6);
    }
  }
}

// InverseCompositional::computeNormalEquations (InverseCompositional.cpp:105-154) without prior.
// rcode (optional, W*H int16): R_NOT_INTEREST / R_INVISIBLE / residual.
static NE view_normal_equations(View & V, const Config & cfg, int16_t * rcode)
{
  const size_t N = V.ipU.size();
  V.r.assign(N, 0.0);
  V.w.assign(N, 0.0);
  double R[9];
  quat_to_R(V.se3.q, R);
  const int W = V.W;
  const int Wt = V.to->W[V.level], Ht = V.to->H[V.level];
  const uint8_t * I = V.to->I[V.level].data();
  const uint8_t * T = V.from->I[V.level].data();
  const double * pcl = V.from->pcl[V.level].data();
  if (rcode)
    for (size_t i = 0; i < (size_t)V.W * V.H; i++) rcode[i] = R_NOT_INTEREST;
  for (size_t k = 0; k < N; k++) {
    const int u = V.ipU[k], v = V.ipV[k];
    double up, vp;
    const bool ok = warp_apply(R, V.se3.t, V.camProj, pcl + 3 * ((size_t)v * W + u), up, vp);
    const bool visible = ok && 1 < up && up < Wt - 1 && 1 < vp && vp < Ht - 1 && std::isfinite(up);
    if (visible) {
      const int iu = (int)std::round(up), iv = (int)std::round(vp);
      V.r[k] = (double)I[(size_t)iv * Wt + iu] - (double)T[(size_t)v * W + u];
      V.w[k] = 1.0;
      if (rcode) rcode[(size_t)v * W + u] = (int16_t)V.r[k];
    } else if (rcode) {
      rcode[(size_t)v * W + u] = R_INVISIBLE;
    }
  }
  NE ne;
  if (cfg.lossKind != LOSS_NONE) {
    const Scale s = compute_scale(cfg, V.r);
    ne.scale = s.scale;
    ne.offset = s.offset;
    for (size_t k = 0; k < N; k++)
      if (V.w[k] > 0.0) V.w[k] = compute_weight(cfg, (V.r[k] - s.offset) / s.scale);
  }
  // NormalEquations(J, r, w) (NormalEquations.cpp:53-61)
  double A[21] = {0}, b[6] = {0}, chi2 = 0;
  for (size_t k = 0; k < N; k++) {
    const double wk = V.w[k];
    if (wk == 0.0) continue;  // exact: contributes +0
    const double * Jk = &V.J[6 * k];
    const double rk = V.r[k];
    int idx = 0;
    for (int i = 0; i < 6; i++) {
      const double wj = wk * Jk[i];
      for (int j = i; j < 6; j++) A[idx++] += wj * Jk[j];
      b[i] += wj * rk;
    }
    chi2 += rk * wk * rk;
  }
  if (cfg.chi2Mode == 0) {
    chi2 = 0;
    const double w0 = N ? V.w[0] : 0.0;
    for (size_t k = 0; k < N; k++) chi2 += (V.r[k] * w0) * V.r[k];
  }
  int idx = 0;
  for (int i = 0; i < 6; i++)
    for (int j = i; j < 6; j++) {
      ne.A(i, j) = A[idx];
      ne.A(j, i) = A[idx++];
    }
  for (int i = 0; i < 6; i++) ne.b[i] = b[i];
  ne.chi2 = chi2;
  ne.n = (long)N;
  if (ne.n > 1) {  // InverseCompositional.cpp:139-143
    const double inv = (double)ne.n;
    for (double & a : ne.A.a) a /= inv;
    for (double & x : ne.b.v) x /= inv;
    ne.chi2 /= inv;
  }
  return ne;
}

// ------------------------------------------------------------------------------------------
// EXTENSION (named by BASELINE.json's north_star, absent from the reference's aligner): forward
// formulation with BILINEAR sampling, per-iteration Jacobians and an optional DEPTH residual.  Semantics follow
// lukas_kanade/src/ForwardAdditive.cpp:52-137 (gradients of the target sampled at the warped position, visibility
// 1 < u' < W-1, scale over the visible residuals) and the dense RGB-D odometry literature odometry/README.md:3-6
// cites (Kerl et al.: geometric residual Z_to(u', v') - q_z).  This restatement is the definition: parity unpinned.
//   q = R p + t, (u', v') = pi(q), S(f) = true bilinear interpolation of plane f at (u', v')
//   r_I = S(I_to) - T(u, v);   g = (S(dIx_to), S(dIy_to)) / 8   (the Sobel planes are 8 x the derivative)
//   right perturbation se3 * exp(xi) (the solvers update se3 <- se3 * exp(-dx)):  dq/dxi = R [I | -p^]
//   J_I = G R, (p x (G R)) with G = g^T dpi/dq = (g_x fx / q_z, g_y fy / q_z, -(g_x fx q_x + g_y fy q_y) / q_z^2)
//   r_Z = S(Z_to) - q_z (all four neighbours valid), gradient of the bilinear interpolant itself,
//   J_Z = G_Z R, (p x (G_Z R)),  G_Z = g_Z^T dpi/dq - (0, 0, 1)
//   A = (sum w_I J_I^T J_I + depthWeight sum w_Z J_Z^T J_Z) / n, b and chi2 alike, n = all interest points
// ------------------------------------------------------------------------------------------
static inline double bilinear_true(const double f00, const double f10, const double f01, const double f11, double a, double b)
{
  return (1.0 - b) * ((1.0 - a) * f00 + a * f10) + b * ((1.0 - a) * f01 + a * f11);
}

static NE view_normal_equations_ext(View & V, const Config & cfg)
{
  const size_t N = V.ipU.size();
  double R[9];
  quat_to_R(V.se3.q, R);
  const int W = V.W, l = V.level;
  const int Wt = V.to->W[l], Ht = V.to->H[l];
  const uint8_t * I = V.to->I[l].data();
  const uint8_t * T = V.from->I[l].data();
  const int16_t * gxT = V.to->dIx[l].data();
  const int16_t * gyT = V.to->dIy[l].data();
  const double * Zt = V.to->Z[l].data();
  const double * pcl = V.from->pcl[l].data();
  const double fx = V.camProj.fx, fy = V.camProj.fy;
  const double NaN = std::numeric_limits<double>::quiet_NaN();
  std::vector<double> rI(N, NaN), rZ(N, NaN), JI(6 * N, 0.0), JZ(6 * N, 0.0);
  std::vector<double> visI, visZ;
  for (size_t k = 0; k < N; k++) {
    const int u = V.ipU[k], v = V.ipV[k];
    const double * p = pcl + 3 * ((size_t)v * W + u);
    double up, vp;
    if (!warp_apply(R, V.se3.t, V.camProj, p, up, vp)) continue;
    if (!(1 < up && up < Wt - 1 && 1 < vp && vp < Ht - 1)) continue;
    const double qx = std::fma(R[2], p[2], std::fma(R[1], p[1], std::fma(R[0], p[0], V.se3.t[0])));
    const double qy = std::fma(R[5], p[2], std::fma(R[4], p[1], std::fma(R[3], p[0], V.se3.t[1])));
    const double qz = std::fma(R[8], p[2], std::fma(R[7], p[1], std::fma(R[6], p[0], V.se3.t[2])));
    const int x0 = (int)std::floor(up), y0 = (int)std::floor(vp);
    const double a = up - x0, b = vp - y0;
    const size_t i00 = (size_t)y0 * Wt + x0, i10 = i00 + 1, i01 = i00 + Wt, i11 = i01 + 1;
    rI[k] = bilinear_true(I[i00], I[i10], I[i01], I[i11], a, b) - (double)T[(size_t)v * W + u];
    visI.push_back(rI[k]);
    const double gx = bilinear_true(gxT[i00], gxT[i10], gxT[i01], gxT[i11], a, b) / 8.0;
    const double gy = bilinear_true(gyT[i00], gyT[i10], gyT[i01], gyT[i11], a, b) / 8.0;
    const double iz = 1.0 / qz;
    auto jac = [&](double g0, double g1, double g2, double * J) {
      const double H[3] = {g0 * R[0] + g1 * R[3] + g2 * R[6], g0 * R[1] + g1 * R[4] + g2 * R[7], g0 * R[2] + g1 * R[5] + g2 * R[8]};
      J[0] = H[0];
      J[1] = H[1];
      J[2] = H[2];
      J[3] = p[1] * H[2] - p[2] * H[1];
      J[4] = p[2] * H[0] - p[0] * H[2];
      J[5] = p[0] * H[1] - p[1] * H[0];
    };
    jac(gx * fx * iz, gy * fy * iz, -(gx * fx * qx + gy * fy * qy) * iz * iz, &JI[6 * k]);
    if (cfg.depthWeight > 0.0) {
      const double z00 = Zt[i00], z10 = Zt[i10], z01 = Zt[i01], z11 = Zt[i11];
      if (z00 > 0 && z10 > 0 && z01 > 0 && z11 > 0 && std::isfinite(z00 + z10 + z01 + z11)) {
        rZ[k] = bilinear_true(z00, z10, z01, z11, a, b) - qz;
        visZ.push_back(rZ[k]);
        const double gzx = (1.0 - b) * (z10 - z00) + b * (z11 - z01), gzy = (1.0 - a) * (z01 - z00) + a * (z11 - z10);
        jac(gzx * fx * iz, gzy * fy * iz, -(gzx * fx * qx + gzy * fy * qy) * iz * iz - 1.0, &JZ[6 * k]);
      }
    }
  }
  NE ne;
  Scale sI{0.0, 1.0}, sZ{0.0, 1.0};
  if (cfg.lossKind != LOSS_NONE) {
    sI = compute_scale(cfg, visI);
    if (!visZ.empty()) sZ = compute_scale(cfg, visZ);
  }
  ne.scale = sI.scale;
  ne.offset = sI.offset;
  double A[21] = {0}, b[6] = {0}, chi2 = 0;
  auto add = [&](const double * Jk, double rk, double wk) {
    int idx = 0;
    for (int i = 0; i < 6; i++) {
      const double wj = wk * Jk[i];
      for (int j = i; j < 6; j++) A[idx++] += wj * Jk[j];
      b[i] += wj * rk;
    }
    chi2 += rk * wk * rk;
  };
  for (size_t k = 0; k < N; k++) {
    if (!std::isnan(rI[k]))
      add(&JI[6 * k], rI[k], cfg.lossKind == LOSS_NONE ? 1.0 : compute_weight(cfg, (rI[k] - sI.offset) / sI.scale));
    if (!std::isnan(rZ[k]))
      add(&JZ[6 * k], rZ[k],
          cfg.depthWeight * (cfg.lossKind == LOSS_NONE ? 1.0 : compute_weight(cfg, (rZ[k] - sZ.offset) / sZ.scale)));
  }
  int idx = 0;
  for (int i = 0; i < 6; i++)
    for (int j = i; j < 6; j++) {
      ne.A(i, j) = A[idx];
      ne.A(j, i) = A[idx++];
    }
  for (int i = 0; i < 6; i++) ne.b[i] = b[i];
  ne.chi2 = chi2;
  ne.n = (long)N;
  if (ne.n > 1) {
    const double inv = (double)ne.n;
    for (double & x : ne.A.a) x /= inv;
    for (double & x : ne.b.v) x /= inv;
    ne.chi2 /= inv;
  }
  return ne;
}

// MotionPrior::apply (odometry/src/direct_image_alignment/SE3Alignment.cpp:37-47)
static void prior_apply(NE & ne, const Vec6 & xPred, const Vec6 & x)
{
  const double normalizer = 1.0 / (255.0 * 255.0);
  for (double & a : ne.A.a) a *= normalizer;
  for (double & v : ne.b.v) v *= normalizer;
  for (int i = 0; i < 6; i++) {
    ne.A(i, i) += 1.0;
    ne.b[i] += xPred[i] - x[i];
  }
}

// ------------------------------------------------------------------------------------------
// GaussNewton::solve (least_squares/src/GaussNewton.cpp:33-102) over a stacked problem of K views
// (lukas_kanade/src/InverseCompositionalStacked.cpp:25-48; K == 1 is the plain problem)
// ------------------------------------------------------------------------------------------
enum Stop { STOP_MAXITER = 0, STOP_CONSTRAINTS = 1, STOP_HESSIAN = 2, STOP_CHI2 = 3, STOP_CONVERGED = 4, STOP_NAN = 5 };
static const int TRACE_STRIDE = 64;
// trace record: 0 valid, 1 n, 2 chi2, 3 stepSize, 4 scale(view0), 5 offset(view0), 6..11 x after update,
// 12..17 b, 18..53 A (row major), 54..59 dx

struct SolveResult
{
  int iteration = -1;  // index of the last pushed NE (Solver.h:59; -1 = never set)
  int nPushed = 0;
  int stop = STOP_MAXITER;
  NE last;
};

static NE stacked_normal_equations(std::vector<View> & views, const Config & cfg, bool stacked)
{
  if (!stacked) {
    NE ne = cfg.sampling == 1 ? view_normal_equations_ext(views[0], cfg) : view_normal_equations(views[0], cfg, nullptr);
    if (views[0].hasPrior) prior_apply(ne, views[0].xPred, se3_log(views[0].se3));
    return ne;
  }
  NE sum;  // NormalEquations(nParameters) zero, then combine (InverseCompositionalStacked.cpp:44-46)
  bool first = true;
  for (auto & V : views) {
    NE ne = cfg.sampling == 1 ? view_normal_equations_ext(V, cfg) : view_normal_equations(V, cfg, nullptr);
    if (V.hasPrior) prior_apply(ne, V.xPred, se3_log(V.se3));
    for (int i = 0; i < 36; i++) sum.A.a[i] += ne.A.a[i];
    for (int i = 0; i < 6; i++) sum.b[i] += ne.b[i];
    sum.chi2 += ne.chi2;
    sum.n += ne.n;
    if (first) {
      sum.scale = ne.scale;
      sum.offset = ne.offset;
      first = false;
    }
  }
  return sum;
}

static SolveResult gauss_newton(
  std::vector<View> & views, const Config & cfg, bool stacked, double * trace /* maxIterations*64 or null */)
{
  SolveResult res;
  const int maxIt = cfg.maxIterations;
  std::vector<double> chi2(maxIt, 0.0), step(maxIt, 0.0);
  std::vector<Vec6> xs(maxIt);
  for (int i = 0; i < maxIt; i++) {
    NE ne = stacked_normal_equations(views, cfg, stacked);
    const double det = det6(ne.A);
    if (ne.n < 6) {
      res.stop = STOP_CONSTRAINTS;
      break;
    }
    if (!std::isfinite(det) || std::fabs(det) < 1e-6) {
      res.stop = STOP_HESSIAN;
      break;
    }
    chi2[i] = ne.chi2;
    const double dChi2 = i > 0 ? chi2[i] - chi2[i - 1] : 0;
    if (i > 0 && dChi2 > 0) {
      // problem->setX(x.row(i-1)): se3 = exp(x) for every view (Warp.cpp:233-237,
      // InverseCompositionalStacked.cpp:35-38)
      for (auto & V : views) V.se3 = se3_exp(xs[i - 1]);
      res.stop = STOP_CHI2;
      break;
    }
    const Vec6 dx = ldlt_solve(ne.A, ne.b);
    // updateX(dx) -> updateCompositional(-dx): se3 = se3 * exp(-dx) (InverseCompositional.cpp:104, Warp.cpp:153-157)
    Vec6 mdx;
    for (int k = 0; k < 6; k++) mdx[k] = -dx[k];
    const SE3 inc = se3_exp(mdx);
    for (auto & V : views) V.se3 = se3_mul(V.se3, inc);
    xs[i] = se3_log(views[0].se3);
    double nrm = 0;
    for (int k = 0; k < 6; k++) nrm += dx[k] * dx[k];
    step[i] = std::sqrt(nrm);
    res.last = ne;
    res.iteration = i;
    res.nPushed++;
    if (trace) {
      double * t = trace + (size_t)i * TRACE_STRIDE;
      t[0] = 1;
      t[1] = (double)ne.n;
      t[2] = ne.chi2;
      t[3] = step[i];
      t[4] = ne.scale;
      t[5] = ne.offset;
      for (int k = 0; k < 6; k++) t[6 + k] = xs[i][k];
      for (int k = 0; k < 6; k++) t[12 + k] = ne.b[k];
      for (int k = 0; k < 36; k++) t[18 + k] = ne.A.a[k];
      for (int k = 0; k < 6; k++) t[54 + k] = dx[k];
    }
    double bmax = ne.b[0];
    for (int k = 1; k < 6; k++) bmax = std::max(bmax, ne.b[k]);
    if (i > 0 && (step[i] < cfg.minStepSize || std::fabs(bmax) < cfg.minStepSize || std::fabs(dChi2) < cfg.minStepSize)) {
      res.stop = STOP_CONVERGED;
      break;
    }
    if (!std::isfinite(step[i])) {
      if (i > 0)
        for (auto & V : views) V.se3 = se3_exp(xs[i - 1]);  // i == 0: row(-1) is UB in the reference
      res.stop = STOP_NAN;
      break;
    }
  }
  return res;
}

// ------------------------------------------------------------------------------------------
// EXTENSION: Levenberg-Marquardt on the same problem (named by BASELINE.json's configs[3]; the reference has no LM
// solver, so this restatement IS the definition - parity unpinned).  Per evaluation i:
//   ne = normal equations at the current pose (normalised, prior-augmented, exactly as for Gauss-Newton)
//   worse than the accepted state (chi2 > chi2_acc):  reject - pose <- accepted pose, lambda *= up (stop at 1e12)
//   else: accept - (A, b, chi2, pose)_acc <- current, lambda <- max(lambda * down, 1e-12) from the second accept on;
//         converged when |chi2 - chi2_acc_before| < minStepSize
//   dx = ldlt(A_acc + lambda diag(A_acc))^-1 b_acc;  se3 <- se3_acc * exp(-dx);  converged when |dx| < minStepSize
// The covariance is the inverse of the last accepted, undamped A (as SE3Alignment.cpp:101 does for Gauss-Newton).
// trace slot 60 holds lambda.
// ------------------------------------------------------------------------------------------
static SolveResult levenberg_marquardt(std::vector<View> & views, const Config & cfg, bool stacked, double * trace)
{
  SolveResult res;
  double lambda = cfg.lmLambda0, chi2Acc = 0.0;
  bool accepted = false;
  Mat6 Aacc;
  Vec6 bacc;
  NE neAcc;
  std::vector<SE3> se3Acc(views.size());
  for (int i = 0; i < cfg.maxIterations; i++) {
    NE ne = stacked_normal_equations(views, cfg, stacked);
    if (ne.n < 6) {
      res.stop = STOP_CONSTRAINTS;
      break;
    }
    if (accepted && ne.chi2 > chi2Acc) {
      for (size_t k = 0; k < views.size(); k++) views[k].se3 = se3Acc[k];
      lambda *= cfg.lmUp;
      if (lambda > 1e12) {
        res.stop = STOP_CONVERGED;
        break;
      }
    } else {
      const bool converged = accepted && std::fabs(ne.chi2 - chi2Acc) < cfg.minStepSize;
      if (accepted) lambda = std::max(lambda * cfg.lmDown, 1e-12);
      Aacc = ne.A;
      bacc = ne.b;
      chi2Acc = ne.chi2;
      neAcc = ne;
      for (size_t k = 0; k < views.size(); k++) se3Acc[k] = views[k].se3;
      accepted = true;
      res.last = ne;
      if (converged) {
        res.iteration = std::max(res.iteration, 0);
        res.stop = STOP_CONVERGED;
        break;
      }
    }
    Mat6 D = Aacc;
    for (int k = 0; k < 6; k++) D(k, k) += lambda * Aacc(k, k);
    const double det = det6(D);
    if (!std::isfinite(det) || std::fabs(det) < 1e-6) {
      res.stop = STOP_HESSIAN;
      break;
    }
    const Vec6 dx = ldlt_solve(D, bacc);
    Vec6 mdx;
    for (int k = 0; k < 6; k++) mdx[k] = -dx[k];
    const SE3 inc = se3_exp(mdx);
    for (size_t k = 0; k < views.size(); k++) views[k].se3 = se3_mul(se3Acc[k], inc);
    double nrm = 0;
    for (int k = 0; k < 6; k++) nrm += dx[k] * dx[k];
    const double step = std::sqrt(nrm);
    res.iteration = i;
    res.nPushed++;
    if (trace) {
      double * t = trace + (size_t)i * TRACE_STRIDE;
      const Vec6 x = se3_log(views[0].se3);
      t[0] = 1;
      t[1] = (double)neAcc.n;
      t[2] = chi2Acc;
      t[3] = step;
      t[4] = neAcc.scale;
      t[5] = neAcc.offset;
      for (int k = 0; k < 6; k++) t[6 + k] = x[k];
      for (int k = 0; k < 6; k++) t[12 + k] = bacc[k];
      for (int k = 0; k < 36; k++) t[18 + k] = Aacc.a[k];
      for (int k = 0; k < 6; k++) t[54 + k] = dx[k];
      t[60] = lambda;
    }
    if (!std::isfinite(step)) {
      for (size_t k = 0; k < views.size(); k++) views[k].se3 = se3Acc[k];
      res.stop = STOP_NAN;
      break;
    }
    if (step < cfg.minStepSize) {
      res.stop = STOP_CONVERGED;
      break;
    }
  }
  return res;
}

// ------------------------------------------------------------------------------------------
// SE3Alignment::align, single (SE3Alignment.cpp:64-105) and multi keyframe (:106-146)
// ------------------------------------------------------------------------------------------
struct AlignOut
{
  SE3 pose;
  Mat6 cov;
  int status = 0;          // 0 ok, 1 = a level finished without a single completed iteration (reference: UB)
  int iters[16] = {0};     // pushed iterations per level
  int stop[16] = {0};
};

static AlignOut align(
  const Config & cfg, const std::vector<const Frame *> & from, const std::vector<SE3> & fromPose,
  const Frame * to, const SE3 & toPose, bool multi, double * trace /* nLevels*maxIt*64 */)
{
  AlignOut out;
  out.pose = toPose;
  for (int i = 0; i < 36; i++) out.cov.a[i] = (i % 7 == 0) ? 1.0 : 0.0;
  const int nLevels = from[0]->nLevels;
  const size_t K = multi ? from.size() : 1;
  std::vector<Vec6> xPred(K);
  for (size_t k = 0; k < K; k++) xPred[k] = se3_log(se3_mul(toPose, se3_inv(fromPose[k])));
  for (int level = nLevels - 1; level >= 0; level--) {
    std::vector<View> views(K);
    for (size_t k = 0; k < K; k++) {
      view_init(views[k], cfg, from[k], to, level, out.pose, fromPose[k], multi);
      views[k].hasPrior = cfg.includePrior != 0;
      views[k].xPred = xPred[k];
    }
    double * tr = trace ? trace + (size_t)level * cfg.maxIterations * TRACE_STRIDE : nullptr;
    SolveResult r = cfg.solverKind == 1 ? levenberg_marquardt(views, cfg, multi, tr) : gauss_newton(views, cfg, multi, tr);
    out.iters[level] = r.nPushed;
    out.stop[level] = r.stop;
    if (r.iteration < 0) {
      out.status = 1;
      out.pose = toPose;
      for (int i = 0; i < 36; i++) out.cov.a[i] = (i % 7 == 0) ? 1.0 : 0.0;
      return out;
    }
    out.cov = inverse6(r.last.A);
    out.pose = se3_mul(views[0].se3, views[0].poseRef);  // WarpSE3::poseCur (Warp.cpp:239)
  }
  return out;
}

}  // namespace vo

// ------------------------------------------------------------------------------------------
// C entry points (ctypes; see oracle/oracle.py)
// ------------------------------------------------------------------------------------------
extern "C" {

struct vo_config
{
  double min_gradient;
  int loss_kind;
  double huber_c;
  double loss_v;
  int scaler_kind;
  double scaler_v;
  double min_step_size;
  int max_iterations;
  int include_prior;
  int f32_storage;
  int solver_kind;
  double lm_lambda0, lm_up, lm_down;
  int sampling;
  double depth_weight;
  int chi2_mode;
};

static vo::Config to_cfg(const vo_config * c)
{
  return {c->min_gradient, c->loss_kind, c->huber_c, c->loss_v, c->scaler_kind, c->scaler_v,
          c->min_step_size, c->max_iterations, c->include_prior, c->f32_storage, c->solver_kind, c->lm_lambda0, c->lm_up,
          c->lm_down, c->sampling, c->depth_weight, c->chi2_mode};
}
static vo::SE3 to_se3(const double * p)
{
  vo::SE3 s;
  for (int i = 0; i < 4; i++) s.q[i] = p[i];
  for (int i = 0; i < 3; i++) s.t[i] = p[4 + i];
  return s;
}

void vo_pyr_down_u8(const uint8_t * src, int W, int H, uint8_t * dst) { vo::pyr_down_u8(src, W, H, dst); }
void vo_rgb_to_gray_u8(const uint8_t * rgb, long n, int bgr, uint8_t * gray) { vo::rgb_to_gray_u8(rgb, (size_t)n, bgr, gray); }
void vo_depth_ingest(const double * in, long n, double divisor, double * out) { vo::depth_ingest(in, (size_t)n, divisor, out); }
void vo_blur3_u8(const uint8_t * src, int W, int H, uint8_t * dst) { vo::blur3_u8(src, W, H, dst); }
void vo_sobel3_s16(const uint8_t * g, int W, int H, int16_t * dx, int16_t * dy) { vo::sobel3_s16(g, W, H, dx, dy); }
long vo_depth_down(const double * src, int W, int H, double * dst, int f32_storage)
{
  long n2 = 0;
  vo::depth_down(src, W, H, dst, f32_storage, &n2);
  return n2;
}
double vo_bilinear_f64(const double * m, int W, double x, double y) { return vo::bilinear_ref(m, W, x, y); }
double vo_bilinear_u8(const uint8_t * m, int W, double x, double y) { return vo::bilinear_ref(m, W, x, y); }
double vo_median(double * v, int n)
{
  std::vector<double> x(v, v + n);
  return vo::median_ref(x, nullptr);
}
void vo_se3_exp(const double * xi, double * pose7)
{
  vo::Vec6 x;
  for (int i = 0; i < 6; i++) x[i] = xi[i];
  vo::SE3 s = vo::se3_exp(x);
  for (int i = 0; i < 4; i++) pose7[i] = s.q[i];
  for (int i = 0; i < 3; i++) pose7[4 + i] = s.t[i];
}
void vo_se3_log(const double * pose7, double * xi)
{
  vo::Vec6 x = vo::se3_log(to_se3(pose7));
  for (int i = 0; i < 6; i++) xi[i] = x[i];
}
void vo_se3_mul(const double * a, const double * b, double * o)
{
  vo::SE3 s = vo::se3_mul(to_se3(a), to_se3(b));
  for (int i = 0; i < 4; i++) o[i] = s.q[i];
  for (int i = 0; i < 3; i++) o[4 + i] = s.t[i];
}
void vo_se3_inv(const double * a, double * o)
{
  vo::SE3 s = vo::se3_inv(to_se3(a));
  for (int i = 0; i < 4; i++) o[i] = s.q[i];
  for (int i = 0; i < 3; i++) o[4 + i] = s.t[i];
}
void vo_quat_to_R(const double * q, double * R) { vo::quat_to_R(q, R); }
// Camera::camera2image / image2camera (core/src/Camera.cpp:23-35)
int vo_camera2image(double fx, double fy, double cx, double cy, const double * p, double * uv)
{
  const double I[9] = {1, 0, 0, 0, 1, 0, 0, 0, 1}, z[3] = {0, 0, 0};
  vo::Camera c{fx, fy, cx, cy};
  // camera2image alone has no "p.z > 0" pre-test on the source point, only on the projected one
  if (p[2] <= 0) {
    uv[0] = uv[1] = std::numeric_limits<double>::quiet_NaN();
    return 0;
  }
  return vo::warp_apply(I, z, c, p, uv[0], uv[1]) ? 1 : 0;
}
void vo_image2camera(double fx, double fy, double cx, double cy, int u, int v, double d, double * p)
{
  vo::Camera c{fx, fy, cx, cy};
  vo::image2camera(c, u, v, d, p);
}

void * vo_frame_create(
  int W, int H, const uint8_t * gray, const double * depth, double fx, double fy, double cx, double cy,
  int nLevels, int f32_storage)
{
  return vo::frame_create(W, H, gray, depth, fx, fy, cx, cy, nLevels, f32_storage);
}
void vo_frame_destroy(void * f) { delete (vo::Frame *)f; }
long vo_frame_n2_hits(void * f) { return ((vo::Frame *)f)->n2_hits; }
void vo_frame_dims(void * fp, int level, int * out4)
{
  auto f = (vo::Frame *)fp;
  out4[0] = f->W[level];
  out4[1] = f->H[level];
  out4[2] = f->Wd[level];
  out4[3] = f->Hd[level];
}
void vo_frame_get_level(void * fp, int level, uint8_t * I, double * Z, int16_t * dx, int16_t * dy, double * pcl, double * cam4)
{
  auto f = (vo::Frame *)fp;
  if (I) std::memcpy(I, f->I[level].data(), f->I[level].size());
  if (Z) std::memcpy(Z, f->Z[level].data(), f->Z[level].size() * sizeof(double));
  if (dx) std::memcpy(dx, f->dIx[level].data(), f->dIx[level].size() * sizeof(int16_t));
  if (dy) std::memcpy(dy, f->dIy[level].data(), f->dIy[level].size() * sizeof(int16_t));
  if (pcl) std::memcpy(pcl, f->pcl[level].data(), f->pcl[level].size() * sizeof(double));
  if (cam4) {
    cam4[0] = f->cam[level].fx;
    cam4[1] = f->cam[level].fy;
    cam4[2] = f->cam[level].cx;
    cam4[3] = f->cam[level].cy;
  }
}

// One evaluation of InverseCompositional::computeNormalEquations at a given relative transform
// se3 (pose7) on one level.  out: rcode plane (W*H int16), ne64 = [n, chi2, scale, offset, b[6], A[36]]
// (normalised, no prior), J (optional, n x 6), ip count returned.
long vo_evaluate(
  const vo_config * c, void * from, void * to, int level, const double * se3_pose7, int multi_mask,
  int16_t * rcode, double * ne64, double * Jout)
{
  vo::Config cfg = to_cfg(c);
  vo::View V;
  vo::SE3 I;
  vo::view_init(V, cfg, (vo::Frame *)from, (vo::Frame *)to, level, to_se3(se3_pose7), I, multi_mask != 0);
  V.se3 = to_se3(se3_pose7);
  vo::NE ne = vo::view_normal_equations(V, cfg, rcode);
  if (ne64) {
    ne64[0] = (double)ne.n;
    ne64[1] = ne.chi2;
    ne64[2] = ne.scale;
    ne64[3] = ne.offset;
    for (int i = 0; i < 6; i++) ne64[4 + i] = ne.b[i];
    for (int i = 0; i < 36; i++) ne64[10 + i] = ne.A.a[i];
  }
  if (Jout) std::memcpy(Jout, V.J.data(), V.J.size() * sizeof(double));
  return (long)V.ipU.size();
}

// SE3Alignment::align.  from_poses: nFrom x 7, out_pose 7, out_cov 36, iters/stop: nLevels ints,
// trace: nLevels*max_iterations*64 doubles (zero-filled by the caller) or NULL.  multi != 0 selects the
// vector overload (SE3Alignment.cpp:106-146) even for nFrom == 1.
int vo_align(
  const vo_config * c, int nFrom, void ** from, const double * from_poses, void * to, const double * to_pose,
  int multi, double * out_pose, double * out_cov, int * iters, int * stop, double * trace)
{
  vo::Config cfg = to_cfg(c);
  std::vector<const vo::Frame *> fr(nFrom);
  std::vector<vo::SE3> fp(nFrom);
  for (int i = 0; i < nFrom; i++) {
    fr[i] = (const vo::Frame *)from[i];
    fp[i] = to_se3(from_poses + 7 * i);
  }
  const vo::Frame * t = (const vo::Frame *)to;
  for (int l = 0; l < fr[0]->nLevels; l++)
    for (auto f : fr)
      if (f->W[l] != f->Wd[l] || f->H[l] != f->Hd[l] || t->W[l] != f->W[l] || t->H[l] != f->H[l]) return -2;
  vo::AlignOut o = vo::align(cfg, fr, fp, t, to_se3(to_pose), multi != 0, trace);
  for (int i = 0; i < 4; i++) out_pose[i] = o.pose.q[i];
  for (int i = 0; i < 3; i++) out_pose[4 + i] = o.pose.t[i];
  std::memcpy(out_cov, o.cov.a, sizeof(o.cov.a));
  for (int l = 0; l < fr[0]->nLevels; l++) {
    if (iters) iters[l] = o.iters[l];
    if (stop) stop[l] = o.stop[l];
  }
  return o.status;
}

// CPU baseline helper: frame preparation + single-keyframe alignment of nPairs consecutive pairs with
// nThreads OpenMP threads (pair-level parallelism).  gray: (nPairs+1) frames, depth likewise (float).
// poses_out: nPairs x 7 relative poses.  Returns total pushed iterations.
long vo_baseline_sequence(
  const vo_config * c, int nFrames, int W, int H, const uint8_t * gray, const float * depth, double fx,
  double fy, double cx, double cy, int nLevels, int nThreads, double * poses_out)
{
  vo::Config cfg = to_cfg(c);
  long total = 0;
  const size_t npx = (size_t)W * H;
#pragma omp parallel for schedule(dynamic, 1) num_threads(nThreads) reduction(+ : total)
  for (int i = 0; i < nFrames - 1; i++) {
    std::vector<double> d0(npx), d1(npx);
    for (size_t k = 0; k < npx; k++) {
      d0[k] = depth[(size_t)i * npx + k];
      d1[k] = depth[(size_t)(i + 1) * npx + k];
    }
    std::unique_ptr<vo::Frame> f0(vo::frame_create(W, H, gray + (size_t)i * npx, d0.data(), fx, fy, cx, cy, nLevels, cfg.f32Storage));
    std::unique_ptr<vo::Frame> f1(vo::frame_create(W, H, gray + (size_t)(i + 1) * npx, d1.data(), fx, fy, cx, cy, nLevels, cfg.f32Storage));
    vo::SE3 I;
    vo::AlignOut o = vo::align(cfg, {f0.get()}, {I}, f1.get(), I, false, nullptr);
    for (int k = 0; k < 4; k++) poses_out[7 * i + k] = o.pose.q[k];
    for (int k = 0; k < 3; k++) poses_out[7 * i + 4 + k] = o.pose.t[k];
    for (int l = 0; l < nLevels; l++) total += o.iters[l];
  }
  return total;
}

// Independent pairs out of one frame array: pair i aligns frame i * stride + 1 against frame i * stride (stride 1: the
// consecutive pairs of a sequence, as vo_baseline_sequence; stride 2: BASELINE configs[2], frames stored ref0, cur0,
// ref1, cur1, ...).  iters_out (may be null): nPairs x nLevels completed Gauss-Newton steps.  start_pose7 (may be null:
// identity): the initial guess of every pair's second frame.
long vo_baseline_batch(
  const vo_config * c, int nPairs, int stride, int W, int H, const uint8_t * gray, const float * depth, double fx,
  double fy, double cx, double cy, int nLevels, int nThreads, double * poses_out, int * iters_out, const double * start_pose7)
{
  const vo::SE3 start = start_pose7 ? to_se3(start_pose7) : vo::SE3();
  vo::Config cfg = to_cfg(c);
  long total = 0;
  const size_t npx = (size_t)W * H;
#pragma omp parallel for schedule(dynamic, 1) num_threads(nThreads) reduction(+ : total)
  for (int i = 0; i < nPairs; i++) {
    const size_t a = (size_t)i * stride, b = a + 1;
    std::vector<double> d0(npx), d1(npx);
    for (size_t k = 0; k < npx; k++) {
      d0[k] = depth[a * npx + k];
      d1[k] = depth[b * npx + k];
    }
    std::unique_ptr<vo::Frame> f0(vo::frame_create(W, H, gray + a * npx, d0.data(), fx, fy, cx, cy, nLevels, cfg.f32Storage));
    std::unique_ptr<vo::Frame> f1(vo::frame_create(W, H, gray + b * npx, d1.data(), fx, fy, cx, cy, nLevels, cfg.f32Storage));
    vo::SE3 I;
    vo::AlignOut o = vo::align(cfg, {f0.get()}, {I}, f1.get(), start, false, nullptr);
    for (int k = 0; k < 4; k++) poses_out[7 * i + k] = o.pose.q[k];
    for (int k = 0; k < 3; k++) poses_out[7 * i + 4 + k] = o.pose.t[k];
    for (int l = 0; l < nLevels; l++) {
      total += o.iters[l];
      if (iters_out) iters_out[(size_t)i * nLevels + l] = o.iters[l];
    }
  }
  return total;
}

}  // extern "C"
