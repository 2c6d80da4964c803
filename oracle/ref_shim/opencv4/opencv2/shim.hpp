// Stand-in for the OpenCV calls of the reference's core/src/Frame.cpp (eigen2cv / cv2eigen, GaussianBlur 3x3, Sobel 3x3
// into CV_16S, buildPyramid), so that Frame.cpp compiles UNCHANGED (oracle/build_ref.py).  The arithmetic is the
// oracle's restatement of those OpenCV operators (oracle/vslam_oracle.cpp: vo_blur3_u8, vo_sobel3_s16, vo_pyr_down_u8),
// which tests/test_oracle_cpu.py pins bit for bit against outputs of cv2 committed under tests/golden/.
// Test infrastructure; not OpenCV.
#ifndef REF_SHIM_OPENCV
#define REF_SHIM_OPENCV
#include <Eigen/Dense>
#include <cstdint>
#include <stdexcept>
#include <vector>

extern "C" {
void vo_pyr_down_u8(const uint8_t * src, int W, int H, uint8_t * dst);
void vo_blur3_u8(const uint8_t * src, int W, int H, uint8_t * dst);
void vo_sobel3_s16(const uint8_t * g, int W, int H, int16_t * dx, int16_t * dy);
}

#define CV_8U 0
#define CV_16S 3

namespace cv
{
enum { BORDER_DEFAULT = 4 };
struct Size
{
  Size(int w, int h) : width(w), height(h) {}
  int width, height;
};
struct Mat  // row-major, one channel, u8 or s16
{
  int rows = 0, cols = 0, depth = CV_8U;
  std::vector<uint8_t> u8;
  std::vector<int16_t> s16;
};

template <class S, int R, int C>
void eigen2cv(const Eigen::Matrix<S, R, C> & src, Mat & dst)
{
  static_assert(sizeof(S) == 1, "only 8-bit images travel into the stand-in");
  dst.rows = (int)src.rows();
  dst.cols = (int)src.cols();
  dst.depth = CV_8U;
  dst.u8.resize((size_t)dst.rows * dst.cols);
  for (int v = 0; v < dst.rows; v++)
    for (int u = 0; u < dst.cols; u++) dst.u8[(size_t)v * dst.cols + u] = src(v, u);
}
template <class S, int R, int C>
void cv2eigen(const Mat & src, Eigen::Matrix<S, R, C> & dst)
{
  dst.resize(src.rows, src.cols);
  for (int v = 0; v < src.rows; v++)
    for (int u = 0; u < src.cols; u++)
      dst(v, u) = src.depth == CV_8U ? (S)src.u8[(size_t)v * src.cols + u] : (S)src.s16[(size_t)v * src.cols + u];
}
inline void GaussianBlur(const Mat & src, Mat & dst, Size k, double sx, double sy, int border)
{
  if (src.depth != CV_8U || k.width != 3 || k.height != 3 || sx != 0 || sy != 0 || border != BORDER_DEFAULT)
    throw std::runtime_error("ref_shim: only GaussianBlur(u8, 3x3, sigma 0, BORDER_DEFAULT)");
  Mat out;
  out.rows = src.rows;
  out.cols = src.cols;
  out.u8.resize(src.u8.size());
  vo_blur3_u8(src.u8.data(), src.cols, src.rows, out.u8.data());
  dst = out;
}
inline void Sobel(const Mat & src, Mat & dst, int ddepth, int dx, int dy, int ksize)
{
  if (src.depth != CV_8U || ddepth != CV_16S || ksize != 3 || dx + dy != 1)
    throw std::runtime_error("ref_shim: only Sobel(u8 -> s16, first derivative, 3x3)");
  std::vector<int16_t> gx(src.u8.size()), gy(src.u8.size());
  vo_sobel3_s16(src.u8.data(), src.cols, src.rows, gx.data(), gy.data());
  Mat out;
  out.rows = src.rows;
  out.cols = src.cols;
  out.depth = CV_16S;
  out.s16 = dx ? gx : gy;
  dst = out;
}
inline void buildPyramid(const Mat & src, std::vector<Mat> & dst, int maxlevel)
{
  dst.assign(1, src);
  for (int i = 1; i <= maxlevel; i++) {
    const Mat & p = dst.back();
    Mat d;
    d.rows = (p.rows + 1) / 2;
    d.cols = (p.cols + 1) / 2;
    d.u8.resize((size_t)d.rows * d.cols);
    vo_pyr_down_u8(p.u8.data(), p.cols, p.rows, d.u8.data());
    dst.push_back(d);
  }
}
}  // namespace cv
#endif
