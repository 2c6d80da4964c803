// Stand-in for the reference's utils/include/utils/utils.h: logging, plotting and timing (easylogging++, matplotlib,
// OpenCV windows in the reference) are out of scope and become no-ops with the same spelling,
// so the reference's translation units compile unchanged.  Nothing streamed into a log is evaluated for effect.
#ifndef REF_SHIM_UTILS_H
#define REF_SHIM_UTILS_H
#include <core/core.h>

#include <memory>
#include <string>

namespace ref_shim
{
struct NullLog
{
  template <class T>
  NullLog & operator<<(const T &)
  {
    return *this;
  }
  NullLog & operator<<(std::ostream & (*)(std::ostream &)) { return *this; }
};
}  // namespace ref_shim

#define CLOG(level, name) ::ref_shim::NullLog()
#define LOG(level) ::ref_shim::NullLog()
#define SYSTEM(loglevel) CLOG(loglevel, "system")
#define IMAGE_ALIGNMENT(loglevel) CLOG(loglevel, "image_alignment")
#define SOLVER(loglevel) CLOG(loglevel, "solver")
#define LOG_IMG(name) ::ref_shim::NullLog()
#define LOG_PLT(name) ::ref_shim::NullLog()
#define TIMED_FUNC(name)
#define TIMED_SCOPE(name, label)

namespace pd::vslam
{
struct Log
{
  static void get(const std::string &) {}
};
namespace vis
{
struct Histogram
{
  template <class... A>
  explicit Histogram(A &&...)
  {
  }
};
struct PlotGaussNewton
{
  template <class... A>
  explicit PlotGaussNewton(A &&...)
  {
  }
};
}  // namespace vis
}  // namespace pd::vslam
#endif
