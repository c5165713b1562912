// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for aocommon/imagecoordinates.h (ours); only
// the sky-model text writer (out of scope) uses it, so a small-angle l,m -> ra,dec suffices.
#ifndef AOCOMMON_IMAGECOORDINATES_H_
#define AOCOMMON_IMAGECOORDINATES_H_
#include <cmath>
#include <cstddef>
namespace aocommon {
class ImageCoordinates {
 public:
  template <typename T>
  static void XYToLM(size_t x, size_t y, T pixel_size_x, T pixel_size_y, size_t width,
                     size_t height, T& l, T& m) {
    const T mid_x = T(width) / 2.0, mid_y = T(height) / 2.0;
    l = (mid_x - T(x)) * pixel_size_x;
    m = (T(y) - mid_y) * pixel_size_y;
  }
  template <typename T>
  static void LMToRaDec(T l, T m, T phase_centre_ra, T phase_centre_dec, T& ra, T& dec) {
    const T n = std::sqrt(T(1.0) - l * l - m * m);
    const T cos_dec0 = std::cos(phase_centre_dec), sin_dec0 = std::sin(phase_centre_dec);
    ra = phase_centre_ra + std::atan2(l, n * cos_dec0 - m * sin_dec0);
    dec = std::asin(m * cos_dec0 + n * sin_dec0);
  }
};
}  // namespace aocommon
#endif
