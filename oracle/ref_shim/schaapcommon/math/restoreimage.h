// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for schaapcommon/math/restoreimage.h (ours).
// Local-RMS maps (math/rms_image.cc) are caller-side and off by default: throws.
#ifndef SCHAAPCOMMON_MATH_RESTOREIMAGE_H_
#define SCHAAPCOMMON_MATH_RESTOREIMAGE_H_
#include <cstddef>
#include <stdexcept>
namespace schaapcommon::math {
inline void RestoreImage(float*, const float*, size_t, size_t, long double, long double,
                         long double, long double, long double) {
  throw std::runtime_error("oracle/_ref: RestoreImage (local RMS maps) is not restated");
}
}  // namespace schaapcommon::math
#endif
