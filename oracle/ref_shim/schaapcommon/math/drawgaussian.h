// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for schaapcommon/math/drawgaussian.h (ours).
// Only AspAlgorithm (out of scope) draws Gaussians: throws.
#ifndef SCHAAPCOMMON_MATH_DRAWGAUSSIAN_H_
#define SCHAAPCOMMON_MATH_DRAWGAUSSIAN_H_
#include <cstddef>
#include <stdexcept>

#include "ellipse.h"
namespace schaapcommon::math {
inline void DrawGaussianToXy(float*, size_t, size_t, double, double, const Ellipse&, long double) {
  throw std::runtime_error("oracle/_ref: DrawGaussianToXy (ASP) is not restated");
}
}  // namespace schaapcommon::math
#endif
