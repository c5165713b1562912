// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for schaapcommon/math/convolution.h
// (schaapcommon @ 7628fc36, not vendored): forwards to the oracle's restated FFT convolution
// (oracle/fft.cc). Semantics: SURVEY Appendix B. Ours, nothing copied.
#ifndef SCHAAPCOMMON_MATH_CONVOLUTION_H_
#define SCHAAPCOMMON_MATH_CONVOLUTION_H_
#include <cstddef>

#include "../../../oracle.h"

namespace schaapcommon::math {
inline void MakeFftwfPlannerThreadSafe() {}
inline void PrepareSmallConvolutionKernel(float* dest, size_t image_width, size_t image_height,
                                          const float* kernel, size_t kernel_size) {
  oracle::PrepareSmallConvolutionKernel(dest, image_width, image_height, kernel, kernel_size);
}
inline void PrepareConvolutionKernel(float* dest, const float* source, size_t image_width,
                                     size_t image_height) {
  oracle::PrepareConvolutionKernel(dest, source, image_width, image_height);
}
inline void Convolve(float* image, const float* kernel, size_t image_width,
                     size_t image_height) {
  oracle::Convolve(image, kernel, image_width, image_height);
}
}  // namespace schaapcommon::math
#endif
