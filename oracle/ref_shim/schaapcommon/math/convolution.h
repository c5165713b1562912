// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for schaapcommon/math/convolution.h
// (schaapcommon @ 7628fc36, not vendored): forwards to the oracle's restated FFT convolution
// (oracle/fft.cc). Semantics: SURVEY Appendix B. Ours, nothing copied.
#ifndef SCHAAPCOMMON_MATH_CONVOLUTION_H_
#define SCHAAPCOMMON_MATH_CONVOLUTION_H_
#include <chrono>
#include <cstddef>
#include <map>
#include <mutex>
#include <utility>

#include "../../../oracle.h"

namespace schaapcommon::math {
// bench.py's fairness leg: how many convolutions of which size a call made and how long they took
// with the oracle's own FFT, so that the same count can be re-priced with MKL.
struct ConvolveStats {
  std::mutex mutex;
  std::map<std::pair<size_t, size_t>, size_t> calls;  // (width, height) -> count
  double seconds = 0.0;
  static ConvolveStats& Get() {
    static ConvolveStats s;
    return s;
  }
};
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
  const auto t0 = std::chrono::steady_clock::now();
  oracle::Convolve(image, kernel, image_width, image_height);
  const double dt = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
  ConvolveStats& s = ConvolveStats::Get();
  const std::lock_guard<std::mutex> lock(s.mutex);
  ++s.calls[{image_width, image_height}];
  s.seconds += dt;
}
}  // namespace schaapcommon::math
#endif
