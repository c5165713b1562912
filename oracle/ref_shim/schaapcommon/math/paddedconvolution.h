// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for schaapcommon/math/paddedconvolution.h
// (ours): Untrim both -> PrepareConvolutionKernel -> Convolve -> Trim. Only the optional
// component optimisers (out of scope) reach it.
#ifndef SCHAAPCOMMON_MATH_PADDEDCONVOLUTION_H_
#define SCHAAPCOMMON_MATH_PADDEDCONVOLUTION_H_
#include <aocommon/image.h>

#include "convolution.h"

namespace schaapcommon::math {
inline void PaddedConvolution(aocommon::Image& image, const aocommon::Image& kernel,
                              size_t padded_width, size_t padded_height) {
  const size_t w = image.Width(), h = image.Height();
  aocommon::Image pi = image.Untrim(padded_width, padded_height);
  aocommon::Image pk = kernel.Untrim(padded_width, padded_height);
  aocommon::Image prepared(padded_width, padded_height);
  PrepareConvolutionKernel(prepared.Data(), pk.Data(), padded_width, padded_height);
  Convolve(pi.Data(), prepared.Data(), padded_width, padded_height);
  image = pi.Trim(w, h);
}
}  // namespace schaapcommon::math
#endif
