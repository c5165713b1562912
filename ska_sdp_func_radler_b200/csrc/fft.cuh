// FFT convolution on the device (schaapcommon::math::Convolve & friends restated
// for sm_100a) and the reference's FFT-size helpers.
#ifndef RDL_FFT_CUH_
#define RDL_FFT_CUH_

#include <cstddef>
#include <cstdint>

#include "handle.h"

namespace rdl {

// utils/fft_size_calculations.h:15-50
size_t CalculateGoodFFTSize(size_t minimum_size);
size_t GetConvolutionSize(double convolution_scale, size_t original_size, double padding);

// In-place circular convolution engine for one (width, height).
class FftConvolver {
 public:
  FftConvolver(size_t width, size_t height);
  ~FftConvolver();
  size_t Width() const { return width_; }
  size_t Height() const { return height_; }
  // Transforms `kernel` (already wrapped around the origin) and keeps its spectrum.
  void SetKernel(const float* kernel, cudaStream_t stream);
  // out = irfft2(rfft2(in) * kernel_spectrum) / (W*H); in may equal out.
  void Convolve(const float* in, float* out, cudaStream_t stream);
  uint64_t TakeLaunchCount() {
    const uint64_t n = launches_;
    launches_ = 0;
    return n;
  }

 private:
  struct Impl;
  Impl* impl_;
  size_t width_, height_;
  uint64_t launches_ = 0;
};

FftConvolver& GetConvolver(rdl_handle& h, size_t width, size_t height);

}  // namespace rdl
#endif
