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
  // content_width x content_height: the largest source / destination ever passed to
  // ConvolveWith (0 = the whole frame). Only used when the frame has no native plan.
  FftConvolver(size_t width, size_t height, size_t content_width = 0, size_t content_height = 0);
  ~FftConvolver();
  // true when width x height is transformed directly; false when the circular convolution is
  // evaluated through a larger smooth frame (see fft.cu, "embedded frames")
  bool Native() const { return inner_ == nullptr; }
  static bool HasNativePlan(size_t width, size_t height);
  size_t Width() const { return width_; }
  size_t Height() const { return height_; }
  // Transforms `kernel` (already wrapped around the origin) and keeps its spectrum.
  void SetKernel(const float* kernel, cudaStream_t stream);
  // out = irfft2(rfft2(in) * kernel_spectrum) / (W*H); in may equal out.
  void Convolve(const float* in, float* out, cudaStream_t stream);

  // -- building blocks used by the multiscale / sub-minor code --------------------------
  // Number of float2 elements of one kernel spectrum of this frame size.
  size_t SpectrumSize() const;
  // spectrum_out = scaled half spectrum of a W x H kernel image (origin at pixel 0,0).
  void MakeKernelSpectrum(const float* kernel_image, float2* spectrum_out, cudaStream_t stream);
  // PrepareSmallConvolutionKernel (n x n kernel wrapped around the origin) + spectrum.
  void MakeSmallKernelSpectrum(const float* small_kernel, size_t n, float2* spectrum_out,
                               cudaStream_t stream);
  // PrepareConvolutionKernel(Untrim(psf, pw x ph -> W x H)) + spectrum.
  void MakePsfKernelSpectrum(const float* psf, size_t pw, size_t ph, float2* spectrum_out,
                             cudaStream_t stream);
  // Untrim(src, sw x sh) -> convolve -> Trim(dw x dh), fused epilogue:
  // mode 0: dst = result, 1: dst -= result, 2: dst += result. src may equal dst.
  void ConvolveWith(const float* src, size_t sw, size_t sh, const float2* kernel_spectrum,
                    float* dst, size_t dw, size_t dh, int mode, cudaStream_t stream);
  // The same with explicit positions of the source and the destination inside the frame.
  void ConvolveAt(const float* src, size_t sw, size_t sh, size_t sox, size_t soy,
                  const float2* kernel_spectrum, float* dst, size_t dw, size_t dh, size_t dox,
                  size_t doy, int mode, cudaStream_t stream);
  uint64_t TakeLaunchCount() {
    const uint64_t n = launches_ + (inner_ ? inner_->TakeLaunchCount() : 0);
    launches_ = 0;
    return n;
  }

 private:
  void ForwardRows(const float* src, size_t sw, size_t sh, size_t ox, size_t oy, cudaStream_t stream);
  void ExpandKernel(float2* spectrum_out, cudaStream_t stream);
  struct Impl;
  Impl* impl_ = nullptr;
  FftConvolver* inner_ = nullptr;  // embedded frames: the engine of the larger smooth frame
  size_t width_, height_, content_w_, content_h_;
  uint64_t launches_ = 0;
};

// content_*: see FftConvolver's constructor
FftConvolver& GetConvolver(rdl_handle& h, size_t width, size_t height, size_t content_width = 0,
                           size_t content_height = 0);

}  // namespace rdl
#endif
