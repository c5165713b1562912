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
  uint64_t TakeLaunchCount() {
    const uint64_t n = launches_;
    launches_ = 0;
    return n;
  }

 private:
  void ForwardRows(const float* src, size_t sw, size_t sh, cudaStream_t stream);
  struct Impl;
  Impl* impl_;
  size_t width_, height_;
  uint64_t launches_ = 0;
};

FftConvolver& GetConvolver(rdl_handle& h, size_t width, size_t height);

}  // namespace rdl
#endif
