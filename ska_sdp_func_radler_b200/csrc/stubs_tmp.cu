// TEMPORARY stubs until fft.cu / multiscale.cu / subminor.cu land.
#include <cmath>
#include "fft.cuh"
#include "multiscale.cuh"
#include "subminor.cuh"
namespace rdl {
class MultiScaleState {};
size_t CalculateGoodFFTSize(size_t minimum_size) {
  size_t best = 2 * minimum_size;
  for (size_t f2 = 2; f2 < best; f2 *= 2)
    for (size_t f23 = f2; f23 < best; f23 *= 3)
      for (size_t f235 = f23; f235 < best; f235 *= 5)
        for (size_t f2357 = f235; f2357 < best; f2357 *= 7)
          if (f2357 >= minimum_size) best = f2357;
  return best;
}
size_t GetConvolutionSize(double scale, size_t original_size, double padding) {
  return CalculateGoodFFTSize(size_t(std::ceil(padding * (scale * 1.5 + double(original_size)))));
}
struct FftConvolver::Impl {};
FftConvolver::FftConvolver(size_t w, size_t h) : impl_(nullptr), width_(w), height_(h) {}
FftConvolver::~FftConvolver() {}
void FftConvolver::SetKernel(const float*, cudaStream_t) { throw std::domain_error("fft not built"); }
void FftConvolver::Convolve(const float*, float*, cudaStream_t) { throw std::domain_error("fft not built"); }
FftConvolver& GetConvolver(rdl_handle&, size_t, size_t) { throw std::domain_error("fft not built"); }
std::shared_ptr<MultiScaleState> CloneMultiScaleState(const MultiScaleState&, cudaStream_t) { return {}; }
uint64_t MultiScaleScaleCount(const MultiScaleState&) { return 0; }
rdl_scale_info MultiScaleScaleInfo(const MultiScaleState&, uint64_t) { return {}; }
void MultiScaleCopyScaleMask(const MultiScaleState&, uint64_t, uint8_t*) {}
void MultiScaleTransform(rdl_handle&, float* const*, uint64_t, uint64_t, uint64_t, float, int32_t, cudaStream_t) { throw std::domain_error("multiscale not built"); }
rdl_result RunMultiScale(rdl_handle&, DeviceCube&, cudaStream_t) { throw std::domain_error("multiscale not built"); }
rdl_result RunGenericCleanSubMinor(rdl_handle&, DeviceCube&, cudaStream_t) { throw std::domain_error("sub-minor loop not built"); }
}
