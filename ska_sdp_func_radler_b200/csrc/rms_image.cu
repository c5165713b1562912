// Local-RMS maps on the device: cpp/math/rms_image.cc of the reference,
// the producer of the image that SetRmsFactorImage hands to the minor loops.
//   :39-78   SlidingMinimum / SlidingMaximum   -> RmsSlidingMinimum      (bit-exact)
//   :97-123  MakeRmsFactorImage                -> RmsMakeFactorImage     (bit-exact for strength 0
//            and 1; std::pow otherwise, within an ulp of the device's pow)
//   :16-37   Make, :80-95 MakeWithNegativityLimit -> RmsMake: sqrt(norm * (image^2 (*) Gaussian
//            window)), max(|sliding minimum| * 0.3). The Gaussian window is
//            schaapcommon::math::RestoreImage in the reference (third-party, absent from the
//            tree): restated here as a circular FFT convolution with a unit-peak elliptical
//            Gaussian of FWHM beam * window, position angle from the +y axis towards +x. That
//            part of the parity is UNPINNED; the normalisation, the square root, the sliding
//            minimum and the combination follow the reference's lines.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <stdexcept>
#include <vector>

#include "fft.cuh"
#include "handle.h"
#include "rms_image.cuh"

namespace rdl {
namespace {

constexpr int kRowTile = 256;

// temp[y][x] = min of in[y][left, right), left = max(x, w2) - w2, right = min(x, W - w2) + w2
// (rms_image.cc:47-55). Strict '<' from the left: the value std::min_element returns.
template <bool NEGATE>
__global__ void __launch_bounds__(kRowTile) SlidingRowKernel(const float* __restrict__ in,
                                                             float* __restrict__ out, int width,
                                                             int w2) {
  extern __shared__ float tile[];  // kRowTile + 2 * w2
  const int y = blockIdx.y, x0 = blockIdx.x * kRowTile;
  const float* row = in + size_t(y) * width;
  const int lo = x0 - w2, span = kRowTile + 2 * w2;
  for (int i = threadIdx.x; i < span; i += kRowTile) {
    const int x = lo + i;
    const float v = (x >= 0 && x < width) ? row[x] : 0.0f;
    tile[i] = NEGATE ? -v : v;
  }
  __syncthreads();
  const int x = x0 + threadIdx.x;
  if (x >= width) return;
  const int left = max(x, w2) - w2, right = min(x, width - w2) + w2;
  float m = tile[left - lo];
  for (int i = left + 1; i < right; ++i) {
    const float v = tile[i - lo];
    if (v < m) m = v;
  }
  out[size_t(y) * width + x] = m;
}

// out[y][x] = min of temp[top, bottom)[x] (rms_image.cc:58-76); NEGATE flips the result back
template <bool NEGATE>
__global__ void SlidingColumnKernel(const float* __restrict__ temp, float* __restrict__ out, int width,
                                    int height, int w2) {
  const int x = blockIdx.x * blockDim.x + threadIdx.x, y = blockIdx.y;
  if (x >= width) return;
  const int top = max(y, w2) - w2, bottom = min(y, height - w2) + w2;
  float m = temp[size_t(top) * width + x];
  for (int r = top + 1; r < bottom; ++r) {
    const float v = temp[size_t(r) * width + x];
    if (v < m) m = v;
  }
  out[size_t(y) * width + x] = NEGATE ? -m : m;
}

__global__ void MinKernel(const float* a, size_t n, unsigned* out) {
  // ordered-integer encoding: smaller float <=> smaller unsigned
  unsigned best = 0xffffffffu;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    const unsigned b = __float_as_uint(a[i]);
    const unsigned o = (b & 0x80000000u) ? ~b : (b | 0x80000000u);
    best = min(best, o);
  }
  for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
  if ((threadIdx.x & 31) == 0) atomicMin(out, best);
}

// value != 0: value = stddev / value (strength 1), pow(stddev / value, strength) otherwise
__global__ void FactorKernel(float* a, size_t n, double stddev, double strength, int mode) {
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    const float v = a[i];
    if (mode == 0) {
      a[i] = 1.0f;
    } else if (v != 0.0f) {
      const double q = stddev / double(v);
      a[i] = mode == 1 ? float(q) : float(pow(q, strength));
    }
  }
}

__global__ void SquareKernel(const float* in, float* out, size_t n) {
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    out[i] = __fmul_rn(in[i], in[i]);
}
// val = sqrt(val * norm) in double, as the reference's float * double expression
__global__ void SqrtNormKernel(float* a, size_t n, double norm) {
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    a[i] = float(sqrt(double(a[i]) * norm));
}
// rms = max(rms, |sliding_minimum| * (1.5 / 5.0))   (rms_image.cc:90-93)
__global__ void NegativityKernel(float* rms, const float* sliding_min, size_t n) {
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    const float limit = float(double(fabsf(sliding_min[i])) * (1.5 / 5.0));
    rms[i] = fmaxf(rms[i], limit);
  }
}

int Grid(size_t n) { return int(std::clamp<size_t>((n + 255) / 256, 1, size_t(kSmCount) * 16)); }

}  // namespace

void RmsSlidingMinimum(rdl_handle& h, const float* in, float* out, size_t width, size_t height,
                       size_t window_size, bool maximum, cudaStream_t stream) {
  const size_t w2 = window_size / 2;
  if (w2 == 0) throw std::invalid_argument("sliding window must span at least two pixels");
  if (w2 > width || w2 > height) throw std::invalid_argument("sliding window larger than the image");
  if (w2 > 8192) throw std::domain_error("sliding windows beyond 16384 pixels are not supported");
  if (in == out) throw std::invalid_argument("sliding minimum is out of place");
  h.rms_temp.Reserve(width * height);
  const dim3 row_grid(unsigned((width + kRowTile - 1) / kRowTile), unsigned(height));
  const size_t smem = (kRowTile + 2 * w2) * sizeof(float);
  const dim3 col_grid(unsigned((width + 255) / 256), unsigned(height));
  if (maximum) {
    if (smem > size_t(48) * 1024)
      RDL_CUDA(cudaFuncSetAttribute(SlidingRowKernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    int(smem)));
    SlidingRowKernel<true><<<row_grid, kRowTile, smem, stream>>>(in, h.rms_temp.get(), int(width), int(w2));
    SlidingColumnKernel<true><<<col_grid, 256, 0, stream>>>(h.rms_temp.get(), out, int(width), int(height),
                                                            int(w2));
  } else {
    if (smem > size_t(48) * 1024)
      RDL_CUDA(cudaFuncSetAttribute(SlidingRowKernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                    int(smem)));
    SlidingRowKernel<false><<<row_grid, kRowTile, smem, stream>>>(in, h.rms_temp.get(), int(width), int(w2));
    SlidingColumnKernel<false><<<col_grid, 256, 0, stream>>>(h.rms_temp.get(), out, int(width),
                                                             int(height), int(w2));
  }
  RDL_CUDA(cudaGetLastError());
  h.launches += 2;
}

double RmsMakeFactorImage(rdl_handle& h, float* image, size_t n, double strength, cudaStream_t stream) {
  if (n == 0) throw std::invalid_argument("empty RMS image");
  h.rms_scalar.Reserve(1);
  const unsigned init = 0xffffffffu;
  RDL_CUDA(cudaMemcpyAsync(h.rms_scalar.get(), &init, sizeof(init), cudaMemcpyHostToDevice, stream));
  MinKernel<<<Grid(n), 256, 0, stream>>>(image, n, h.rms_scalar.get());
  unsigned o = 0;
  RDL_CUDA(cudaMemcpyAsync(&o, h.rms_scalar.get(), sizeof(o), cudaMemcpyDeviceToHost, stream));
  RDL_CUDA(cudaStreamSynchronize(stream));
  const unsigned bits = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
  float min_value;
  memcpy(&min_value, &bits, sizeof(float));
  const double stddev = min_value;
  if (stddev < 0.0)
    throw std::invalid_argument("RMS image can only contain values >= 0, but contains values < 0.0");
  const int mode = strength == 1.0 ? 1 : (strength == 0.0 ? 0 : 2);
  FactorKernel<<<Grid(n), 256, 0, stream>>>(image, n, stddev, strength, mode);
  RDL_CUDA(cudaGetLastError());
  h.launches += 2;
  return stddev;
}

void RmsMake(rdl_handle& h, const float* in, float* out, size_t width, size_t height, double window_size,
             double beam_major, double beam_minor, double beam_pa, double pixel_scale_l,
             double pixel_scale_m, bool negativity_limit, cudaStream_t stream) {
  if (!(beam_major > 0.0) || !(beam_minor > 0.0) || !(pixel_scale_l > 0.0) || !(pixel_scale_m > 0.0) ||
      !(window_size > 0.0))
    throw std::invalid_argument("local RMS needs a positive beam, pixel scale and window");
  if (in == out) throw std::invalid_argument("local RMS is out of place");
  const size_t n = width * height;
  // the Gaussian window: unit peak at (width / 2, height / 2), FWHM = beam * window
  const double fwhm_to_sigma = 1.0 / (2.0 * std::sqrt(2.0 * std::log(2.0)));
  const double s_maj = beam_major * window_size * fwhm_to_sigma / pixel_scale_l;  // pixels
  const double s_min = beam_minor * window_size * fwhm_to_sigma / pixel_scale_m;
  std::vector<float> kernel(n);
  const double sn = std::sin(beam_pa), cs = std::cos(beam_pa);
  for (size_t y = 0; y != height; ++y)
    for (size_t x = 0; x != width; ++x) {
      const double dx = double(x) - double(width / 2), dy = double(y) - double(height / 2);
      const double u = dx * sn + dy * cs, v = dx * cs - dy * sn;  // along / across the major axis
      const double e = -0.5 * (u * u / (s_maj * s_maj) + v * v / (s_min * s_min));
      kernel[y * width + x] = e < -60.0 ? 0.0f : float(std::exp(e));
    }
  h.rms_kernel.Reserve(n);
  RDL_CUDA(cudaMemcpyAsync(h.rms_kernel.get(), kernel.data(), n * sizeof(float), cudaMemcpyHostToDevice,
                           stream));
  RDL_CUDA(cudaStreamSynchronize(stream));  // `kernel` is pageable host memory
  FftConvolver& conv = GetConvolver(h, width, height);
  h.fft_spectrum.Reserve(conv.SpectrumSize());
  conv.MakePsfKernelSpectrum(h.rms_kernel.get(), width, height, h.fft_spectrum.get(), stream);
  SquareKernel<<<Grid(n), 256, 0, stream>>>(in, out, n);
  conv.ConvolveWith(out, width, height, h.fft_spectrum.get(), out, width, height, 0, stream);
  // rms_image.cc:28-36 (long double in the reference; the value is rounded to double there too)
  const long double sigma_maj = static_cast<long double>(beam_major) / (2.0L * sqrtl(2.0L * logl(2.0L)));
  const long double sigma_min = static_cast<long double>(beam_minor) / (2.0L * sqrtl(2.0L * logl(2.0L)));
  const double s = std::sqrt(2.0 * M_PI);
  const double norm = double(1.0L / (s * sigma_maj / pixel_scale_l * window_size * s * sigma_min /
                                     pixel_scale_l * window_size));
  SqrtNormKernel<<<Grid(n), 256, 0, stream>>>(out, n, norm);
  h.launches += 2 + conv.TakeLaunchCount();
  if (negativity_limit) {
    const double beam_in_pixels = std::max(beam_major / pixel_scale_l, 1.0);
    h.rms_min.Reserve(n);
    RmsSlidingMinimum(h, in, h.rms_min.get(), width, height, size_t(window_size * beam_in_pixels), false,
                      stream);
    NegativityKernel<<<Grid(n), 256, 0, stream>>>(out, h.rms_min.get(), n);
    ++h.launches;
  }
  RDL_CUDA(cudaGetLastError());
}

}  // namespace rdl
