// IuwtDeconvolution::ExecuteMajorIteration on the device:
//   cpp/algorithms/iuwt_deconvolution.h:21-38                 the DeconvolutionAlgorithm wrapper
//   cpp/algorithms/iuwt_deconvolution_algorithm.cc:803-918    PerformMajorIteration
//   :414-498 FindAndDeconvolveStructure, :500-626 FillAndDeconvolveStructure,
//   :326-407 RunConjugateGradient, :42-102 MeasureRMSPerScale, :180-262 BoundingBox / AdjustBox.
//
// The host keeps the reference's scalar control flow (scale bookkeeping, thresholds, the
// conjugate-gradient stopping rules) and drives device work: decompositions (iuwt.cu, including
// the reference's input-aliased Decompose(x, x), which is what the CG loop really computes),
// structure selection / Mad / arg-max (iuwt_analysis.cu), FFT convolutions (fft.cu) and the
// reductions below. Every image stays in HBM; per step the host reads a few scalars.
//
// Where parity is by tolerance and why: DotProduct, Snr and Image::RMS are float32 SEQUENTIAL
// sums in the reference; here they are fixed-order double-precision tree sums (deterministic,
// run to run). The step sizes therefore agree to ~1e-7 relative and the images to the FFT
// tolerance; the stopping rules compare those sums against constants (SNR > 100, SNR < previous,
// rmsAfter > rmsBefore), so a run can only branch differently at a knife edge.
//
// Cubes with several images (:628-801, PerformSubImageFitAll / FitSingle / ComponentFit[Boxed]): the
// structure model of the integrated image is split into its connected components (row-major
// seed scan + FloodFill2D), every component is refitted per image inside its own box, and the
// images get component * factor(image) / factor(integrated). The component sums are float32
// sequential sums in flood-fill order in the reference, fixed-order double sums here.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <limits>
#include <vector>

#include "fft.cuh"
#include "iuwt.cuh"

namespace rdl {
namespace {

constexpr int kRedBlocks = 592;  // 4 per SM
constexpr int kRedThreads = 256;

__device__ __forceinline__ double BlockSum(double v, double* red) {
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
  __syncthreads();
  if (threadIdx.x < 32) {
    v = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.0;
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  }
  __syncthreads();
  return v;
}

// partial[b] = sum over the block's pixels of a*b (fixed grid => fixed summation order)
__global__ void __launch_bounds__(kRedThreads) DotKernel(const float* a, const float* b, size_t n,
                                                         double* partial) {
  __shared__ double red[32];
  double acc = 0.0;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    acc += double(a[i]) * double(b[i]);
  acc = BlockSum(acc, red);
  if (threadIdx.x == 0) partial[blockIdx.x] = acc;
}

// Snr (iuwt_deconvolution_algorithm.cc:308-323) over masked coefficient planes: both
// decompositions carry the same mask, outside of which they are zero.
__global__ void __launch_bounds__(kRedThreads) SnrKernel(const float* noisy, const float* model,
                                                         const uint8_t* mask, size_t n,
                                                         double* partial) {
  __shared__ double red[32];
  double m_sum = 0.0, n_sum = 0.0;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    if (!mask[i]) continue;
    const double m = model[i], d = m - double(noisy[i]);
    m_sum += m * m;
    n_sum += d * d;
  }
  m_sum = BlockSum(m_sum, red);
  n_sum = BlockSum(n_sum, red);
  if (threadIdx.x == 0) {
    partial[2 * blockIdx.x] = m_sum;
    partial[2 * blockIdx.x + 1] = n_sum;
  }
}

__global__ void FinishSumKernel(const double* partial, int n_partial, int n_values, double* out) {
  // one warp per value, fixed order
  const int v = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (v >= n_values) return;
  double acc = 0.0;
  for (int i = lane; i < n_partial; i += 32) acc += partial[size_t(i) * n_values + v];
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if (lane == 0) out[v] = acc;
}

// y = fma(x, f, y)  (Image::AddWithFactor)
__global__ void AxpyKernel(float* y, const float* x, float f, size_t n) {
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    y[i] = __fmaf_rn(x[i], f, y[i]);
}
// g = fma(g, f, r): gradient_{i+1} = residual_{i+1} + gradstep * gradient_i (:372-375)
__global__ void GradientKernel(float* g, const float* r, float f, size_t n) {
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    g[i] = __fmaf_rn(g[i], f, r[i]);
}
__global__ void ScaleKernel(float* y, float f, size_t n) {
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    y[i] = __fmul_rn(y[i], f);
}
// y += x (add = true) or y -= x
__global__ void AddSubKernel(float* y, const float* x, bool add, size_t n) {
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    y[i] = add ? __fadd_rn(y[i], x[i]) : __fsub_rn(y[i], x[i]);
}

// min and max of an image: out[0] = max bits, out[1] = min bits (ordered-integer encoding)
__device__ __forceinline__ unsigned Ordered(float v) {
  const unsigned b = __float_as_uint(v);
  return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__host__ __device__ inline float Unordered(unsigned o) {
  const unsigned b = (o & 0x80000000u) ? (o & 0x7fffffffu) : ~o;
#ifdef __CUDA_ARCH__
  return __uint_as_float(b);
#else
  float f;
  memcpy(&f, &b, 4);
  return f;
#endif
}
__global__ void MinMaxKernel(const float* a, size_t n, unsigned* out) {
  unsigned mx = 0u, mn = 0xffffffffu;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    const unsigned o = Ordered(a[i]);
    mx = max(mx, o);
    mn = min(mn, o);
  }
  for (int o = 16; o > 0; o >>= 1) {
    mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
  }
  if ((threadIdx.x & 31) == 0) {
    atomicMax(out, mx);
    atomicMin(out + 1, mn);
  }
}

// per row: first and last x with |v| > level (width / -1 when none): BoundingBox (:180-214)
__global__ void RowExtentKernel(const float* a, int width, int height, double level, int* first,
                                int* last) {
  const int y = blockIdx.x;
  __shared__ int s_first, s_last;
  if (threadIdx.x == 0) {
    s_first = width;
    s_last = -1;
  }
  __syncthreads();
  int f = width, l = -1;
  for (int x = threadIdx.x; x < width; x += blockDim.x)
    if (double(fabsf(a[size_t(y) * width + x])) > level) {
      f = min(f, x);
      l = max(l, x);
    }
  if (f < width) {
    atomicMin(&s_first, f);
    atomicMax(&s_last, l);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    first[y] = s_first;
    last[y] = s_last;
  }
}

// dst (bw x bh) = box of src at (x1, y1); mask (optional, same box of a frame-sized byte plane)
// zeroes the pixels outside it (CreateTrimmed of a decomposition whose mask was applied)
__global__ void TrimKernel(float* dst, const float* src, int src_width, int x1, int y1, int bw, int bh,
                           const uint8_t* mask) {
  const size_t n = size_t(bw) * bh;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    const int y = int(i / bw), x = int(i - size_t(y) * bw);
    const size_t j = size_t(y + y1) * src_width + x + x1;
    dst[i] = (mask == nullptr || mask[j]) ? src[j] : 0.0f;
  }
}
__global__ void TrimBytesKernel(uint8_t* dst, const uint8_t* src, int src_width, int x1, int y1, int bw,
                                int bh) {
  const size_t n = size_t(bw) * bh;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    const int y = int(i / bw), x = int(i - size_t(y) * bw);
    dst[i] = src[size_t(y + y1) * src_width + x + x1];
  }
}
// Untrim (:274-306): dst (w x h) = 0 with the box placed at (x1, y1)
__global__ void UntrimKernel(float* dst, int width, int height, const float* box, int x1, int y1, int bw,
                             int bh) {
  const size_t n = size_t(width) * height;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    const int y = int(i / width), x = int(i - size_t(y) * width);
    const bool in = x >= x1 && x < x1 + bw && y >= y1 && y < y1 + bh;
    dst[i] = in ? box[size_t(y - y1) * bw + (x - x1)] : 0.0f;
  }
}

int GridFor(size_t n) { return int(std::clamp<size_t>((n + 255) / 256, 1, size_t(kSmCount) * 16)); }

// ---- PerformSubImageFitSingle (:672-742): components of the structure model -------------------
// first row-major pixel with !mask && |model| > threshold (the seed scan of :688-693)
__global__ void FindSeedKernel(const float* model, const uint8_t* mask, size_t n, float threshold,
                               unsigned* first) {
  unsigned best = 0xffffffffu;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    if (!mask[i] && fabsf(model[i]) > threshold) best = min(best, unsigned(i));
  for (int o = 16; o > 0; o >>= 1) best = min(best, __shfl_xor_sync(0xffffffffu, best, o));
  if ((threadIdx.x & 31) == 0 && best != 0xffffffffu) atomicMin(first, best);
}
// area = pixels the flood fill has just set; sub_dirty = model inside the area, 0 elsewhere;
// box[0..3] = min x, max x, min y, max y of the area
__global__ void AreaKernel(const float* model, const uint8_t* mask_after, const uint8_t* mask_before,
                           int width, size_t n, uint8_t* area, float* sub_dirty, unsigned* box) {
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    const bool in = mask_after[i] && !mask_before[i];
    area[i] = in ? 1 : 0;
    sub_dirty[i] = in ? model[i] : 0.0f;
    if (in) {
      const unsigned y = unsigned(i / width), x = unsigned(i - size_t(y) * width);
      atomicMin(box + 0, x);
      atomicMax(box + 1, x);
      atomicMin(box + 2, y);
      atomicMax(box + 3, y);
    }
  }
}
// partial sums over the area of model and masked_dirty, both bw x bh boxes at (x1, y1) of the
// frame the area mask lives in (PerformSubImageComponentFit, :790-794)
__global__ void __launch_bounds__(kRedThreads) AreaSumsKernel(const float* model, const float* masked_dirty,
                                                              const uint8_t* area, int area_width, int x1,
                                                              int y1, int bw, int bh, double* partial) {
  __shared__ double red[32];
  double m_sum = 0.0, d_sum = 0.0;
  const size_t n = size_t(bw) * bh;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    const int y = int(i / bw), x = int(i - size_t(y) * bw);
    if (!area[size_t(y + y1) * area_width + x + x1]) continue;
    m_sum += double(model[i]);
    d_sum += double(masked_dirty[i]);
  }
  m_sum = BlockSum(m_sum, red);
  d_sum = BlockSum(d_sum, red);
  if (threadIdx.x == 0) {
    partial[2 * blockIdx.x] = m_sum;
    partial[2 * blockIdx.x + 1] = d_sum;
  }
}
// fitted[i] += model[i] * factor / integrated_factor inside the area (:729-733)
__global__ void FittedAddKernel(float* fitted, const float* model, const uint8_t* area, float factor,
                                float integrated, size_t n) {
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    if (area[i]) fitted[i] = __fadd_rn(fitted[i], __fdiv_rn(__fmul_rn(model[i], factor), integrated));
}

int EndScale(size_t max_image_dimension) {  // iuwt_decomposition.h:182-184
  return std::max(int(std::log2(double(max_image_dimension))) - 3, 2);
}
size_t MinImageDimension(int end_scale) { return size_t(1) << (end_scale + 3); }

struct ScaleResponse {
  float rms = 0.0f, peak_response = 0.0f, peak_response_to_next_scale = 0.0f;
};
struct ValComponent {
  size_t x = 0, y = 0;
  int scale = 0;
  float val = 0.0f;
};

// One IuwtDecomposition: n_scales coefficient planes of w x h (the smooth remainder is never kept:
// every Decompose of the algorithm passes include_largest = false).
struct Decomposition {
  int n_scales = 0;
  size_t w = 0, h = 0;
  DeviceBuffer<float> planes;
  void Shape(int scales, size_t width, size_t height) {
    n_scales = scales;
    w = width;
    h = height;
    planes.Reserve(size_t(scales) * width * height);
  }
  float* Plane(int s) const { return planes.get() + size_t(s) * w * h; }
};

class IuwtRun {
 public:
  IuwtRun(rdl_handle& h, DeviceCube& cube, cudaStream_t stream)
      : h_(h), cube_(cube), stream_(stream), width_(cube.width), height_(cube.height) {
    const Settings& s = h.settings;
    minor_loop_gain_ = s.minor_loop_gain;
    major_loop_gain_ = s.major_loop_gain;
    clean_border_ = s.clean_border_ratio;
    allow_negative_ = s.allow_negative_components;
    absolute_threshold_ = s.threshold;
    mask_ = h.mask.get();
  }

  float PerformMajorIteration(uint64_t& iter_counter, uint64_t n_iter, bool& reached_major_threshold);

 private:
  // ---- scalar read-backs ----------------------------------------------------------------------
  void Sync() { RDL_CUDA(cudaStreamSynchronize(stream_)); }
  float ReadPixel(const float* image, size_t index) {
    float v = 0.0f;
    RDL_CUDA(cudaMemcpyAsync(&v, image + index, sizeof(float), cudaMemcpyDeviceToHost, stream_));
    Sync();
    return v;
  }
  double Dot(const float* a, const float* b, size_t n) {
    partial_.Reserve(2 * kRedBlocks + 8);
    DotKernel<<<kRedBlocks, kRedThreads, 0, stream_>>>(a, b, n, partial_.get());
    FinishSumKernel<<<1, 32, 0, stream_>>>(partial_.get(), kRedBlocks, 1, partial_.get() + 2 * kRedBlocks);
    h_.launches += 2;
    double v = 0.0;
    RDL_CUDA(cudaMemcpyAsync(&v, partial_.get() + 2 * kRedBlocks, sizeof(double), cudaMemcpyDeviceToHost, stream_));
    Sync();
    return v;
  }
  float DotProduct(const float* a, const float* b, size_t n) { return float(Dot(a, b, n)); }
  float Rms(const float* a, size_t n) { return float(std::sqrt(Dot(a, a, n) / double(n))); }
  void Copy(float* dst, const float* src, size_t n) {
    RDL_CUDA(cudaMemcpyAsync(dst, src, n * sizeof(float), cudaMemcpyDeviceToDevice, stream_));
  }
  void Zero(float* dst, size_t n) { RDL_CUDA(cudaMemsetAsync(dst, 0, n * sizeof(float), stream_)); }
  void Axpy(float* y, const float* x, float f, size_t n) {
    AxpyKernel<<<GridFor(n), 256, 0, stream_>>>(y, x, f, n);
    ++h_.launches;
  }

  // ---- decomposition helpers --------------------------------------------------------------------
  void Decompose(Decomposition& d, const float* input) {
    IuwtDecompose(h_, input, d.w, d.h, d.n_scales, false, d.planes.get(), stream_);
  }
  void DecomposeAliased(Decomposition& d, float* x) {
    IuwtDecomposeAliased(h_, x, d.w, d.h, d.n_scales, d.planes.get(), stream_);
  }
  // ApplyMask + Recompose(out, false)
  void RecomposeMasked(const Decomposition& d, const uint8_t* masks, float* out) {
    IuwtRecompose(h_, d.planes.get(), d.w, d.h, d.n_scales, false, masks, out, stream_);
  }
  // PrepareConvolutionKernel + the kernel's spectrum for a w x h frame
  const float2* KernelSpectrum(const float* psf, size_t w, size_t h, DeviceBuffer<float2>& store) {
    FftConvolver& conv = GetConvolver(h_, w, h);
    store.Reserve(conv.SpectrumSize());
    conv.MakePsfKernelSpectrum(psf, w, h, store.get(), stream_);
    h_.launches += conv.TakeLaunchCount();
    return store.get();
  }
  void Convolve(float* image, const float2* spectrum, size_t w, size_t h) {
    FftConvolver& conv = GetConvolver(h_, w, h);
    conv.ConvolveWith(image, w, h, spectrum, image, w, h, 0, stream_);
    h_.launches += conv.TakeLaunchCount();
  }

  void MeasureRMSPerScale(const float* psf, int end_scale);
  void BoundingBox(size_t& x1, size_t& y1, size_t& x2, size_t& y2, const float* image, size_t width,
                   size_t height);
  static void AdjustBox(size_t& x1, size_t& y1, size_t& x2, size_t& y2, size_t width, size_t height,
                        int end_scale);
  float Snr(const Decomposition& noisy, const Decomposition& model, const uint8_t* masks);
  bool RunConjugateGradient(Decomposition& iuwt, const uint8_t* masks, float* masked_dirty,
                            float* structure_model, float* scratch, const float2* psf_kernel,
                            size_t width, size_t height);
  bool FindAndDeconvolveStructure(Decomposition& iuwt, float* dirty, const float* psf,
                                  const float2* psf_kernel, float* scratch, size_t cur_end_scale,
                                  size_t cur_min_scale, std::vector<ValComponent>& max_components);
  bool FillAndDeconvolveStructure(Decomposition& iuwt, float* dirty, float* structure_model_full,
                                  float* scratch, const float* psf, const float2* psf_kernel,
                                  size_t cur_end_scale, size_t cur_min_scale, size_t width,
                                  size_t height, const std::vector<float>& thresholds,
                                  const ValComponent& max_comp, bool allow_trimming,
                                  const uint8_t* prior_mask);
  // cubes with several images: the structure found in the integrated image is refitted per image
  void PerformSubImageFitAll(Decomposition& iuwt, const uint8_t* masks, const float* structure_model,
                             float* scratch_a, float* scratch_b, const ValComponent& max_comp,
                             float* fitted_model, const float* psf, const float* dirty, size_t width,
                             size_t height);
  void PerformSubImageFitSingle(Decomposition& iuwt, const uint8_t* masks, const float* structure_model,
                                float* scratch_b, const ValComponent& max_comp, const float* psf,
                                float* sub_dirty, float* fitted_sub_model,
                                std::vector<float>& correction_factor, size_t width, size_t height);
  float PerformSubImageComponentFitBoxed(Decomposition& iuwt, const uint8_t* masks, float* model,
                                         const float* masked_dirty, const float* psf,
                                         const float2* psf_kernel, size_t x1, size_t y1, size_t x2,
                                         size_t y2, size_t width, size_t height);
  float PerformSubImageComponentFit(Decomposition& iuwt, const uint8_t* masks, float* model,
                                    const float* masked_dirty, const float2* psf_kernel, size_t x_offset,
                                    size_t y_offset, size_t width, size_t height, size_t area_width);

  rdl_handle& h_;
  DeviceCube& cube_;
  cudaStream_t stream_;
  size_t width_, height_;
  float minor_loop_gain_, major_loop_gain_, clean_border_, absolute_threshold_;
  const float threshold_sigma_level_ = 4.0f, tolerance_ = 0.75f;  // iuwt_deconvolution_algorithm.h:24-25
  bool allow_negative_;
  const uint8_t* mask_;
  std::vector<float> rmses_;
  std::vector<ScaleResponse> psf_response_;
  DeviceBuffer<double> partial_;
  DeviceBuffer<unsigned> minmax_;
  DeviceBuffer<int> row_first_, row_last_;
  DeviceBuffer<float> structure_model_;  // the (single) image of the structure model ImageSet
  // the trimmed recursion level of FillAndDeconvolveStructure
  Decomposition trimmed_iuwt_, initial_iuwt_;
  DeviceBuffer<float> small_dirty_, small_psf_, small_scratch_, small_model_, masked_dirty_,
      cg_model_, gradient_;
  DeviceBuffer<float2> small_spectrum_;
  DeviceBuffer<uint8_t> masks_full_, masks_small_, small_prior_;
  // PerformSubImageFit*: current bounding box in the full frame (cur_box_*), work planes
  size_t cur_box_x1_ = 0, cur_box_y1_ = 0, cur_box_x2_ = 0, cur_box_y2_ = 0;
  DeviceBuffer<uint8_t> mask2d_, mask2d_before_, area_, fit_masks_;
  DeviceBuffer<unsigned> fit_scalars_;
  DeviceBuffer<float> fit_sub_dirty_, fit_psf_, fit_model_, fit_masked_dirty_, fit_small_psf_;
  DeviceBuffer<float2> fit_spectrum_, fit_small_spectrum_;
  Decomposition fit_iuwt_;
};

// MeasureRMSPerScale (:42-102). The Gaussian fits and the flood fills at its end only produce
// b_major / b_minor / b_pa / convolved_area, which nothing reads: left out.
void IuwtRun::MeasureRMSPerScale(const float* psf, int end_scale) {
  const size_t n = width_ * height_;
  Decomposition d;
  d.Shape(end_scale, width_, height_);
  Decompose(d, psf);
  psf_response_.assign(end_scale, ScaleResponse{});
  const size_t centre = width_ / 2 + (height_ / 2) * width_;
  for (int s = 0; s != end_scale; ++s) {
    psf_response_[s].rms = Rms(d.Plane(s), n);
    psf_response_[s].peak_response = ReadPixel(d.Plane(s), centre);
  }
  // imageIUWT.Decompose(imageIUWT[1].Coefficients().Data(), scratch, false)
  DeviceBuffer<float> scale1(n);
  Copy(scale1.get(), d.Plane(1), n);
  Decompose(d, scale1.get());
  for (int s = 0; s != end_scale; ++s)
    psf_response_[s].peak_response_to_next_scale = ReadPixel(d.Plane(s), centre);
}

void IuwtRun::BoundingBox(size_t& x1, size_t& y1, size_t& x2, size_t& y2, const float* image,
                          size_t width, size_t height) {
  const size_t n = width * height;
  minmax_.Reserve(2);
  const unsigned init[2] = {0u, 0xffffffffu};
  RDL_CUDA(cudaMemcpyAsync(minmax_.get(), init, sizeof(init), cudaMemcpyHostToDevice, stream_));
  MinMaxKernel<<<GridFor(n), 256, 0, stream_>>>(image, n, minmax_.get());
  unsigned mm[2];
  RDL_CUDA(cudaMemcpyAsync(mm, minmax_.get(), sizeof(mm), cudaMemcpyDeviceToHost, stream_));
  Sync();
  const float mP = Unordered(mm[0]), mN = Unordered(mm[1]);
  const float m = std::max(mP, -mN);
  row_first_.Reserve(height);
  row_last_.Reserve(height);
  RowExtentKernel<<<unsigned(height), 256, 0, stream_>>>(image, int(width), int(height), double(m) * 0.01,
                                                         row_first_.get(), row_last_.get());
  h_.launches += 2;
  std::vector<int> first(height), last(height);
  RDL_CUDA(cudaMemcpyAsync(first.data(), row_first_.get(), height * sizeof(int), cudaMemcpyDeviceToHost, stream_));
  RDL_CUDA(cudaMemcpyAsync(last.data(), row_last_.get(), height * sizeof(int), cudaMemcpyDeviceToHost, stream_));
  Sync();
  // the reference's loops (:186-213), on the per-row extents
  x1 = width;
  x2 = 0;
  y1 = height;
  y2 = 0;
  for (size_t y = 0; y != height; ++y) {
    if (first[y] < int(width)) {
      if (size_t(first[y]) < x1) x1 = size_t(first[y]);
      // "for (x = width - 1; x != x2; --x)": only pixels to the right of x2 are looked at
      if (size_t(last[y]) > x2) x2 = size_t(last[y]);
    }
  }
  x2++;
  for (size_t y = 0; y != height; ++y) {
    if (first[y] < int(width)) {
      if (y1 > y) y1 = y;
      if (y2 < y) y2 = y + 1;  // sic: a row equal to the current y2 does not move it
    }
  }
}

void IuwtRun::AdjustBox(size_t& x1, size_t& y1, size_t& x2, size_t& y2, size_t width, size_t height,
                        int end_scale) {  // :216-262
  const int min_box_size = std::max<int>(128, int(MinImageDimension(end_scale)) * 3 / 2);
  const int box_width = int(x2) - int(x1);
  const int box_height = int(y2) - int(y1);
  int new_x1 = int(double(x1) - 0.5 * box_width);
  int new_x2 = int(double(x2) + 0.5 * box_width);
  int new_y1 = int(double(y1) - 0.5 * box_height);
  int new_y2 = int(double(y2) + 0.5 * box_height);
  if (new_x2 - new_x1 < min_box_size) {
    const int mid = int(0.5 * (int(x1) + int(x2)));
    new_x1 = mid - min_box_size / 2;
    new_x2 = mid + min_box_size / 2;
  }
  if (new_y2 - new_y1 < min_box_size) {
    const int mid = int(0.5 * (int(y1) + int(y2)));
    new_y1 = mid - min_box_size / 2;
    new_y2 = mid + min_box_size / 2;
  }
  x1 = new_x1 >= 0 ? size_t(new_x1) : 0;
  x2 = new_x2 < int(width) ? size_t(new_x2) : width;
  y1 = new_y1 >= 0 ? size_t(new_y1) : 0;
  y2 = new_y2 < int(height) ? size_t(new_y2) : height;
  while ((x2 - x1) % 8 != 0) x2--;
  while ((y2 - y1) % 8 != 0) y2--;
}

float IuwtRun::Snr(const Decomposition& noisy, const Decomposition& model, const uint8_t* masks) {
  const size_t n = size_t(noisy.n_scales) * noisy.w * noisy.h;
  partial_.Reserve(2 * kRedBlocks + 8);
  SnrKernel<<<kRedBlocks, kRedThreads, 0, stream_>>>(noisy.planes.get(), model.planes.get(), masks, n,
                                                     partial_.get());
  FinishSumKernel<<<1, 64, 0, stream_>>>(partial_.get(), kRedBlocks, 2, partial_.get() + 2 * kRedBlocks);
  h_.launches += 2;
  double v[2];
  RDL_CUDA(cudaMemcpyAsync(v, partial_.get() + 2 * kRedBlocks, sizeof(v), cudaMemcpyDeviceToHost, stream_));
  Sync();
  return float(v[0]) / float(v[1]);
}

// RunConjugateGradient (:326-407)
bool IuwtRun::RunConjugateGradient(Decomposition& iuwt, const uint8_t* masks, float* masked_dirty,
                                   float* structure_model, float* scratch, const float2* psf_kernel,
                                   size_t width, size_t height) {
  const size_t n = width * height;
  gradient_.Reserve(n);
  float* gradient = gradient_.get();
  Copy(gradient, masked_dirty, n);
  float model_snr = 0.0f;
  // IuwtDecomposition initialDirtyIUWT(iuwt): the masked decomposition of the dirty image
  initial_iuwt_.Shape(iuwt.n_scales, width, height);
  Copy(initial_iuwt_.planes.get(), iuwt.planes.get(), size_t(iuwt.n_scales) * n);

  for (size_t minor_iter = 0; minor_iter != 20; ++minor_iter) {
    // scratch = IUWT^-1 mask IUWT (gradient (x) psf)
    Copy(scratch, gradient, n);
    Convolve(scratch, psf_kernel, width, height);
    DecomposeAliased(iuwt, scratch);
    RecomposeMasked(iuwt, masks, scratch);

    const float gradient_dot_scratch = DotProduct(gradient, scratch, n);
    if (gradient_dot_scratch == 0.0f) return false;
    const float dirty_dot = DotProduct(masked_dirty, masked_dirty, n);
    const float step_size = dirty_dot / gradient_dot_scratch;
    Axpy(structure_model, gradient, step_size, n);
    const float grad_step_den = dirty_dot;
    if (grad_step_den == 0.0f) return false;
    Axpy(masked_dirty, scratch, -step_size, n);
    const float grad_step = DotProduct(masked_dirty, masked_dirty, n) / grad_step_den;
    GradientKernel<<<GridFor(n), 256, 0, stream_>>>(gradient, masked_dirty, grad_step, n);
    ++h_.launches;

    // mask IUWT (PSF (x) model)
    Copy(scratch, structure_model, n);
    Convolve(scratch, psf_kernel, width, height);
    DecomposeAliased(iuwt, scratch);

    const float previous_snr = model_snr;
    model_snr = Snr(iuwt, initial_iuwt_, masks);
    if (model_snr > 100 && minor_iter > 2) return true;
    if (model_snr < previous_snr && minor_iter > 5 && model_snr > 3) return true;
  }
  if (model_snr <= 3.0f) {
    Zero(structure_model, n);
    return false;
  }
  return true;
}

// FindAndDeconvolveStructure (:414-498)
bool IuwtRun::FindAndDeconvolveStructure(Decomposition& iuwt, float* dirty, const float* psf,
                                         const float2* psf_kernel, float* scratch,
                                         size_t cur_end_scale, size_t cur_min_scale,
                                         std::vector<ValComponent>& max_components) {
  const size_t n = width_ * height_;
  Decompose(iuwt, dirty);
  std::vector<float> thresholds(cur_end_scale);
  rmses_.resize(cur_end_scale);
  for (size_t scale = 0; scale != cur_end_scale; ++scale) {
    const float r = IuwtMad(h_, iuwt.Plane(int(scale)), n, stream_);
    rmses_[scale] = r;
    thresholds[scale] = float(r * (threshold_sigma_level_ * 4.0 / 5.0));
  }
  max_components.resize(cur_end_scale);
  for (size_t scale = 0; scale != cur_end_scale; ++scale) {
    size_t x = 0, y = 0;
    const float max_abs = IuwtMaxAbs(h_, iuwt.Plane(int(scale)), width_, height_, clean_border_,
                                     allow_negative_, mask_, x, y, stream_);
    max_components[scale].x = x;
    max_components[scale].y = y;
    max_components[scale].scale = int(scale);
    max_components[scale].val = max_abs;
  }

  float max_val = -1.0f;
  size_t max_x = 0, max_y = 0;
  int max_val_scale = -1;
  for (size_t scale = 0; scale != cur_end_scale; ++scale) {
    const ValComponent& val = max_components[scale];
    const float abs_coef = val.val / psf_response_[scale].rms;
    if (scale >= cur_min_scale && abs_coef > max_val &&
        val.val > rmses_[scale] * threshold_sigma_level_ &&
        val.val > rmses_[scale] / rmses_[0] * absolute_threshold_) {
      max_x = val.x;
      max_y = val.y;
      max_val_scale = int(scale);
      if (scale == 0) {
        const float lowest_rms = std::min(psf_response_[0].rms, psf_response_[1].rms);
        max_val = val.val / lowest_rms * psf_response_[1].peak_response /
                  psf_response_[0].peak_response_to_next_scale;
      } else {
        max_val = abs_coef;
      }
    }
  }
  if (max_val_scale == -1) return false;  // "No significant pixel found."

  max_val = ReadPixel(iuwt.Plane(max_val_scale), max_x + max_y * width_);
  if (std::fabs(max_val) < thresholds[max_val_scale]) return false;  // "... is in the noise"

  const float scale_max_abs_val = std::fabs(max_val);
  for (size_t scale = 0; scale != cur_end_scale; ++scale) {
    if (thresholds[scale] < tolerance_ * scale_max_abs_val) thresholds[scale] = tolerance_ * scale_max_abs_val;
    if (max_val < 0.0f) thresholds[scale] = -thresholds[scale];
  }
  ValComponent max_comp;
  max_comp.x = max_x;
  max_comp.y = max_y;
  max_comp.scale = max_val_scale;
  structure_model_.Reserve(n);
  return FillAndDeconvolveStructure(iuwt, dirty, structure_model_.get(), scratch, psf, psf_kernel,
                                    cur_end_scale, cur_min_scale, width_, height_, thresholds,
                                    max_comp, true, mask_);
}

// FillAndDeconvolveStructure (:500-626)
bool IuwtRun::FillAndDeconvolveStructure(Decomposition& iuwt, float* dirty, float* structure_model_full,
                                         float* scratch, const float* psf, const float2* psf_kernel,
                                         size_t cur_end_scale, size_t cur_min_scale, size_t width,
                                         size_t height, const std::vector<float>& thresholds,
                                         const ValComponent& max_comp, bool allow_trimming,
                                         const uint8_t* prior_mask) {
  const size_t n = width * height;
  DeviceBuffer<uint8_t>& mask_store = allow_trimming ? masks_full_ : masks_small_;
  mask_store.Reserve(cur_end_scale * n);
  uint8_t* masks = mask_store.get();
  RDL_CUDA(cudaMemsetAsync(masks, 0, cur_end_scale * n, stream_));
  IuwtSelectStructures(h_, iuwt.planes.get(), masks, width, height, int(cur_end_scale),
                       thresholds.data(), int(cur_min_scale), int(cur_end_scale), clean_border_,
                       prior_mask, stream_);
  // iuwt.ApplyMask(mask); iuwt.Recompose(scratch, false)
  Decomposition view;
  view.n_scales = int(cur_end_scale);

  size_t x1 = 0, y1 = 0, x2 = width, y2 = height;
  if (allow_trimming) {
    IuwtRecompose(h_, iuwt.planes.get(), width, height, int(cur_end_scale), false, masks, scratch, stream_);
    BoundingBox(x1, y1, x2, y2, scratch, width, height);
    AdjustBox(x1, y1, x2, y2, width, height, max_comp.scale + 1);
  }
  if (allow_trimming && ((x2 - x1) < width || (y2 - y1) < height)) {
    const size_t new_width = x2 - x1, new_height = y2 - y1, small_n = new_width * new_height;
    if (x2 <= x1 || y2 <= y1 || new_width < 2 || new_height < 2)
      throw std::domain_error("IUWT bounding box is degenerate");
    small_dirty_.Reserve(small_n);
    TrimKernel<<<GridFor(small_n), 256, 0, stream_>>>(small_dirty_.get(), dirty, int(width), int(x1),
                                                      int(y1), int(new_width), int(new_height), nullptr);
    // TrimPsf: the centred new_width x new_height part of the PSF
    small_psf_.Reserve(small_n);
    TrimKernel<<<GridFor(small_n), 256, 0, stream_>>>(
        small_psf_.get(), psf, int(width), int((width - new_width) / 2), int((height - new_height) / 2),
        int(new_width), int(new_height), nullptr);
    h_.launches += 2;
    const float2* small_kernel = KernelSpectrum(small_psf_.get(), new_width, new_height, small_spectrum_);
    small_scratch_.Reserve(small_n);

    const int max_scale = std::max(EndScale(std::min(new_width, new_height)), max_comp.scale + 1);
    if (max_scale < int(cur_end_scale)) cur_end_scale = size_t(max_scale);
    // iuwt.CreateTrimmed: the box of the (masked) coefficient planes
    trimmed_iuwt_.Shape(int(cur_end_scale), new_width, new_height);
    for (size_t s = 0; s != cur_end_scale; ++s) {
      TrimKernel<<<GridFor(small_n), 256, 0, stream_>>>(trimmed_iuwt_.Plane(int(s)), iuwt.Plane(int(s)),
                                                        int(width), int(x1), int(y1), int(new_width),
                                                        int(new_height), masks + s * n);
      ++h_.launches;
    }
    // structure_model_full.Trim(...): every image of the set
    cur_box_x1_ = x1;
    cur_box_x2_ = x2;
    cur_box_y1_ = y1;
    cur_box_y2_ = y2;
    const size_t n_img = cube_.n_images;
    small_model_.Reserve(n_img * small_n);
    for (size_t i = 0; i != n_img; ++i)
      TrimKernel<<<GridFor(small_n), 256, 0, stream_>>>(small_model_.get() + i * small_n,
                                                        structure_model_full + i * n, int(width), int(x1),
                                                        int(y1), int(new_width), int(new_height), nullptr);
    h_.launches += n_img - 1;
    const uint8_t* small_prior = nullptr;
    if (prior_mask) {
      small_prior_.Reserve(small_n);
      TrimBytesKernel<<<GridFor(small_n), 256, 0, stream_>>>(small_prior_.get(), prior_mask, int(width),
                                                             int(x1), int(y1), int(new_width),
                                                             int(new_height));
      small_prior = small_prior_.get();
    }
    h_.launches += 2;
    ValComponent new_max = max_comp;
    new_max.x = max_comp.x - x1;
    new_max.y = max_comp.y - y1;
    const bool result = FillAndDeconvolveStructure(
        trimmed_iuwt_, small_dirty_.get(), small_model_.get(), small_scratch_.get(), small_psf_.get(),
        small_kernel, cur_end_scale, cur_min_scale, new_width, new_height, thresholds, new_max, false,
        small_prior);
    for (size_t i = 0; i != n_img; ++i)
      UntrimKernel<<<GridFor(n), 256, 0, stream_>>>(structure_model_full + i * n, int(width), int(height),
                                                    small_model_.get() + i * small_n, int(x1), int(y1),
                                                    int(new_width), int(new_height));
    h_.launches += n_img;
    cur_box_x1_ = 0;
    cur_box_x2_ = width;
    cur_box_y1_ = 0;
    cur_box_y2_ = height;
    RDL_CUDA(cudaGetLastError());
    return result;
  }

  // get undeconvolved dirty
  iuwt.Shape(int(cur_end_scale), width, height);
  Decompose(iuwt, dirty);
  masked_dirty_.Reserve(n);
  RecomposeMasked(iuwt, masks, scratch);
  Copy(masked_dirty_.get(), scratch, n);
  // `iuwt` now stands for the MASKED decomposition (ApplyMask): RunConjugateGradient copies it as
  // its reference point and Snr only looks inside the mask

  cg_model_.Reserve(n);
  Zero(cg_model_.get(), n);
  const bool success = RunConjugateGradient(iuwt, masks, masked_dirty_.get(), cg_model_.get(), scratch,
                                            psf_kernel, width, height);
  if (!success) return false;

  const float rms_before = Rms(dirty, n);
  Copy(scratch, cg_model_.get(), n);
  Convolve(scratch, psf_kernel, width, height);
  Copy(masked_dirty_.get(), dirty, n);  // maskedDirty as temporary
  Axpy(masked_dirty_.get(), scratch, -minor_loop_gain_, n);
  const float rms_after = Rms(masked_dirty_.get(), n);
  if (rms_after > rms_before) return false;  // "RMS got worse"

  if (cube_.n_images == 1) {
    // PerformSubImageFitAll with a single image: fitted_model.SetImage(0, structure_model)
    Copy(structure_model_full, cg_model_.get(), n);
  } else {
    PerformSubImageFitAll(iuwt, masks, cg_model_.get(), scratch, masked_dirty_.get(), max_comp,
                          structure_model_full, psf, dirty, width, height);
  }
  RDL_CUDA(cudaGetLastError());
  return true;
}

// PerformSubImageFitAll (:628-670): with several images the structure model found in the
// integrated image is refitted image by image. First the per-component correction factors of the
// integrated image itself, then every image of the set (trimmed to the current box, with its own
// PSF): fitted[img] += component * factor(img) / factor(integrated).
void IuwtRun::PerformSubImageFitAll(Decomposition& iuwt, const uint8_t* masks,
                                    const float* structure_model, float* scratch_a, float* scratch_b,
                                    const ValComponent& max_comp, float* fitted_model, const float* psf,
                                    const float* dirty, size_t width, size_t height) {
  const size_t n = width * height;
  std::vector<float> correction_factors;
  Copy(scratch_a, dirty, n);
  PerformSubImageFitSingle(iuwt, masks, structure_model, scratch_b, max_comp, psf, scratch_a, nullptr,
                           correction_factors, width, height);
  Zero(fitted_model, cube_.n_images * n);
  const size_t bw = cur_box_x2_ - cur_box_x1_, bh = cur_box_y2_ - cur_box_y1_;
  if (bw != width || bh != height) throw std::logic_error("IUWT box does not match the sub-image");
  for (size_t img = 0; img != cube_.n_images; ++img) {
    TrimKernel<<<GridFor(n), 256, 0, stream_>>>(scratch_a, cube_.residual.p[img], int(width_),
                                                int(cur_box_x1_), int(cur_box_y1_), int(bw), int(bh),
                                                nullptr);
    ++h_.launches;
    const float* sub_psf = cube_.psf.p[cube_.PsfIndex(img)];
    if (width_ != width || height_ != height) {
      fit_psf_.Reserve(n);
      TrimKernel<<<GridFor(n), 256, 0, stream_>>>(fit_psf_.get(), sub_psf, int(width_),
                                                  int((width_ - width) / 2), int((height_ - height) / 2),
                                                  int(width), int(height), nullptr);
      ++h_.launches;
      sub_psf = fit_psf_.get();
    }
    PerformSubImageFitSingle(iuwt, masks, structure_model, scratch_b, max_comp, sub_psf, scratch_a,
                             fitted_model + img * n, correction_factors, width, height);
  }
}

// PerformSubImageFitSingle (:672-742)
void IuwtRun::PerformSubImageFitSingle(Decomposition& iuwt, const uint8_t* masks,
                                       const float* structure_model, float* scratch_b,
                                       const ValComponent& max_comp, const float* psf, float* sub_dirty,
                                       float* fitted_sub_model, std::vector<float>& correction_factor,
                                       size_t width, size_t height) {
  const size_t n = width * height;
  const float2* psf_kernel = KernelSpectrum(psf, width, height, fit_spectrum_);
  float* masked_dirty = scratch_b;
  DecomposeAliased(iuwt, sub_dirty);  // iuwt.Decompose(sub_dirty, sub_dirty, false)
  RecomposeMasked(iuwt, masks, masked_dirty);
  mask2d_.Reserve(n);
  mask2d_before_.Reserve(n);
  area_.Reserve(n);
  fit_scalars_.Reserve(8);
  RDL_CUDA(cudaMemsetAsync(mask2d_.get(), 0, n, stream_));
  const float peak_level = std::fabs(ReadPixel(structure_model, size_t(max_comp.y) * width + max_comp.x));
  const float threshold = peak_level * 1e-4f;
  size_t component_index = 0;
  for (;;) {
    // next seed in row-major order
    const unsigned none = 0xffffffffu;
    RDL_CUDA(cudaMemcpyAsync(fit_scalars_.get(), &none, sizeof(none), cudaMemcpyHostToDevice, stream_));
    FindSeedKernel<<<GridFor(n), 256, 0, stream_>>>(structure_model, mask2d_.get(), n, threshold,
                                                    fit_scalars_.get());
    ++h_.launches;
    unsigned seed = none;
    RDL_CUDA(cudaMemcpyAsync(&seed, fit_scalars_.get(), sizeof(seed), cudaMemcpyDeviceToHost, stream_));
    Sync();
    if (seed == none) break;
    RDL_CUDA(cudaMemcpyAsync(mask2d_before_.get(), mask2d_.get(), n, cudaMemcpyDeviceToDevice, stream_));
    IuwtFloodfill2D(h_, structure_model, mask2d_.get(), width, height, threshold, true, seed % width,
                    seed / width, stream_);
    // the area, the model inside it, its bounding box
    const unsigned box_init[4] = {0xffffffffu, 0u, 0xffffffffu, 0u};
    RDL_CUDA(cudaMemcpyAsync(fit_scalars_.get() + 4, box_init, sizeof(box_init), cudaMemcpyHostToDevice,
                             stream_));
    AreaKernel<<<GridFor(n), 256, 0, stream_>>>(structure_model, mask2d_.get(), mask2d_before_.get(),
                                                int(width), n, area_.get(), sub_dirty,
                                                fit_scalars_.get() + 4);
    ++h_.launches;
    unsigned box[4];
    RDL_CUDA(cudaMemcpyAsync(box, fit_scalars_.get() + 4, sizeof(box), cudaMemcpyDeviceToHost, stream_));
    Sync();
    size_t box_x1 = box[0], box_x2 = box[1], box_y1 = box[2], box_y2 = box[3];  // maxima inclusive (sic)
    AdjustBox(box_x1, box_y1, box_x2, box_y2, width, height, iuwt.n_scales);
    const float factor = PerformSubImageComponentFitBoxed(iuwt, masks, sub_dirty, masked_dirty, psf,
                                                          psf_kernel, box_x1, box_y1, box_x2, box_y2,
                                                          width, height);
    if (fitted_sub_model != nullptr) {
      const float integrated = correction_factor.at(component_index);
      if (std::isfinite(factor) && std::isfinite(integrated) && integrated != 0.0f) {
        FittedAddKernel<<<GridFor(n), 256, 0, stream_>>>(fitted_sub_model, structure_model, area_.get(),
                                                         factor, integrated, n);
        ++h_.launches;
      }
      ++component_index;
    } else {
      correction_factor.push_back(factor);
    }
  }
  RDL_CUDA(cudaGetLastError());
}

// PerformSubImageComponentFitBoxed (:744-777)
float IuwtRun::PerformSubImageComponentFitBoxed(Decomposition& iuwt, const uint8_t* masks, float* model,
                                                const float* masked_dirty, const float* psf,
                                                const float2* psf_kernel, size_t x1, size_t y1,
                                                size_t x2, size_t y2, size_t width, size_t height) {
  if (x1 > 0 || y1 > 0 || x2 < width || y2 < height) {
    const size_t new_width = x2 - x1, new_height = y2 - y1, small_n = new_width * new_height;
    const size_t n = width * height;
    fit_iuwt_.Shape(iuwt.n_scales, new_width, new_height);
    fit_masks_.Reserve(size_t(iuwt.n_scales) * small_n);
    for (int s = 0; s != iuwt.n_scales; ++s)
      TrimBytesKernel<<<GridFor(small_n), 256, 0, stream_>>>(fit_masks_.get() + size_t(s) * small_n,
                                                             masks + size_t(s) * n, int(width), int(x1),
                                                             int(y1), int(new_width), int(new_height));
    fit_model_.Reserve(small_n);
    TrimKernel<<<GridFor(small_n), 256, 0, stream_>>>(fit_model_.get(), model, int(width), int(x1), int(y1),
                                                      int(new_width), int(new_height), nullptr);
    fit_small_psf_.Reserve(small_n);
    TrimKernel<<<GridFor(small_n), 256, 0, stream_>>>(fit_small_psf_.get(), psf, int(width),
                                                      int((width - new_width) / 2),
                                                      int((height - new_height) / 2), int(new_width),
                                                      int(new_height), nullptr);
    const float2* small_kernel = KernelSpectrum(fit_small_psf_.get(), new_width, new_height,
                                                fit_small_spectrum_);
    fit_masked_dirty_.Reserve(small_n);
    TrimKernel<<<GridFor(small_n), 256, 0, stream_>>>(fit_masked_dirty_.get(), masked_dirty, int(width),
                                                      int(x1), int(y1), int(new_width), int(new_height),
                                                      nullptr);
    h_.launches += 3 + iuwt.n_scales;
    return PerformSubImageComponentFit(fit_iuwt_, fit_masks_.get(), fit_model_.get(),
                                       fit_masked_dirty_.get(), small_kernel, x1, y1, new_width,
                                       new_height, width);
  }
  return PerformSubImageComponentFit(iuwt, masks, model, masked_dirty, psf_kernel, 0, 0, width, height,
                                     width);
}

// PerformSubImageComponentFit (:779-801): IUWT^-1 mask IUWT (model (x) PSF), then the ratio of the
// sums of masked dirty and model over the component's area. (float32 sequential sums over the
// area in flood-fill order in the reference; fixed-order double sums here.)
float IuwtRun::PerformSubImageComponentFit(Decomposition& iuwt, const uint8_t* masks, float* model,
                                           const float* masked_dirty, const float2* psf_kernel,
                                           size_t x_offset, size_t y_offset, size_t width,
                                           size_t height, size_t area_width) {
  Convolve(model, psf_kernel, width, height);
  DecomposeAliased(iuwt, model);
  RecomposeMasked(iuwt, masks, model);
  partial_.Reserve(2 * kRedBlocks + 8);
  AreaSumsKernel<<<kRedBlocks, kRedThreads, 0, stream_>>>(model, masked_dirty, area_.get(), int(area_width),
                                                          int(x_offset), int(y_offset), int(width),
                                                          int(height), partial_.get());
  FinishSumKernel<<<1, 64, 0, stream_>>>(partial_.get(), kRedBlocks, 2, partial_.get() + 2 * kRedBlocks);
  h_.launches += 2;
  double sums[2];
  RDL_CUDA(cudaMemcpyAsync(sums, partial_.get() + 2 * kRedBlocks, sizeof(sums), cudaMemcpyDeviceToHost,
                           stream_));
  Sync();
  const float model_sum = float(sums[0]), dirty_sum = float(sums[1]);
  if (model_sum == 0.0f || !std::isfinite(dirty_sum) || !std::isfinite(model_sum)) return 0.0f;
  return dirty_sum / model_sum;
}

// PerformMajorIteration (:803-918)
float IuwtRun::PerformMajorIteration(uint64_t& iter_counter, uint64_t n_iter,
                                     bool& reached_major_threshold) {
  reached_major_threshold = false;
  if (iter_counter == n_iter) return 0.0f;
  const size_t n = width_ * height_;
  DeviceBuffer<float> dirty(n), psf(n), scratch(n), dirty_before(n);
  ConstPtrTable residual{}, psf_table{};
  for (size_t i = 0; i != cube_.n_images; ++i) residual.p[i] = cube_.residual.p[i];
  for (size_t p = 0; p != cube_.n_channels; ++p) psf_table.p[p] = cube_.psf.p[p];
  const IntegSpec linear = MakeLinearSpec(cube_);
  LaunchIntegrate(linear, residual, n, dirty.get(), cube_.vec, stream_);
  LaunchIntegrate(MakeIntegratedPsfSpec(cube_), psf_table, n, psf.get(), cube_.vec, stream_);
  h_.launches += 2;

  const int max_scale = EndScale(std::min(width_, height_));
  int cur_end_scale = 2;
  DeviceBuffer<float2> psf_spectrum, image_psf_spectrum;
  const float2* psf_kernel = KernelSpectrum(psf.get(), width_, height_, psf_spectrum);
  // "Measuring PSF..."; the PSF convolved with itself only feeds the dead outputs
  MeasureRMSPerScale(psf.get(), max_scale);

  const size_t n_img = cube_.n_images;
  structure_model_.Reserve(n_img * n);
  Zero(structure_model_.get(), n_img * n);
  cur_box_x1_ = 0;
  cur_box_x2_ = width_;
  cur_box_y1_ = 0;
  cur_box_y2_ = height_;
  Decomposition iuwt;
  iuwt.Shape(max_scale, width_, height_);  // room for every scale count this call can reach
  iuwt.n_scales = cur_end_scale;

  float max_value = 0.0f;
  size_t cur_min_scale = 0;
  bool do_continue = true;
  std::vector<ValComponent> initial_components;
  do {
    Copy(dirty_before.get(), dirty.get(), n);
    std::vector<ValComponent> max_components;
    iuwt.n_scales = cur_end_scale;
    const bool succeeded = FindAndDeconvolveStructure(iuwt, dirty.get(), psf.get(), psf_kernel,
                                                      scratch.get(), size_t(cur_end_scale),
                                                      cur_min_scale, max_components);
    if (succeeded) {
      ScaleKernel<<<GridFor(n_img * n), 256, 0, stream_>>>(structure_model_.get(), minor_loop_gain_,
                                                           n_img * n);
      size_t spectrum_of = size_t(-1);
      const float2* kernel = nullptr;
      for (size_t i = 0; i != n_img; ++i) {
        float* part = structure_model_.get() + i * n;
        AddSubKernel<<<GridFor(n), 256, 0, stream_>>>(cube_.model.p[i], part, true, n);
        // dirty = dirty - structureModel (x) psf, per image with its own PSF
        Copy(scratch.get(), part, n);
        if (cube_.PsfIndex(i) != spectrum_of) {
          spectrum_of = cube_.PsfIndex(i);
          kernel = KernelSpectrum(cube_.psf.p[spectrum_of], width_, height_, image_psf_spectrum);
        }
        Convolve(scratch.get(), kernel, width_, height_);
        AddSubKernel<<<GridFor(n), 256, 0, stream_>>>(cube_.residual.p[i], scratch.get(), false, n);
        h_.launches += 2;
      }
      LaunchIntegrate(linear, residual, n, dirty.get(), cube_.vec, stream_);
      h_.launches += 2;

      while (max_components.size() > initial_components.size())
        initial_components.push_back(max_components[initial_components.size()]);
      max_value = 0.0f;
      for (size_t c = 0; c != initial_components.size(); ++c) {
        max_value = std::max(max_value, max_components[c].val);
        if (std::fabs(max_components[c].val) <
            std::fabs(initial_components[c].val) * (1.0 - major_loop_gain_))
          reached_major_threshold = true;
      }
      if (reached_major_threshold) break;
    } else {
      if (int(cur_min_scale) + 1 < cur_end_scale) {
        ++cur_min_scale;
      } else {
        cur_min_scale = 0;
        if (cur_end_scale != max_scale)
          ++cur_end_scale;
        else
          do_continue = false;  // "Max scale reached: finished all scales, quiting."
      }
      Copy(dirty.get(), dirty_before.get(), n);
    }
    ++iter_counter;
  } while (iter_counter != n_iter && do_continue);
  Sync();
  RDL_CUDA(cudaGetLastError());
  return max_value;
}

}  // namespace

// IuwtDeconvolution::ExecuteMajorIteration (iuwt_deconvolution.h:21-38)
rdl_result RunIuwt(rdl_handle& h, DeviceCube& cube, cudaStream_t stream) {
  h.trace_host.clear();
  IuwtRun run(h, cube, stream);
  uint64_t iteration_number = h.iteration_number;
  bool reached = false;
  rdl_result r{};
  r.final_peak_value = run.PerformMajorIteration(iteration_number, h.settings.max_iterations, reached);
  r.another_iteration_required = reached ? 1 : 0;
  h.iteration_number = iteration_number;
  if (h.iteration_number >= h.settings.max_iterations) r.another_iteration_required = 0;
  r.iteration_number = h.iteration_number;
  r.has_starting_peak = 0;
  return r;
}

}  // namespace rdl
