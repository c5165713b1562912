// The steps of Radler::Perform either side of ExecuteMajorIteration, on the device (SURVEY 8(f),
// rank 1), so that a caller holding the cube in device memory need not bring a plane back for them:
//   radler.cc:162-169   integrated.MedianAndStdDevFromMAD()  -> ImageMedianAndStdDevFromMad
//   image_set.cc:105-140 ImageSet::LoadAndAverage            -> ImageLoadAndAverage
//   image_set.cc:142-205 ImageSet::LoadAndAveragePsfs        -> ImageLoadAndAverage (psf_rule)
//   radler.cc:254-266   auto-mask from the model images      -> ImageMaskFromModel
//   image_set.cc:208-283 ImageSet::InterpolateAndStoreModel  -> ImageInterpolateModel
// MedianAndStdDevFromMAD is aocommon's (third-party, absent from the tree): restated as the
// median over the finite pixels ((lower + upper middle) * 0.5 for an even count), then
// 1.4826 x the same median of |pixel - median| -- oracle/ref_shim/aocommon/image.h holds the same
// restatement; that part of the parity is UNPINNED. Order statistics have one right answer, so
// the radix select below returns the very float std::nth_element leaves in place.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <stdexcept>
#include <vector>

#include "handle.h"
#include "image_statistics.cuh"
#include "iuwt.cuh"

namespace rdl {
namespace {

// ---- exact selection: three histogram passes over the 32 key bits (11 + 11 + 10) -------------
// A median needs the upper middle element and, for an even count, the lower middle one: both
// ranks are narrowed together. They share one histogram until they fall into different bins
// (adjacent order statistics rarely do before the last pass); from then on a pixel is counted in
// the histogram of whichever prefix it matches.

// h.iuwt_scalars words used here (the IUWT entry points use the same scratch; calls on one
// handle are ordered): [2] / [6] key prefix found so far for the lower / upper rank, [3] / [7]
// rank left inside it, [5] finite pixels; the two histograms (2048 unsigned each) sit at
// kIuwtHistWord and behind it.
constexpr int kPrefixLo = 2, kRankLo = 3, kPrefixHi = 6, kRankHi = 7;
constexpr int kCountWord = 5;
constexpr int kHistogramBins = 2048;
constexpr int kSelectWords = kIuwtHistWord + 2 * kHistogramBins / 2;

struct SelectPass {
  int shift, bits, prefix_shift;  // bin = (key >> shift) & mask; prefix = key >> prefix_shift
  int deviation;                  // 0: key orders the pixel itself; 1: |pixel - centre|
  float centre;
};

// larger float <=> larger unsigned
__device__ __forceinline__ unsigned OrderedBits(float v) {
  const unsigned u = __float_as_uint(v);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

__device__ __forceinline__ void Count(float v, const SelectPass& pass, unsigned prefix_lo,
                                      unsigned prefix_hi, unsigned mask, unsigned* bins) {
  if ((__float_as_uint(v) & 0x7f800000u) == 0x7f800000u) return;  // std::isfinite
  const unsigned key = pass.deviation ? __float_as_uint(fabsf(v - pass.centre)) : OrderedBits(v);
  const unsigned bin = (key >> pass.shift) & mask;
  if (pass.prefix_shift >= 32) {
    atomicAdd(&bins[bin], 1u);
    return;
  }
  const unsigned prefix = key >> pass.prefix_shift;
  if (prefix == prefix_lo)
    atomicAdd(&bins[bin], 1u);
  else if (prefix == prefix_hi)
    atomicAdd(&bins[kHistogramBins + bin], 1u);
}

__global__ void __launch_bounds__(512)
SelectHistogramKernel(const float* __restrict__ image, size_t n, SelectPass pass,
                      const unsigned long long* __restrict__ scalars, unsigned* histogram) {
  __shared__ unsigned bins[2 * kHistogramBins];
  for (int i = threadIdx.x; i < 2 * kHistogramBins; i += blockDim.x) bins[i] = 0;
  __syncthreads();
  const unsigned prefix_lo = unsigned(scalars[kPrefixLo]);
  const unsigned prefix_hi = unsigned(scalars[kPrefixHi]);
  const unsigned mask = (1u << pass.bits) - 1u;
  const size_t stride = size_t(gridDim.x) * blockDim.x;
  const size_t first = size_t(blockIdx.x) * blockDim.x + threadIdx.x;
  if ((reinterpret_cast<uintptr_t>(image) & 15) == 0) {
    const float4* image4 = reinterpret_cast<const float4*>(image);
    for (size_t i = first; i < n / 4; i += stride) {
      const float4 v = __ldg(image4 + i);
      Count(v.x, pass, prefix_lo, prefix_hi, mask, bins);
      Count(v.y, pass, prefix_lo, prefix_hi, mask, bins);
      Count(v.z, pass, prefix_lo, prefix_hi, mask, bins);
      Count(v.w, pass, prefix_lo, prefix_hi, mask, bins);
    }
    for (size_t i = n / 4 * 4 + first; i < n; i += stride)
      Count(image[i], pass, prefix_lo, prefix_hi, mask, bins);
  } else {
    for (size_t i = first; i < n; i += stride) Count(image[i], pass, prefix_lo, prefix_hi, mask, bins);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2 * kHistogramBins; i += blockDim.x)
    if (bins[i]) atomicAdd(&histogram[i], bins[i]);
}

// Two warps, one per rank: each finds the bin holding its rank (in the shared histogram while
// the prefixes are equal), appends it to its prefix; then both histograms are cleared. With
// middle_ranks set (first pass of the first selection) the ranks are (count - 1) / 2 and
// count / 2 of the pixels counted, and the count is kept for the host.
__global__ void __launch_bounds__(64)
SelectNarrowKernel(unsigned long long* scalars, unsigned* histogram, int bits, int middle_ranks) {
  __shared__ unsigned long long chunk_sum[2][32];
  const int which = threadIdx.x / 32, lane = threadIdx.x % 32;
  const unsigned long long prefix_lo = scalars[kPrefixLo], prefix_hi = scalars[kPrefixHi];
  const unsigned* bins = histogram + ((which == 1 && prefix_hi != prefix_lo) ? kHistogramBins : 0);
  const int n_bins = 1 << bits, per = (n_bins + 31) / 32;
  unsigned long long sum = 0;
  for (int b = lane * per; b < min(n_bins, (lane + 1) * per); ++b) sum += bins[b];
  chunk_sum[which][lane] = sum;
  __syncthreads();
  if (lane == 0) {
    unsigned long long rank = scalars[which ? kRankHi : kRankLo];
    if (middle_ranks) {
      unsigned long long total = 0;
      for (int c = 0; c < 32; ++c) total += chunk_sum[which][c];
      if (which == 0) scalars[kCountWord] = total;
      rank = which == 0 ? (total ? (total - 1) / 2 : 0) : total / 2;
    }
    int c = 0;
    while (c < 31 && rank >= chunk_sum[which][c]) rank -= chunk_sum[which][c++];
    int b = c * per;
    while (b < n_bins - 1 && rank >= bins[b]) rank -= bins[b++];
    scalars[which ? kPrefixHi : kPrefixLo] = ((which ? prefix_hi : prefix_lo) << bits) | unsigned(b);
    scalars[which ? kRankHi : kRankLo] = rank;
  }
  __syncthreads();
  for (int b = threadIdx.x; b < 2 * kHistogramBins; b += 64) histogram[b] = 0;
}

struct Selected {
  unsigned key_lo, key_hi;  // of ranks (count - 1) / 2 and count / 2
  uint64_t n_finite;
};

// The keys of the two middle ranks among the finite pixels; n_finite = 0: counted here.
Selected SelectMiddle(rdl_handle& h, const float* image, size_t n, int deviation, float centre,
                      uint64_t n_finite, cudaStream_t stream) {
  h.iuwt_scalars.Reserve(kSelectWords);
  unsigned long long* scalars = h.iuwt_scalars.get();
  unsigned* histogram = reinterpret_cast<unsigned*>(scalars + kIuwtHistWord);
  RDL_CUDA(cudaMemsetAsync(scalars, 0, kSelectWords * sizeof(unsigned long long), stream));
  if (n_finite) {
    unsigned long long init[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    init[kRankLo] = (n_finite - 1) / 2;
    init[kRankHi] = n_finite / 2;
    init[kCountWord] = n_finite;
    RDL_CUDA(cudaMemcpyAsync(scalars, init, sizeof(init), cudaMemcpyHostToDevice, stream));
  }
  const SelectPass passes[3] = {{21, 11, 32, deviation, centre},
                                {10, 11, 21, deviation, centre},
                                {0, 10, 10, deviation, centre}};
  const unsigned grid = unsigned(std::min<size_t>((n / 4 + 511) / 512 + 1, size_t(kSmCount) * 4));
  for (int p = 0; p != 3; ++p) {
    SelectHistogramKernel<<<grid, 512, 0, stream>>>(image, n, passes[p], scalars, histogram);
    SelectNarrowKernel<<<1, 64, 0, stream>>>(scalars, histogram, passes[p].bits,
                                             p == 0 && !n_finite);
  }
  RDL_CUDA(cudaGetLastError());
  unsigned long long back[8];
  RDL_CUDA(cudaMemcpyAsync(back, scalars, sizeof(back), cudaMemcpyDeviceToHost, stream));
  RDL_CUDA(cudaStreamSynchronize(stream));
  return {unsigned(back[kPrefixLo]), unsigned(back[kPrefixHi]), back[kCountWord]};
}

float FromBits(unsigned bits) {
  float v;
  static_assert(sizeof(v) == sizeof(bits), "float is 32 bits");
  std::memcpy(&v, &bits, sizeof(v));
  return v;
}
float FromOrderedBits(unsigned key) {
  return FromBits((key & 0x80000000u) ? (key ^ 0x80000000u) : ~key);
}

// median of the finite pixels (deviation = 0) or of |pixel - centre| (deviation = 1)
float Median(rdl_handle& h, const float* image, size_t n, int deviation, float centre,
             uint64_t& n_finite, cudaStream_t stream) {
  const Selected s = SelectMiddle(h, image, n, deviation, centre, n_finite, stream);
  n_finite = s.n_finite;
  if (!n_finite) return 0.0f;
  const float hi = deviation ? FromBits(s.key_hi) : FromOrderedBits(s.key_hi);
  if (n_finite % 2 == 1) return hi;
  const float lo = deviation ? FromBits(s.key_lo) : FromOrderedBits(s.key_lo);
  return (lo + hi) * 0.5f;
}

}  // namespace

MadStatistics ImageMedianAndStdDevFromMad(rdl_handle& h, const float* image, size_t n,
                                          cudaStream_t stream) {
  if (n >= (size_t(1) << 32)) throw std::invalid_argument("MedianAndStdDevFromMAD: plane too large");
  MadStatistics result;
  if (n == 0) return result;
  uint64_t n_finite = 0;
  result.median = Median(h, image, n, 0, 0.0f, n_finite, stream);
  result.n_finite = n_finite;
  if (!n_finite) return result;
  result.stddev = 1.4826f * Median(h, image, n, 1, result.median, n_finite, stream);
  return result;
}

// ---- LoadAndAverage ---------------------------------------------------------------------------

namespace {

template <int VEC> struct VecOf;
template <> struct VecOf<1> { using type = float; };
template <> struct VecOf<4> { using type = float4; };

bool Aligned(const void* p, size_t a) { return reinterpret_cast<uintptr_t>(p) % a == 0; }

constexpr int kAverageTerms = 16;  // input planes one launch folds into an output plane

struct AverageParams {
  const float* in[kAverageTerms];
  float weight[kAverageTerms];
  double weight_d[kAverageTerms];  // the PSF rule multiplies in double
  int n_terms;
  int first, last;  // first: start from zero instead of the output's pixels; last: scale
  float scale;
  double scale_d;
  float* out;
};

// image = 0; image = fma(input, weight, image) per input in order (Image::AddWithFactor);
// image *= 1 / sum of weights (image_set.cc:105-140). One pass over the inputs of an output
// plane instead of a read-modify-write of the output per input.
// PSF_RULE: ImageSet::LoadAndAveragePsfs (image_set.cc:142-205) instead, whose weights stay
// double: psf = float(double(psf) + double(input) * weight), contracted by the reference's build
// to one double fma, and psf = float(double(psf) * (1 / sum of weights)).
template <int VEC, bool PSF_RULE>
__global__ void __launch_bounds__(256) LoadAndAverageKernel(AverageParams p, size_t n) {
  using V = typename VecOf<VEC>::type;
  const size_t stride = size_t(gridDim.x) * blockDim.x;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    float acc[VEC];
    if (p.first) {
#pragma unroll
      for (int k = 0; k != VEC; ++k) acc[k] = 0.0f;
    } else {
      const V a = reinterpret_cast<const V*>(p.out)[i];
#pragma unroll
      for (int k = 0; k != VEC; ++k) acc[k] = reinterpret_cast<const float*>(&a)[k];
    }
#pragma unroll 4
    for (int t = 0; t < p.n_terms; ++t) {
      const V v = __ldg(reinterpret_cast<const V*>(p.in[t]) + i);
      if (PSF_RULE) {
        const double w = p.weight_d[t];
#pragma unroll
        for (int k = 0; k != VEC; ++k)
          acc[k] = float(fma(double(reinterpret_cast<const float*>(&v)[k]), w, double(acc[k])));
      } else {
        const float w = p.weight[t];
#pragma unroll
        for (int k = 0; k != VEC; ++k)
          acc[k] = __fmaf_rn(reinterpret_cast<const float*>(&v)[k], w, acc[k]);
      }
    }
    if (p.last) {
#pragma unroll
      for (int k = 0; k != VEC; ++k)
        acc[k] = PSF_RULE ? float(double(acc[k]) * p.scale_d) : acc[k] * p.scale;
    }
    V o;
#pragma unroll
    for (int k = 0; k != VEC; ++k) reinterpret_cast<float*>(&o)[k] = acc[k];
    reinterpret_cast<V*>(p.out)[i] = o;
  }
}

constexpr int kMaskPlanes = 32;
struct MaskParams {
  const float* model[kMaskPlanes];
  int n_models, first;
};

template <int VEC>
__global__ void __launch_bounds__(256) MaskFromModelKernel(MaskParams p, uint8_t* mask, size_t n) {
  const size_t stride = size_t(gridDim.x) * blockDim.x;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint8_t set[VEC];
    if (p.first) {
#pragma unroll
      for (int k = 0; k != VEC; ++k) set[k] = 0;
    } else {
#pragma unroll
      for (int k = 0; k != VEC; ++k) set[k] = mask[i * VEC + k];
    }
    for (int m = 0; m < p.n_models; ++m) {
#pragma unroll
      for (int k = 0; k != VEC; ++k) {
        const float v = __ldg(p.model[m] + i * VEC + k);
        // std::isfinite(v) && v != 0.0
        if ((__float_as_uint(v) & 0x7f800000u) != 0x7f800000u && v != 0.0f) set[k] = 1;
      }
    }
#pragma unroll
    for (int k = 0; k != VEC; ++k) mask[i * VEC + k] = set[k];
  }
}

unsigned GridFor(size_t items) {
  return unsigned(std::max<size_t>(1, std::min<size_t>((items + 255) / 256, size_t(kSmCount) * 8)));
}

}  // namespace

void ImageLoadAndAverage(rdl_handle& h, const float* const* inputs, const double* weights,
                         const uint32_t* output_index, size_t n_inputs, float* const* outputs,
                         size_t n_outputs, size_t n_pixels, bool psf_rule, cudaStream_t stream) {
  (void)h;
  for (size_t i = 0; i != n_inputs; ++i) {
    if (output_index[i] >= n_outputs) throw std::invalid_argument("LoadAndAverage: output index");
    if (!inputs[i]) throw std::invalid_argument("LoadAndAverage: null input plane");
  }
  for (size_t o = 0; o != n_outputs; ++o) {
    if (!outputs[o]) throw std::invalid_argument("LoadAndAverage: null output plane");
    for (size_t i = 0; i != n_inputs; ++i)
      if (inputs[i] == outputs[o]) throw std::invalid_argument("LoadAndAverage: planes alias");
  }
  if (n_pixels == 0) return;
  for (size_t o = 0; o != n_outputs; ++o) {
    // the inputs of this output with a weight, in the caller's order (image_set.cc:126-131)
    std::vector<size_t> terms;
    double sum = 0.0;
    bool vec = n_pixels % 4 == 0 && Aligned(outputs[o], 16);
    for (size_t i = 0; i != n_inputs; ++i) {
      // LoadAndAverage leaves inputs without weight out; LoadAndAveragePsfs adds them x 0
      if (output_index[i] != o || (!psf_rule && weights[i] == 0.0)) continue;
      terms.push_back(i);
      sum += weights[i];
      vec = vec && Aligned(inputs[i], 16);
    }
    // with no weighted input LoadAndAverage scales a zero image by 1 / 0: NaN, reproduced;
    // LoadAndAveragePsfs scales by 0 (image_set.cc:196-198)
    const float scale = float(1.0 / sum);
    const double scale_d = sum == 0.0 ? 0.0 : 1.0 / sum;
    size_t done = 0;
    do {
      AverageParams p;
      p.n_terms = int(std::min<size_t>(kAverageTerms, terms.size() - done));
      for (int t = 0; t != p.n_terms; ++t) {
        p.in[t] = inputs[terms[done + t]];
        p.weight[t] = float(weights[terms[done + t]]);
        p.weight_d[t] = weights[terms[done + t]];
      }
      p.first = done == 0;
      done += size_t(p.n_terms);
      p.last = done == terms.size();
      p.scale = scale;
      p.scale_d = scale_d;
      p.out = outputs[o];
      const unsigned grid = GridFor(vec ? n_pixels / 4 : n_pixels);
      if (vec && psf_rule)
        LoadAndAverageKernel<4, true><<<grid, 256, 0, stream>>>(p, n_pixels / 4);
      else if (vec)
        LoadAndAverageKernel<4, false><<<grid, 256, 0, stream>>>(p, n_pixels / 4);
      else if (psf_rule)
        LoadAndAverageKernel<1, true><<<grid, 256, 0, stream>>>(p, n_pixels);
      else
        LoadAndAverageKernel<1, false><<<grid, 256, 0, stream>>>(p, n_pixels);
    } while (done != terms.size());
  }
  RDL_CUDA(cudaGetLastError());
}

void ImageMaskFromModel(rdl_handle& h, const float* const* models, size_t n_models, uint8_t* mask,
                        size_t n_pixels, cudaStream_t stream) {
  (void)h;
  for (size_t m = 0; m != n_models; ++m)
    if (!models[m]) throw std::invalid_argument("auto-mask from model: null model plane");
  if (n_pixels == 0) return;
  if (n_models == 0) {
    RDL_CUDA(cudaMemsetAsync(mask, 0, n_pixels, stream));
    return;
  }
  for (size_t done = 0; done != n_models;) {
    MaskParams p;
    p.n_models = int(std::min<size_t>(kMaskPlanes, n_models - done));
    bool vec = n_pixels % 4 == 0 && Aligned(mask, 4);
    for (int m = 0; m != p.n_models; ++m) {
      p.model[m] = models[done + m];
      vec = vec && Aligned(p.model[m], 16);
    }
    p.first = done == 0;
    done += size_t(p.n_models);
    if (vec)
      MaskFromModelKernel<4><<<GridFor(n_pixels / 4), 256, 0, stream>>>(p, mask, n_pixels / 4);
    else
      MaskFromModelKernel<1><<<GridFor(n_pixels), 256, 0, stream>>>(p, mask, n_pixels);
  }
  RDL_CUDA(cudaGetLastError());
}

// ---- InterpolateAndStoreModel -------------------------------------------------------------------

namespace {

constexpr int kMaxFitTerms = 8;        // polynomial terms held in registers per pixel
constexpr int kInterpolateOutputs = 32;  // output planes per launch

struct InterpolateParams {
  const float* model[kMaxImages];
  float* out[kInterpolateOutputs];
  const double* solve;   // t_used x n_models: channel values -> polynomial terms
  const double* powers;  // n_out x n_terms: (nu_out / nu_ref - 1)^k
  int n_models, t_used, n_terms, n_out;
};

// Per pixel (image_set.cc:236-283): the spectrum over the deconvolution channels -> polynomial
// terms (fitter.Fit; all-zero spectra give zero terms) -> fitter.Evaluate at every output
// frequency. Same operation order as the polynomial fitter of the clean loops: fma over the
// channels in double, terms rounded to float, then s = s + term * power in double.
__global__ void __launch_bounds__(256) InterpolateModelKernel(InterpolateParams p, size_t n) {
  __shared__ double s_solve[kMaxFitTerms * kMaxImages];
  __shared__ double s_powers[kInterpolateOutputs * kMaxFitTerms];
  for (int i = threadIdx.x; i < p.t_used * p.n_models; i += blockDim.x) s_solve[i] = p.solve[i];
  for (int i = threadIdx.x; i < p.n_out * p.n_terms; i += blockDim.x) s_powers[i] = p.powers[i];
  __syncthreads();
  const size_t stride = size_t(gridDim.x) * blockDim.x;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    double term[kMaxFitTerms];
#pragma unroll
    for (int k = 0; k != kMaxFitTerms; ++k) term[k] = 0.0;
    bool is_zero = true;
    for (int j = 0; j < p.n_models; ++j) {
      const float v = __ldg(p.model[j] + i);
      is_zero = is_zero && v == 0.0f;
      const double dv = double(v);
#pragma unroll
      for (int k = 0; k != kMaxFitTerms; ++k)
        if (k < p.t_used) term[k] = fma(s_solve[k * p.n_models + j], dv, term[k]);
    }
    if (is_zero) {
      for (int o = 0; o < p.n_out; ++o) p.out[o][i] = 0.0f;
      continue;
    }
    double rounded[kMaxFitTerms];
#pragma unroll
    for (int k = 0; k != kMaxFitTerms; ++k) rounded[k] = k < p.t_used ? double(float(term[k])) : 0.0;
    for (int o = 0; o < p.n_out; ++o) {
      double sum = 0.0;
#pragma unroll
      for (int k = 0; k != kMaxFitTerms; ++k)
        if (k < p.n_terms) sum = sum + rounded[k] * s_powers[o * p.n_terms + k];
      p.out[o][i] = float(sum);
    }
  }
}

}  // namespace

void ImageInterpolateModel(rdl_handle& h, const float* const* models, size_t n_models,
                           const double* output_frequencies, size_t n_outputs,
                           float* const* outputs, size_t n_pixels, cudaStream_t stream) {
  if (h.fit_mode != RDL_FIT_POLYNOMIAL)
    throw std::domain_error("interpolating the model needs the polynomial spectral fitter "
                            "(rdl_set_spectral_fitter)");
  if (n_models != h.fit_freqs.size())
    throw std::invalid_argument("interpolate model: one model plane per fitter channel");
  if (h.fit_terms > size_t(kMaxFitTerms))
    throw std::domain_error("interpolate model: at most 8 polynomial terms");
  for (size_t j = 0; j != n_models; ++j)
    if (!models[j]) throw std::invalid_argument("interpolate model: null model plane");
  for (size_t o = 0; o != n_outputs; ++o) {
    if (!outputs[o]) throw std::invalid_argument("interpolate model: null output plane");
    for (size_t j = 0; j != n_models; ++j)
      if (outputs[o] == models[j]) throw std::invalid_argument("interpolate model: planes alias");
  }
  if (n_pixels == 0 || n_outputs == 0) return;
  size_t t_used = 0;
  const size_t n_terms = h.fit_terms;
  std::vector<double> host = PolynomialTermSolve(n_terms, h.fit_freqs, h.fit_weights, t_used);
  const size_t solve_size = host.size();  // t_used x n_models
  // SpectralFitter::Evaluate: powers of nu / nu_ref - 1 by repeated multiplication
  const double reference = h.fit_freqs[h.fit_freqs.size() / 2];
  for (size_t o = 0; o != n_outputs; ++o) {
    const double x = output_frequencies[o] / reference - 1.0;
    double power = 1.0;
    for (size_t k = 0; k != n_terms; ++k) {
      host.push_back(power);
      power = power * x;
    }
  }
  h.interpolation.Reserve(std::max<size_t>(host.size(), 1));
  RDL_CUDA(cudaMemcpyAsync(h.interpolation.get(), host.data(), host.size() * sizeof(double),
                           cudaMemcpyHostToDevice, stream));
  for (size_t done = 0; done != n_outputs;) {
    InterpolateParams p;
    for (size_t j = 0; j != n_models; ++j) p.model[j] = models[j];
    p.n_out = int(std::min<size_t>(kInterpolateOutputs, n_outputs - done));
    for (int o = 0; o != p.n_out; ++o) p.out[o] = outputs[done + o];
    p.solve = h.interpolation.get();
    p.powers = h.interpolation.get() + solve_size + done * n_terms;
    p.n_models = int(n_models);
    p.t_used = int(t_used);
    p.n_terms = int(n_terms);
    done += size_t(p.n_out);
    InterpolateModelKernel<<<GridFor(n_pixels), 256, 0, stream>>>(p, n_pixels);
  }
  RDL_CUDA(cudaGetLastError());
  // the pageable staging vector must outlive the copy
  RDL_CUDA(cudaStreamSynchronize(stream));
}

}  // namespace rdl
