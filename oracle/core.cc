// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h).
// Image helpers, thread pool stand-in, ImageSet integration, peak finder,
// SubtractImage, spectral fitter.
#include "oracle.h"

#include <algorithm>
#include <atomic>
#include <cassert>
#include <cstring>
#include <limits>
#include <condition_variable>
#include <functional>
#include <mutex>
#include <thread>

namespace oracle {

// ---- aocommon::Image::Trim / Untrim (used at
// subminor_loop.cc:199-215) -------------------------------------------------
void Trim(float* dst, size_t dw, size_t dh, const float* src, size_t sw, size_t sh) {
  const size_t sx = (sw - dw) / 2, sy = (sh - dh) / 2;
  for (size_t y = 0; y != dh; ++y)
    std::memcpy(dst + y * dw, src + (y + sy) * sw + sx, dw * sizeof(float));
}

void Untrim(float* dst, size_t dw, size_t dh, const float* src, size_t sw, size_t sh) {
  const size_t sx = (dw - sw) / 2, sy = (dh - sh) / 2;
  std::fill(dst, dst + dw * dh, 0.0f);
  for (size_t y = 0; y != sh; ++y)
    std::memcpy(dst + (y + sy) * dw + sx, src + y * sw, sw * sizeof(float));
}

// ---- thread pool stand-in --------------------------------------------------
static size_t g_threads = 1;
void SetThreadCount(size_t n) { g_threads = std::max<size_t>(1, n); }
size_t ThreadCount() { return g_threads; }

// A persistent pool for top-level loops (the reference keeps an aocommon::ThreadPool alive for
// the whole run, radler.cc:138-142): spawning 16-32 std::threads per loop costs about half a
// millisecond, which is most of a small-frame Hogbom iteration. Loops started from inside a pool
// task (facets -> per-facet row loops) fall back to short-lived threads.
namespace {
class Pool {
 public:
  static Pool& Get() {
    static Pool p;
    return p;
  }
  // runs task(t) for t in [0, n_tasks) on up to n_workers threads and waits
  void Run(size_t n_tasks, size_t n_workers, const std::function<void(size_t)>& task) {
    std::unique_lock<std::mutex> run_lock(run_mutex_);  // one top-level loop at a time
    Grow(n_workers);
    {
      std::lock_guard<std::mutex> lk(m_);
      task_ = &task;
      n_tasks_ = n_tasks;
      next_ = 0;
      pending_ = n_tasks;
      ++generation_;
    }
    cv_.notify_all();
    std::unique_lock<std::mutex> lk(m_);
    done_.wait(lk, [&] { return pending_ == 0; });
    task_ = nullptr;
  }
  static thread_local bool inside;

 private:
  ~Pool() {
    {
      std::lock_guard<std::mutex> lk(m_);
      stop_ = true;
    }
    cv_.notify_all();
    for (auto& t : threads_) t.join();
  }
  void Grow(size_t n) {
    while (threads_.size() < n) threads_.emplace_back([this] { Work(); });
  }
  void Work() {
    inside = true;
    uint64_t seen = 0;
    std::unique_lock<std::mutex> lk(m_);
    for (;;) {
      cv_.wait(lk, [&] { return stop_ || (generation_ != seen && next_ < n_tasks_); });
      if (stop_) return;
      while (next_ < n_tasks_) {
        const size_t t = next_++;
        const std::function<void(size_t)>* task = task_;
        lk.unlock();
        (*task)(t);
        lk.lock();
        if (--pending_ == 0) done_.notify_all();
      }
      seen = generation_;
    }
  }
  std::mutex m_, run_mutex_;
  std::condition_variable cv_, done_;
  std::vector<std::thread> threads_;
  const std::function<void(size_t)>* task_ = nullptr;
  size_t n_tasks_ = 0, next_ = 0, pending_ = 0;
  uint64_t generation_ = 0;
  bool stop_ = false;
};
thread_local bool Pool::inside = false;
}  // namespace

void StaticFor(size_t begin, size_t end, const std::function<void(size_t, size_t)>& f) {
  const size_t n = end - begin;
  const size_t nt = std::min(g_threads, std::max<size_t>(n, 1));
  if (nt <= 1 || n == 0) {
    if (n) f(begin, end);
    return;
  }
  if (!Pool::inside) {
    Pool::Get().Run(nt, nt, [&](size_t t) {
      const size_t lo = begin + n * t / nt, hi = begin + n * (t + 1) / nt;
      if (lo != hi) f(lo, hi);
    });
    return;
  }
  std::vector<std::thread> pool;
  for (size_t t = 0; t != nt; ++t) {
    const size_t lo = begin + n * t / nt, hi = begin + n * (t + 1) / nt;
    if (lo != hi) pool.emplace_back([&f, lo, hi] { f(lo, hi); });
  }
  for (auto& t : pool) t.join();
}

void DynamicFor(size_t begin, size_t end, const std::function<void(size_t)>& f) {
  const size_t n = end - begin;
  const size_t nt = std::min(g_threads, n);
  if (nt <= 1) {
    for (size_t i = begin; i != end; ++i) f(i);
    return;
  }
  if (!Pool::inside) {
    Pool::Get().Run(n, nt, [&](size_t i) { f(begin + i); });
    return;
  }
  std::atomic<size_t> next{begin};
  std::vector<std::thread> pool;
  for (size_t t = 0; t != nt; ++t)
    pool.emplace_back([&] {
      for (;;) {
        const size_t i = next.fetch_add(1);
        if (i >= end) break;
        f(i);
      }
    });
  for (auto& t : pool) t.join();
}

// ---- ImageSet ---------------------------------------------------------------
ImageSet::ImageSet(const CubeMeta& m, size_t w_, size_t h_) : meta(m), w(w_), h(h_) {
  if (meta.weights.empty()) meta.weights.assign(meta.n_channels, 1.0f);
  if (meta.pol_linked.empty()) meta.pol_linked.assign(meta.n_pol, 1);
  images.reserve(meta.n_channels * meta.n_pol);
  for (size_t i = 0; i != meta.n_channels * meta.n_pol; ++i) images.emplace_back(w, h);
}

// The element-wise helpers below restate aocommon::Image ops with the FMA
// contraction a -O3 -march=native GCC build applies; the
// CUDA kernels use the same explicit fmaf so peak order cannot drift.
static void AssignMultiply(Image& lhs, const Image& rhs, float factor) {  // image_set.cc:16-24
  for (size_t i = 0; i != lhs.Size(); ++i) lhs[i] = rhs[i] * factor;
}
static void AddWithFactor(Image& lhs, const Image& rhs, float factor) {
  for (size_t i = 0; i != lhs.Size(); ++i) lhs[i] = std::fmaf(rhs[i], factor, lhs[i]);
}
static void Scale(Image& img, float f) {
  for (float& v : img.d) v *= f;
}

void ImageSet::GetLinearIntegrated(Image& dest) const {  // image_set.h:150-155
  if (meta.squared_joins)
    SquareSquared(dest);
  else
    LinearNormal(dest);
}

void ImageSet::GetSquareIntegrated(Image& dest, Image& scratch) const {  // image_set.h:131-137
  if (meta.squared_joins)
    SquareSquared(dest);
  else
    SquareNormal(dest, scratch);
}

void ImageSet::LinearNormal(Image& dest) const {  // image_set.cc:423-462
  if (meta.n_channels == 1 && meta.n_pol == 1) {
    std::copy(images[0].d.begin(), images[0].d.end(), dest.d.begin());
    return;
  }
  bool is_first = true;
  double weight_sum = 0.0;
  for (size_t ch = 0; ch != meta.n_channels; ++ch) {
    const double group_weight = meta.weights[ch];
    if (group_weight != 0.0) {
      weight_sum += group_weight;
      for (size_t p = 0; p != meta.n_pol; ++p) {
        if (!meta.pol_linked[p]) continue;
        const Image& img = images[ch * meta.n_pol + p];
        if (is_first) {
          AssignMultiply(dest, img, float(group_weight));
          is_first = false;
        } else {
          AddWithFactor(dest, img, float(group_weight));
        }
      }
    }
  }
  if (weight_sum > 0.0)
    Scale(dest, float(meta.pol_norm / weight_sum));
  else
    std::fill(dest.d.begin(), dest.d.end(), 0.0f);
}

void ImageSet::SquareNormal(Image& dest, Image& scratch) const {  // image_set.cc:309-386
  const size_t n = w * h;
  auto pol_quadrature = [&](size_t ch, Image& out) -> bool {
    // out = sum over linked pols of v^2 (first: Square, then AddSquared)
    bool first = true;
    for (size_t p = 0; p != meta.n_pol; ++p) {
      if (!meta.pol_linked[p]) continue;
      const Image& img = images[ch * meta.n_pol + p];
      if (first) {
        for (size_t i = 0; i != n; ++i) out[i] = img[i] * img[i];
        first = false;
      } else {
        for (size_t i = 0; i != n; ++i) out[i] = std::fmaf(img[i], img[i], out[i]);
      }
    }
    return !first;
  };
  if (meta.n_channels == 1) {
    if (meta.n_pol == 1) {
      std::copy(images[0].d.begin(), images[0].d.end(), dest.d.begin());
    } else {
      pol_quadrature(0, dest);
      const float f = std::sqrt(meta.pol_norm);  // SqrtWithFactor
      for (size_t i = 0; i != n; ++i) dest[i] = std::sqrt(dest[i]) * f;
    }
    return;
  }
  double weight_sum = 0.0;
  bool is_first_channel = true;
  for (size_t ch = 0; ch != meta.n_channels; ++ch) {
    const double group_weight = meta.weights[ch];
    if (group_weight != 0.0) {
      weight_sum += group_weight;
      if (meta.n_pol == 1) {
        std::copy(images[ch].d.begin(), images[ch].d.end(), scratch.d.begin());
      } else {
        if (pol_quadrature(ch, scratch))
          for (size_t i = 0; i != n; ++i) scratch[i] = std::sqrt(scratch[i]);
        else
          std::fill(scratch.d.begin(), scratch.d.begin() + n, 0.0f);
      }
    } else {
      std::fill(scratch.d.begin(), scratch.d.begin() + n, 0.0f);
    }
    if (is_first_channel) {
      for (size_t i = 0; i != n; ++i) dest[i] = scratch[i] * float(group_weight);
      is_first_channel = false;
    } else {
      const float gw = float(group_weight);
      for (size_t i = 0; i != n; ++i) dest[i] = std::fmaf(scratch[i], gw, dest[i]);
    }
  }
  const float f = float(std::sqrt(meta.pol_norm) / weight_sum);
  for (size_t i = 0; i != n; ++i) dest[i] *= f;
}

void ImageSet::SquareSquared(Image& dest) const {  // image_set.cc:388-421
  const size_t n = w * h;
  bool is_first = true;
  double weight_sum = 0.0;
  for (size_t ch = 0; ch != meta.n_channels; ++ch) {
    const double group_weight = meta.weights[ch];
    if (group_weight != 0.0) {
      weight_sum += group_weight;
      const float gw = float(group_weight);
      for (size_t p = 0; p != meta.n_pol; ++p) {
        if (!meta.pol_linked[p]) continue;
        const Image& img = images[ch * meta.n_pol + p];
        if (is_first) {  // SquareWithFactor: d *= d * f
          for (size_t i = 0; i != n; ++i) dest[i] = img[i] * (img[i] * gw);
          is_first = false;
        } else {  // AddSquared(rhs, f): d += r * r * f
          for (size_t i = 0; i != n; ++i) dest[i] = std::fmaf(img[i] * img[i], gw, dest[i]);
        }
      }
    }
  }
  if (weight_sum > 0.0) {
    const float f = float(std::sqrt(meta.pol_norm / weight_sum));
    for (size_t i = 0; i != n; ++i) dest[i] = std::sqrt(dest[i]) * f;
  } else {
    std::fill(dest.d.begin(), dest.d.begin() + n, 0.0f);
  }
}

void ImageSet::GetIntegratedPsf(Image& dest, const std::vector<Image>& psfs) const {
  // image_set.cc:499-530
  if (meta.n_channels == 1) {
    std::copy(psfs[0].d.begin(), psfs[0].d.end(), dest.d.begin());
    return;
  }
  bool is_first = true;
  double weight_sum = 0.0;
  for (size_t ch = 0; ch != meta.n_channels; ++ch) {
    const double gw = meta.weights[ch];
    if (gw != 0.0) {
      weight_sum += gw;
      if (is_first) {
        for (size_t i = 0; i != dest.Size(); ++i) dest[i] = psfs[ch][i] * float(gw);
        is_first = false;
      } else {
        AddWithFactor(dest, psfs[ch], float(gw));
      }
    }
  }
  const double factor = weight_sum == 0.0 ? 0.0 : 1.0 / weight_sum;
  Scale(dest, float(factor));
}

// ---- peak finder (math/peak_finder.cc:19-56 Simple, :97-131 FindWithMask) ----
std::optional<float> FindPeak(const float* image, size_t width, size_t height, size_t& x,
                              size_t& y, bool allow_negative, size_t start_y, size_t end_y,
                              size_t h_border, size_t v_border, const bool* mask) {
  float peak_max = std::numeric_limits<float>::min();
  size_t peak_index = width * height;
  size_t xi_start = h_border, xi_end = width - h_border;
  size_t yi_start = std::max(start_y, v_border), yi_end = std::min(end_y, height - v_border);
  if (xi_end < xi_start) xi_end = xi_start;
  if (yi_end < yi_start) yi_end = yi_start;
  for (size_t yi = yi_start; yi != yi_end; ++yi) {
    size_t index = yi * width + xi_start;
    for (size_t xi = xi_start; xi != xi_end; ++xi, ++index) {
      float value = image[index];
      if (allow_negative) value = std::fabs(value);
      if (value > peak_max && (!mask || mask[index])) {
        peak_index = index;
        peak_max = std::fabs(value);
      }
    }
  }
  if (peak_index == width * height) {
    x = width;
    y = height;
    return {};
  }
  x = peak_index % width;
  y = peak_index / width;
  return image[peak_index];
}

std::optional<float> FindPeakRatio(const float* image, size_t width, size_t height, size_t& x,
                                   size_t& y, bool allow_negative, size_t start_y,
                                   size_t end_y, float border_ratio, const bool* mask) {
  // peak_finder.h:99-107: round(width * ratio) in float->double promotion of `round`
  return FindPeak(image, width, height, x, y, allow_negative, start_y, end_y,
                  size_t(std::round(width * border_ratio)),
                  size_t(std::round(height * border_ratio)), mask);
}

// ---- simple_clean -------------------------------------------------------------
void SubtractImageWhole(float* image, const float* psf, size_t width, size_t height, size_t x,
                        size_t y, float factor) {  // simple_clean.cc:17-59
  const int offset_x = int(x) - int(width / 2), offset_y = int(y) - int(height / 2);
  size_t start_x = offset_x > 0 ? offset_x : 0, start_y = offset_y > 0 ? offset_y : 0;
  size_t end_x = std::min(x + width / 2, width);
  if ((end_x - start_x) % 2 != 0) --end_x;
  const size_t end_y = std::min(y + height / 2, height);
  for (size_t yp = start_y; yp != end_y; ++yp) {
    float* img = image + yp * width;
    const float* p = psf + (yp - offset_y) * width - offset_x;
    for (size_t xp = start_x; xp != end_x; ++xp) img[xp] = std::fmaf(-p[xp], factor, img[xp]);
  }
}

void PartialSubtractImage(float* image, size_t image_width, size_t /*image_height*/,
                          const float* psf, size_t psf_width, size_t psf_height, size_t x,
                          size_t y, float factor, size_t start_y, size_t end_y) {
  // simple_clean.cc:98-135; `a - b*f` is contracted to one FMA by the reference build.
  const int offset_x = int(x) - int(psf_width / 2), offset_y = int(y) - int(psf_height / 2);
  const size_t start_x = offset_x > 0 ? offset_x : 0;
  if (offset_y > int(start_y)) start_y = offset_y;
  const size_t end_x = std::min(x + psf_width / 2, image_width);
  end_y = std::min(y + psf_height / 2, end_y);
  for (size_t yp = start_y; yp < end_y; ++yp) {
    float* img = image + yp * image_width;
    const float* p = psf + (yp - offset_y) * psf_width - offset_x;
    for (size_t xp = start_x; xp < end_x; ++xp) img[xp] = std::fmaf(-p[xp], factor, img[xp]);
  }
}

void PartialSubtractImage(float* image, const float* psf, size_t width, size_t height,
                          size_t x, size_t y, float factor, size_t start_y, size_t end_y) {
  // simple_clean.cc:61-96 (same-size variant)
  PartialSubtractImage(image, width, height, psf, width, height, x, y, factor, start_y, end_y);
}

void ThreadedSubtractImage(float* image, const Image& psf, size_t x, size_t y, float factor) {
  // threaded_deconvolution_tools.cc:18-28
  StaticFor(0, psf.h, [&](size_t lo, size_t hi) {
    PartialSubtractImage(image, psf.Data(), psf.w, psf.h, x, y, factor, lo, hi);
  });
}

// ---- spectral fitter ------------------------------------------------------------
namespace {
// xp: powers x_i^k (n x t); returns the term-solve operator T (t x n), empty when no channel
// carries weight. t = min(n_terms, channels with weight).
std::vector<double> SolveForTerms(size_t n_terms, const std::vector<double>& freqs,
                                  const std::vector<float>& weights, std::vector<double>& xp,
                                  size_t& t) {
  const size_t n = freqs.size();
  std::vector<size_t> used;
  for (size_t i = 0; i != n; ++i)
    if (weights[i] > 0.0f) used.push_back(i);
  const size_t m = used.size();
  t = std::min(n_terms, m);
  if (t == 0) return {};
  const double ref = freqs[n / 2];
  // powers x_i^k for all channels (also for zero-weight ones: they are evaluated)
  xp.assign(n * t, 0.0);
  for (size_t i = 0; i != n; ++i) {
    const double x = freqs[i] / ref - 1.0;
    double p = 1.0;
    for (size_t k = 0; k != t; ++k) {
      xp[i * t + k] = p;
      p = p * x;
    }
  }
  // normal matrix N = X^T W X (t x t), rhs B = X^T W (t x n)
  std::vector<double> N(t * t, 0.0), B(t * n, 0.0);
  for (size_t a = 0; a != t; ++a) {
    for (size_t b = 0; b != t; ++b) {
      double s = 0.0;
      for (size_t u : used) s = s + double(weights[u]) * xp[u * t + a] * xp[u * t + b];
      N[a * t + b] = s;
    }
    for (size_t u : used) B[a * n + u] = double(weights[u]) * xp[u * t + a];
  }
  // Gaussian elimination with partial pivoting on [N | B]
  for (size_t col = 0; col != t; ++col) {
    size_t piv = col;
    for (size_t r = col + 1; r != t; ++r)
      if (std::fabs(N[r * t + col]) > std::fabs(N[piv * t + col])) piv = r;
    if (piv != col) {
      for (size_t c = 0; c != t; ++c) std::swap(N[col * t + c], N[piv * t + c]);
      for (size_t c = 0; c != n; ++c) std::swap(B[col * n + c], B[piv * n + c]);
    }
    const double d = N[col * t + col];
    for (size_t r = col + 1; r != t; ++r) {
      const double f = N[r * t + col] / d;
      for (size_t c = col; c != t; ++c) N[r * t + c] = N[r * t + c] - f * N[col * t + c];
      for (size_t c = 0; c != n; ++c) B[r * n + c] = B[r * n + c] - f * B[col * n + c];
    }
  }
  for (size_t col = t; col-- > 0;) {
    const double d = N[col * t + col];
    for (size_t c = 0; c != n; ++c) {
      double s = B[col * n + c];
      for (size_t k = col + 1; k != t; ++k) s = s - N[col * t + k] * B[k * n + c];
      B[col * n + c] = s / d;
    }
  }
  return B;
}
}  // namespace

std::vector<double> PolynomialTermSolve(size_t n_terms, const std::vector<double>& freqs,
                                        const std::vector<float>& weights, size_t& t_used) {
  std::vector<double> xp;
  return SolveForTerms(n_terms, freqs, weights, xp, t_used);
}

double EvaluatePolynomial(const float* terms, size_t n_terms, double x) {
  double p = 1.0, s = 0.0;
  for (size_t k = 0; k != n_terms; ++k) {
    s = s + double(terms[k]) * p;
    p = p * x;
  }
  return s;
}

std::vector<double> PolynomialProjection(size_t n_terms, const std::vector<double>& freqs,
                                         const std::vector<float>& weights) {
  const size_t n = freqs.size();
  std::vector<double> proj(n * n, 0.0), xp;
  size_t t = 0;
  const std::vector<double> B = SolveForTerms(n_terms, freqs, weights, xp, t);
  if (t == 0) return proj;
  // B holds the term-solve operator T (t x n); projection = X T
  for (size_t i = 0; i != n; ++i)
    for (size_t j = 0; j != n; ++j) {
      double s = 0.0;
      for (size_t k = 0; k != t; ++k) s = s + xp[i * t + k] * B[k * n + j];
      proj[i * n + j] = s;
    }
  return proj;
}

SpectralFitter::SpectralFitter(SpectralMode mode, size_t n_terms, const std::vector<double>& freqs,
                               const std::vector<float>& weights)
    : mode_(mode), n_terms_(n_terms), freqs_(freqs), weights_(weights) {
  if (mode_ == SpectralMode::kPolynomial) proj_ = PolynomialProjection(n_terms, freqs, weights);
}

void SpectralFitter::FitAndEvaluate(float* values) const {
  if (mode_ == SpectralMode::kNoFitting) return;
  const size_t n = freqs_.size();
  double out[64];
  assert(n <= 64);
  for (size_t i = 0; i != n; ++i) {
    double s = 0.0;
    for (size_t j = 0; j != n; ++j) s = std::fma(proj_[i * n + j], double(values[j]), s);
    out[i] = s;
  }
  for (size_t i = 0; i != n; ++i) values[i] = float(out[i]);
}

void DeconvolutionAlgorithm::PerformSpectralFit(float* values) const {
  // deconvolution_algorithm.cc:29-46
  if (fitter.Mode() == SpectralMode::kNoFitting) return;
  const size_t n = fitter.NChannels();
  for (size_t p = 0; p != n_polarizations; ++p) {
    for (size_t ch = 0; ch != n; ++ch) std::swap(values[ch * n_polarizations + p], values[ch]);
    fitter.FitAndEvaluate(values);
    for (size_t i = 0; i != n; ++i) {
      const size_t ch = n - i - 1;
      std::swap(values[ch * n_polarizations + p], values[ch]);
    }
  }
}

}  // namespace oracle
