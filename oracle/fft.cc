// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h).
//
// Restatement of the schaapcommon::math convolution helpers that the reference
// calls (multiscale_transforms.cc:18-21, subminor_loop.cc:199-215). schaapcommon
// (pinned 7628fc36) and FFTW3f are NOT in /root/reference; semantics follow the
// call sites and SURVEY Appendix B:
//   Convolve(img, k, W, H)  = in-place circular convolution irfft2(rfft2(img) * rfft2(k)) / (W*H)
//   PrepareConvolutionKernel = circular shift moving (W/2, H/2) to (0, 0)
//   PrepareSmallConvolutionKernel = wrap an n x n kernel (centre n/2) around the origin
// The FFT itself is self-written: Stockham autosort, mixed radix {4,2,3,5,7,11,13},
// Bluestein for lengths with larger prime factors; single precision data with
// twiddles rounded from double. FFT results are not bit-reproducible across
// implementations (FFTW vs this vs the GPU) -> compared by tolerance only.
#include "oracle.h"

#include <algorithm>
#include <complex>
#include <cstring>
#include <map>
#include <mutex>

namespace oracle {

using cf = std::complex<float>;

namespace {

struct Stage {
  size_t radix, n_cur, m, s;
  std::vector<cf> tw;   // tw[p * radix + j] = exp(-2 pi i p j / n_cur)
  std::vector<cf> dft;  // dft[j * radix + k] = exp(-2 pi i j k / radix)
};

struct Plan {
  size_t n = 0;
  bool bluestein = false;
  std::vector<Stage> stages;
  // Bluestein
  size_t m = 0;
  std::vector<cf> chirp;      // exp(-i pi k^2 / n)
  std::vector<cf> chirp_fft;  // FFT_m of conj-chirp wrapped
  const Plan* sub = nullptr;
};

std::vector<size_t> Factorize(size_t n, bool& smooth) {
  std::vector<size_t> f;
  while (n % 4 == 0) { f.push_back(4); n /= 4; }
  for (size_t p : {2, 3, 5, 7, 11, 13})
    while (n % p == 0) { f.push_back(p); n /= p; }
  smooth = (n == 1);
  return f;
}

std::mutex g_plan_mutex;
std::map<size_t, std::unique_ptr<Plan>> g_plans;

const Plan& GetPlan(size_t n);

std::unique_ptr<Plan> MakePlan(size_t n) {
  auto plan = std::make_unique<Plan>();
  plan->n = n;
  bool smooth;
  std::vector<size_t> factors = Factorize(n, smooth);
  const double two_pi = 6.283185307179586476925286766559;
  if (smooth) {
    size_t n_cur = n, s = 1;
    for (size_t r : factors) {
      Stage st;
      st.radix = r;
      st.n_cur = n_cur;
      st.m = n_cur / r;
      st.s = s;
      st.tw.resize(st.m * r);
      for (size_t p = 0; p != st.m; ++p)
        for (size_t j = 0; j != r; ++j) {
          const double a = -two_pi * double((p * j) % n_cur) / double(n_cur);
          st.tw[p * r + j] = cf(float(std::cos(a)), float(std::sin(a)));
        }
      st.dft.resize(r * r);
      for (size_t j = 0; j != r; ++j)
        for (size_t k = 0; k != r; ++k) {
          const double a = -two_pi * double((j * k) % r) / double(r);
          st.dft[j * r + k] = cf(float(std::cos(a)), float(std::sin(a)));
        }
      plan->stages.push_back(std::move(st));
      n_cur /= r;
      s *= r;
    }
  } else {
    plan->bluestein = true;
    size_t m = 1;
    while (m < 2 * n - 1) m *= 2;
    plan->m = m;
    plan->chirp.resize(n);
    for (size_t k = 0; k != n; ++k) {
      const size_t k2 = (k * k) % (2 * n);
      const double a = -two_pi * 0.5 * double(k2) / double(n);
      plan->chirp[k] = cf(float(std::cos(a)), float(std::sin(a)));
    }
    plan->sub = &GetPlan(m);
  }
  return plan;
}

void ExecSmooth(const Plan& plan, cf* x, cf* y);

const Plan& GetPlan(size_t n) {
  {
    std::lock_guard<std::mutex> lock(g_plan_mutex);
    auto it = g_plans.find(n);
    if (it != g_plans.end()) return *it->second;
  }
  std::unique_ptr<Plan> p = MakePlan(n);
  if (p->bluestein) {
    std::vector<cf> b(p->m, cf(0, 0)), tmp(p->m);
    for (size_t k = 0; k != n; ++k) {
      b[k] = std::conj(p->chirp[k]);
      if (k != 0) b[p->m - k] = std::conj(p->chirp[k]);
    }
    ExecSmooth(*p->sub, b.data(), tmp.data());
    p->chirp_fft = b;
  }
  std::lock_guard<std::mutex> lock(g_plan_mutex);
  auto res = g_plans.emplace(n, std::move(p));
  return *res.first->second;
}

// Forward transform, in place in x (y is scratch of the same length).
void ExecSmooth(const Plan& plan, cf* x, cf* y) {
  cf* in = x;
  cf* out = y;
  for (const Stage& st : plan.stages) {
    const size_t r = st.radix, m = st.m, s = st.s;
    if (r == 2) {
      for (size_t p = 0; p != m; ++p) {
        const cf w1 = st.tw[p * 2 + 1];
        for (size_t q = 0; q != s; ++q) {
          const cf a = in[q + s * p], b = in[q + s * (p + m)];
          out[q + s * (2 * p)] = a + b;
          out[q + s * (2 * p + 1)] = (a - b) * w1;
        }
      }
    } else if (r == 4) {
      for (size_t p = 0; p != m; ++p) {
        const cf w1 = st.tw[p * 4 + 1], w2 = st.tw[p * 4 + 2], w3 = st.tw[p * 4 + 3];
        for (size_t q = 0; q != s; ++q) {
          const cf a = in[q + s * p], b = in[q + s * (p + m)], c = in[q + s * (p + 2 * m)],
                   d = in[q + s * (p + 3 * m)];
          const cf apc = a + c, amc = a - c, bpd = b + d;
          const cf bmd = b - d;
          const cf jbmd(bmd.imag(), -bmd.real());  // -i * (b - d)
          out[q + s * (4 * p)] = apc + bpd;
          out[q + s * (4 * p + 1)] = (amc + jbmd) * w1;
          out[q + s * (4 * p + 2)] = (apc - bpd) * w2;
          out[q + s * (4 * p + 3)] = (amc - jbmd) * w3;
        }
      }
    } else {
      cf a[16];
      for (size_t p = 0; p != m; ++p) {
        for (size_t q = 0; q != s; ++q) {
          for (size_t k = 0; k != r; ++k) a[k] = in[q + s * (p + k * m)];
          for (size_t j = 0; j != r; ++j) {
            cf b = a[0];
            for (size_t k = 1; k != r; ++k) b += a[k] * st.dft[j * r + k];
            out[q + s * (r * p + j)] = j == 0 ? b : b * st.tw[p * r + j];
          }
        }
      }
    }
    std::swap(in, out);
  }
  if (in != x) std::memcpy(x, in, plan.n * sizeof(cf));
}

void ExecForward(const Plan& plan, cf* x, cf* scratch /* >= max(n, 2m) */) {
  if (!plan.bluestein) {
    ExecSmooth(plan, x, scratch);
    return;
  }
  const size_t n = plan.n, m = plan.m;
  cf* a = scratch;
  cf* tmp = scratch + m;
  for (size_t k = 0; k != n; ++k) a[k] = x[k] * plan.chirp[k];
  std::fill(a + n, a + m, cf(0, 0));
  ExecSmooth(*plan.sub, a, tmp);
  for (size_t k = 0; k != m; ++k) a[k] = std::conj(a[k] * plan.chirp_fft[k]);
  ExecSmooth(*plan.sub, a, tmp);  // inverse via conj(FFT(conj(.)))
  const float inv = 1.0f / float(m);
  for (size_t k = 0; k != n; ++k) x[k] = std::conj(a[k]) * inv * plan.chirp[k];
}

size_t ScratchSize(const Plan& plan) { return plan.bluestein ? 2 * plan.m : plan.n; }

void ExecInverse(const Plan& plan, cf* x, cf* scratch) {  // unnormalised
  for (size_t k = 0; k != plan.n; ++k) x[k] = std::conj(x[k]);
  ExecForward(plan, x, scratch);
  for (size_t k = 0; k != plan.n; ++k) x[k] = std::conj(x[k]);
}

// rfft2: real W x H -> complex (W/2+1) x H, stored column-major-by-row [y][kx].
// Rows are transformed two at a time packed in one complex FFT.
void Rfft2(const float* in, cf* out, size_t w, size_t h) {
  const size_t wc = w / 2 + 1;
  const Plan& prow = GetPlan(w);
  const Plan& pcol = GetPlan(h);
  StaticFor(0, (h + 1) / 2, [&](size_t lo, size_t hi) {
    std::vector<cf> z(w), scratch(ScratchSize(prow));
    for (size_t pair = lo; pair != hi; ++pair) {
      const size_t y0 = 2 * pair, y1 = 2 * pair + 1;
      const bool has1 = y1 < h;
      const float* r0 = in + y0 * w;
      const float* r1 = has1 ? in + y1 * w : nullptr;
      for (size_t x = 0; x != w; ++x) z[x] = cf(r0[x], has1 ? r1[x] : 0.0f);
      ExecForward(prow, z.data(), scratch.data());
      for (size_t k = 0; k != wc; ++k) {
        const cf zk = z[k], zn = std::conj(z[(w - k) % w]);
        out[y0 * wc + k] = 0.5f * (zk + zn);
        if (has1) {
          const cf d = zk - zn;  // = 2i * B
          out[y1 * wc + k] = cf(0.5f * d.imag(), -0.5f * d.real());
        }
      }
    }
  });
  StaticFor(0, wc, [&](size_t lo, size_t hi) {
    std::vector<cf> col(h), scratch(ScratchSize(pcol));
    for (size_t k = lo; k != hi; ++k) {
      for (size_t y = 0; y != h; ++y) col[y] = out[y * wc + k];
      ExecForward(pcol, col.data(), scratch.data());
      for (size_t y = 0; y != h; ++y) out[y * wc + k] = col[y];
    }
  });
}

// irfft2 (unnormalised): complex (W/2+1) x H -> real W x H. Destroys `in`.
void Irfft2(cf* in, float* out, size_t w, size_t h) {
  const size_t wc = w / 2 + 1;
  const Plan& prow = GetPlan(w);
  const Plan& pcol = GetPlan(h);
  StaticFor(0, wc, [&](size_t lo, size_t hi) {
    std::vector<cf> col(h), scratch(ScratchSize(pcol));
    for (size_t k = lo; k != hi; ++k) {
      for (size_t y = 0; y != h; ++y) col[y] = in[y * wc + k];
      ExecInverse(pcol, col.data(), scratch.data());
      for (size_t y = 0; y != h; ++y) in[y * wc + k] = col[y];
    }
  });
  StaticFor(0, (h + 1) / 2, [&](size_t lo, size_t hi) {
    std::vector<cf> z(w), scratch(ScratchSize(prow));
    for (size_t pair = lo; pair != hi; ++pair) {
      const size_t y0 = 2 * pair, y1 = 2 * pair + 1;
      const bool has1 = y1 < h;
      // Build Z = A + i B from the two half spectra (Hermitian extension).
      for (size_t k = 0; k != w; ++k) {
        cf a, b(0, 0);
        if (k < wc) {
          a = in[y0 * wc + k];
          if (has1) b = in[y1 * wc + k];
        } else {
          a = std::conj(in[y0 * wc + (w - k)]);
          if (has1) b = std::conj(in[y1 * wc + (w - k)]);
        }
        if (k == 0 || (w % 2 == 0 && k == w / 2)) {  // enforce real DC / Nyquist like c2r
          a = cf(a.real(), 0.0f);
          b = cf(b.real(), 0.0f);
        }
        z[k] = a + cf(-b.imag(), b.real());
      }
      ExecInverse(prow, z.data(), scratch.data());
      float* r0 = out + y0 * w;
      float* r1 = has1 ? out + y1 * w : nullptr;
      for (size_t x = 0; x != w; ++x) {
        r0[x] = z[x].real();
        if (has1) r1[x] = z[x].imag();
      }
    }
  });
}

}  // namespace

void Fft1D(std::vector<std::pair<float, float>>& data, bool inverse) {
  const Plan& p = GetPlan(data.size());
  std::vector<cf> x(data.size()), scratch(ScratchSize(p));
  for (size_t i = 0; i != data.size(); ++i) x[i] = cf(data[i].first, data[i].second);
  if (inverse)
    ExecInverse(p, x.data(), scratch.data());
  else
    ExecForward(p, x.data(), scratch.data());
  for (size_t i = 0; i != data.size(); ++i) data[i] = {x[i].real(), x[i].imag()};
}

void PrepareSmallConvolutionKernel(float* dst, size_t w, size_t h, const float* kernel,
                                   size_t n) {
  // dst is pre-zeroed by the caller (multiscale_transforms.cc:16,29).
  const size_t half = n / 2;
  for (size_t ky = 0; ky != n; ++ky) {
    const size_t y = (ky + h - half) % h;
    for (size_t kx = 0; kx != n; ++kx) {
      const size_t x = (kx + w - half) % w;
      dst[y * w + x] = kernel[ky * n + kx];
    }
  }
}

void PrepareConvolutionKernel(float* dst, const float* src, size_t w, size_t h) {
  const size_t cx = w / 2, cy = h / 2;
  for (size_t y = 0; y != h; ++y) {
    const size_t sy = (y + cy) % h;
    for (size_t x = 0; x != w; ++x) dst[y * w + x] = src[sy * w + (x + cx) % w];
  }
}

void Convolve(float* image, const float* kernel, size_t w, size_t h) {
  const size_t wc = w / 2 + 1;
  std::vector<cf> fi(wc * h), fk(wc * h);
  Rfft2(image, fi.data(), w, h);
  Rfft2(kernel, fk.data(), w, h);
  const float fact = 1.0f / float(w * h);
  StaticFor(0, h, [&](size_t lo, size_t hi) {
    for (size_t i = lo * wc; i != hi * wc; ++i) fi[i] = fi[i] * fk[i] * fact;
  });
  Irfft2(fi.data(), image, w, h);
}

size_t CalculateGoodFFTSize(size_t minimum_size) {  // fft_size_calculations.h:15-29
  size_t best_factor = 2 * minimum_size;
  for (size_t f2 = 2; f2 < best_factor; f2 *= 2)
    for (size_t f23 = f2; f23 < best_factor; f23 *= 3)
      for (size_t f235 = f23; f235 < best_factor; f235 *= 5)
        for (size_t f2357 = f235; f2357 < best_factor; f2357 *= 7)
          if (f2357 >= minimum_size) best_factor = f2357;
  return best_factor;
}

size_t GetConvolutionSize(double convolution_scale, size_t original_size, double padding) {
  // fft_size_calculations.h:39-50
  return CalculateGoodFFTSize(
      size_t(std::ceil(padding * (convolution_scale * 1.5 + double(original_size)))));
}

}  // namespace oracle
