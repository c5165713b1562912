// ORACLE — TEST INFRASTRUCTURE ONLY (never linked into or called by the product).
// CPU restatement of the reference's IUWT decomposition:
//   cpp/algorithms/iuwt/iuwt_decomposition.cc:9-53   DecomposeMt
//   cpp/algorithms/iuwt/iuwt_decomposition.cc:55-76  convolveMT / differenceMT
//   cpp/algorithms/iuwt/iuwt_decomposition.cc:78-131 convolveHorizontalFast (tap order h2,h1,h0,h3,h4)
//   cpp/algorithms/iuwt/iuwt_decomposition.cc:163-237 convolveVerticalPartialFast (tap order h0..h4)
//   cpp/algorithms/iuwt/iuwt_decomposition.h:249-267 convolve ("simple": taps h0..h4 accumulated
//                                                     onto a zeroed image, both directions)
//   cpp/algorithms/iuwt/iuwt_decomposition.h:121-147 Recompose, :162-169 ApplyMask, :57-63 Convolve
// B3-spline taps {1,4,6,4,1}/16 at offsets (2^scale - 1) * {-2..2}; taps that fall outside the
// image are dropped (zero extension, no renormalisation). Only the 6/16 product is inexact, so
// what matters is where the reference's -O3 build (GCC -ffp-contract=fast) fuses it; the
// contractions are written explicitly here (this file is compiled with -ffp-contract=off). Pinned bit-exact against the reference's own sources (oracle/_ref) by
// tests/test_iuwt_oracle.py.
#include <algorithm>
#include <cmath>
#include <cstring>
#include <vector>

#include "oracle.h"

namespace oracle {
namespace {
constexpr float kH0 = 1.0f / 16.0f, kH1 = 4.0f / 16.0f, kH2 = 6.0f / 16.0f;

// One output value from up to five taps v[0..4] (present[k] false = tap outside the image).
// MT flavour (convolveHorizontalFast / convolveVerticalPartialFast): one expression
// "p_a + p_b + p_c ..." in the order given; GCC's contraction fuses the FIRST product with the
// second one left as a plain (rounded) product, every later term becomes fma(v, h, acc).
// Simple flavour (convolveComponent*): "out += in * h" per tap on a zeroed image, h0..h4.
inline float Taps(const float v[5], const bool present[5], const int order[5], bool simple) {
  static const float h[5] = {kH0, kH1, kH2, kH1, kH0};
  float tv[5], th[5];
  int n = 0;
  for (int i = 0; i != 5; ++i) {
    const int k = order[i];
    if (present[k]) {
      tv[n] = v[k];
      th[n] = h[k];
      ++n;
    }
  }
  if (simple) {
    float acc = 0.0f;
    for (int i = 0; i != n; ++i) acc = std::fmaf(tv[i], th[i], acc);
    return acc;
  }
  if (n == 1) return tv[0] * th[0];
  float acc = std::fmaf(tv[0], th[0], tv[1] * th[1]);
  for (int i = 2; i != n; ++i) acc = std::fmaf(tv[i], th[i], acc);
  return acc;
}
constexpr int kOrderMtHorizontal[5] = {2, 1, 0, 3, 4};
constexpr int kOrderNatural[5] = {0, 1, 2, 3, 4};
}  // namespace

// out = B3 smoothing of in at `scale` (>= 1). mt = the DecomposeMt flavour (convolveMT),
// otherwise the "simple" flavour (convolve) used by Recompose and Convolve.
void IuwtSmooth(float* out, const float* in, float* scratch, size_t width, size_t height, int scale,
                bool mt) {
  const long d = (1L << scale) - 1;
  const long w = long(width), h = long(height);
  StaticFor(0, height, [&](size_t y0, size_t y1) {
    for (size_t y = y0; y != y1; ++y) {
      const float* row = in + y * width;
      float* dst = scratch + y * width;
      for (long x = 0; x != w; ++x) {
        float v[5];
        bool p[5];
        for (int k = 0; k != 5; ++k) {
          const long xx = x + (k - 2) * d;
          p[k] = xx >= 0 && xx < w;
          v[k] = p[k] ? row[xx] : 0.0f;
        }
        dst[x] = Taps(v, p, mt ? kOrderMtHorizontal : kOrderNatural, !mt);
      }
    }
  });
  StaticFor(0, height, [&](size_t y0, size_t y1) {
    for (size_t y = y0; y != y1; ++y) {
      float* dst = out + y * width;
      for (long x = 0; x != w; ++x) {
        float v[5];
        bool p[5];
        for (int k = 0; k != 5; ++k) {
          const long yy = long(y) + (k - 2) * d;
          p[k] = yy >= 0 && yy < h;
          v[k] = p[k] ? scratch[yy * w + x] : 0.0f;
        }
        dst[x] = Taps(v, p, kOrderNatural, !mt);
      }
    }
  });
}

// iuwt_decomposition.cc:9-53. coefficients: n_scales planes (+1 when include_largest).
void IuwtDecompose(const float* input, size_t width, size_t height, int n_scales,
                   bool include_largest, float* coefficients) {
  const size_t a = width * height;
  std::vector<float> scratch(a), i0(input, input + a), i1(a), i2(a);
  for (int s = 0; s != n_scales; ++s) {
    IuwtSmooth(i1.data(), i0.data(), scratch.data(), width, height, s + 1, true);
    IuwtSmooth(i2.data(), i1.data(), scratch.data(), width, height, s + 1, true);
    float* c = coefficients + size_t(s) * a;
    for (size_t i = 0; i != a; ++i) c[i] = i0[i] - i2[i];
    i0 = i1;
  }
  if (include_largest) std::memcpy(coefficients + size_t(n_scales) * a, i1.data(), a * sizeof(float));
}

// ApplyMask (iuwt_decomposition.h:162-169) when masks != nullptr, then Recompose (:121-147).
void IuwtRecompose(const float* coefficients, size_t width, size_t height, int n_scales,
                   bool include_largest, const uint8_t* masks, float* out) {
  const size_t a = width * height;
  std::vector<float> scratch(a), tmp(a);
  bool is_zero = true;
  if (include_largest && !masks) {
    std::memcpy(out, coefficients + size_t(n_scales) * a, a * sizeof(float));
    is_zero = false;
  }
  for (int s = n_scales - 1; s != -1; --s) {
    const float* c = coefficients + size_t(s) * a;
    const uint8_t* m = masks ? masks + size_t(s) * a : nullptr;
    if (is_zero) {
      for (size_t i = 0; i != a; ++i) out[i] = (!m || m[i]) ? c[i] : 0.0f;
      is_zero = false;
    } else {
      IuwtSmooth(tmp.data(), out, scratch.data(), width, height, s + 1, false);
      for (size_t i = 0; i != a; ++i) out[i] = tmp[i] + ((!m || m[i]) ? c[i] : 0.0f);
    }
  }
}

}  // namespace oracle
