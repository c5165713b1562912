// Host-side helpers: integration recipes (image_set.cc:309-530) and the spectral-fit
// projection matrix. Compiled with -ffp-contract=off: every operation below rounds once.
#include <cmath>

#include "handle.h"

namespace rdl {

namespace {
IntegSpec ZeroSpec() {
  IntegSpec s{};
  s.mode = kIntegZero;
  return s;
}

IntegSpec MakeSquaredSpec(const DeviceCube& c) {  // image_set.cc:388-421
  IntegSpec s{};
  s.mode = kIntegSquared;
  double weight_sum = 0.0;
  for (size_t ch = 0; ch != c.n_channels; ++ch) {
    const double gw = c.weights[ch];
    if (gw == 0.0) continue;
    weight_sum += gw;
    for (size_t p = 0; p != c.n_pol; ++p) {
      if (!c.pol_linked[p]) continue;
      s.img[s.n_terms] = int(ch * c.n_pol + p);
      s.w[s.n_terms] = float(gw);
      ++s.n_terms;
    }
  }
  if (!(weight_sum > 0.0) || s.n_terms == 0) return ZeroSpec();
  s.factor = float(std::sqrt(c.pol_norm / weight_sum));
  return s;
}
}  // namespace

IntegSpec MakeLinearSpec(const DeviceCube& c) {  // image_set.cc:423-462
  if (c.squared_joins) return MakeSquaredSpec(c);
  IntegSpec s{};
  if (c.n_channels == 1 && c.n_pol == 1) {
    s.mode = kIntegCopy;
    s.n_terms = 1;
    s.img[0] = 0;
    return s;
  }
  s.mode = kIntegLinear;
  double weight_sum = 0.0;
  for (size_t ch = 0; ch != c.n_channels; ++ch) {
    const double gw = c.weights[ch];
    if (gw == 0.0) continue;
    weight_sum += gw;
    for (size_t p = 0; p != c.n_pol; ++p) {
      if (!c.pol_linked[p]) continue;
      s.img[s.n_terms] = int(ch * c.n_pol + p);
      s.w[s.n_terms] = float(gw);
      ++s.n_terms;
    }
  }
  if (!(weight_sum > 0.0) || s.n_terms == 0) return ZeroSpec();
  s.factor = float(c.pol_norm / weight_sum);
  return s;
}

IntegSpec MakeSquareSpec(const DeviceCube& c) {  // image_set.cc:309-386
  if (c.squared_joins) return MakeSquaredSpec(c);
  IntegSpec s{};
  size_t n_linked = 0;
  for (size_t p = 0; p != c.n_pol; ++p) n_linked += c.pol_linked[p] ? 1 : 0;
  if (c.n_channels == 1) {
    if (c.n_pol == 1) {
      s.mode = kIntegCopy;
      s.n_terms = 1;
      s.img[0] = 0;
      return s;
    }
    if (n_linked == 0) return ZeroSpec();
    s.mode = kIntegSquareSingle;
    s.n_terms = 1;
    s.n_pol_used = int(n_linked);
    int k = 0;
    for (size_t p = 0; p != c.n_pol; ++p)
      if (c.pol_linked[p]) s.img[k++] = int(p);
    s.factor = std::sqrt(c.pol_norm);  // float sqrt, as SqrtWithFactor(std::sqrt(float))
    return s;
  }
  double weight_sum = 0.0;
  if (c.n_pol == 1) {
    // scratch = image; dest = scratch*w / fma: same arithmetic as the linear recipe
    s.mode = kIntegLinear;
    for (size_t ch = 0; ch != c.n_channels; ++ch) {
      const double gw = c.weights[ch];
      if (gw == 0.0) continue;
      weight_sum += gw;
      s.img[s.n_terms] = int(ch);
      s.w[s.n_terms] = float(gw);
      ++s.n_terms;
    }
  } else {
    if (n_linked == 0) return ZeroSpec();
    s.mode = kIntegSquareNormal;
    s.n_pol_used = int(n_linked);
    for (size_t ch = 0; ch != c.n_channels; ++ch) {
      const double gw = c.weights[ch];
      if (gw == 0.0) continue;
      weight_sum += gw;
      int k = s.n_terms * s.n_pol_used;
      for (size_t p = 0; p != c.n_pol; ++p)
        if (c.pol_linked[p]) s.img[k++] = int(ch * c.n_pol + p);
      s.w[s.n_terms] = float(gw);
      ++s.n_terms;
    }
  }
  if (!(weight_sum > 0.0) || s.n_terms == 0) return ZeroSpec();
  s.factor = float(std::sqrt(c.pol_norm) / weight_sum);
  return s;
}

IntegSpec MakeIntegratedPsfSpec(const DeviceCube& c) {  // image_set.cc:499-530
  IntegSpec s{};
  if (c.n_channels == 1) {
    s.mode = kIntegCopy;
    s.n_terms = 1;
    s.img[0] = 0;
    return s;
  }
  s.mode = kIntegLinear;
  double weight_sum = 0.0;
  for (size_t ch = 0; ch != c.n_channels; ++ch) {
    const double gw = c.weights[ch];
    if (gw == 0.0) continue;
    weight_sum += gw;
    s.img[s.n_terms] = int(ch);
    s.w[s.n_terms] = float(gw);
    ++s.n_terms;
  }
  if (s.n_terms == 0) return ZeroSpec();
  s.factor = float(weight_sum == 0.0 ? 0.0 : 1.0 / weight_sum);
  return s;
}

bool IsSimpleSpec(const IntegSpec& s, size_t n_images) {
  if (s.mode == kIntegCopy) return n_images == 1;
  if (s.mode != kIntegLinear || size_t(s.n_terms) != n_images) return false;
  for (int k = 0; k != s.n_terms; ++k)
    if (s.img[k] != k) return false;
  return true;
}

// Weighted polynomial least squares over x_i = nu_i/nu_ref - 1 (nu_ref = freqs[n/2]),
// evaluated back at every channel: P = X (X^T W X)^-1 X^T W. Terms are clamped to the
// number of channels with weight > 0. Fixed operation order (normal equations,
// Gaussian elimination with partial pivoting) so the CPU checker and this library
// agree to the bit.
namespace {
// pw = n x t powers of x_i; returns the t x n operator (X^T W X)^-1 X^T W that maps channel
// values to polynomial terms (t = 0: no channel carries weight).
std::vector<double> SolveTerms(size_t n_terms, const std::vector<double>& freqs,
                               const std::vector<float>& weights, std::vector<double>& pw,
                               size_t& t_used) {
  const size_t n = freqs.size();
  std::vector<size_t> used;
  for (size_t i = 0; i != n; ++i)
    if (weights[i] > 0.0f) used.push_back(i);
  t_used = n_terms < used.size() ? n_terms : used.size();
  const size_t t = t_used;
  if (t == 0) return {};
  const double ref = freqs[n / 2];
  pw.assign(n * t, 0.0);
  for (size_t i = 0; i != n; ++i) {
    const double x = freqs[i] / ref - 1.0;
    double p = 1.0;
    for (size_t k = 0; k != t; ++k) {
      pw[i * t + k] = p;
      p = p * x;
    }
  }
  std::vector<double> nm(t * t, 0.0), rhs(t * n, 0.0);
  for (size_t a = 0; a != t; ++a) {
    for (size_t b = 0; b != t; ++b) {
      double s = 0.0;
      for (size_t u : used) s = s + double(weights[u]) * pw[u * t + a] * pw[u * t + b];
      nm[a * t + b] = s;
    }
    for (size_t u : used) rhs[a * n + u] = double(weights[u]) * pw[u * t + a];
  }
  for (size_t col = 0; col != t; ++col) {
    size_t piv = col;
    for (size_t r = col + 1; r != t; ++r)
      if (std::fabs(nm[r * t + col]) > std::fabs(nm[piv * t + col])) piv = r;
    if (piv != col) {
      for (size_t c = 0; c != t; ++c) std::swap(nm[col * t + c], nm[piv * t + c]);
      for (size_t c = 0; c != n; ++c) std::swap(rhs[col * n + c], rhs[piv * n + c]);
    }
    const double d = nm[col * t + col];
    for (size_t r = col + 1; r != t; ++r) {
      const double f = nm[r * t + col] / d;
      for (size_t c = col; c != t; ++c) nm[r * t + c] = nm[r * t + c] - f * nm[col * t + c];
      for (size_t c = 0; c != n; ++c) rhs[r * n + c] = rhs[r * n + c] - f * rhs[col * n + c];
    }
  }
  for (size_t col = t; col-- > 0;) {
    const double d = nm[col * t + col];
    for (size_t c = 0; c != n; ++c) {
      double s = rhs[col * n + c];
      for (size_t k = col + 1; k != t; ++k) s = s - nm[col * t + k] * rhs[k * n + c];
      rhs[col * n + c] = s / d;
    }
  }
  return rhs;
}
}  // namespace

std::vector<double> PolynomialTermSolve(size_t n_terms, const std::vector<double>& freqs,
                                        const std::vector<float>& weights, size_t& t_used) {
  std::vector<double> pw;
  return SolveTerms(n_terms, freqs, weights, pw, t_used);
}

std::vector<double> PolynomialProjection(size_t n_terms, const std::vector<double>& freqs,
                                         const std::vector<float>& weights) {
  const size_t n = freqs.size();
  std::vector<double> proj(n * n, 0.0), pw;
  size_t t = 0;
  const std::vector<double> rhs = SolveTerms(n_terms, freqs, weights, pw, t);
  if (t == 0) return proj;
  for (size_t i = 0; i != n; ++i)
    for (size_t j = 0; j != n; ++j) {
      double s = 0.0;
      for (size_t k = 0; k != t; ++k) s = s + pw[i * t + k] * rhs[k * n + j];
      proj[i * n + j] = s;
    }
  return proj;
}

}  // namespace rdl
