// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for schaapcommon/fitters/spectralfitter.h
// (ours). kNoFitting and kPolynomial only; the polynomial fit is the oracle's documented
// recipe (oracle/core.cc PolynomialProjection: weighted LSQ of sum_k t_k (nu/nu_ref - 1)^k,
// nu_ref = freqs[n/2], terms clamped to the number of positive-weight channels), so the
// reference's own sources and the restated oracle see identical fitted values.
#ifndef SCHAAPCOMMON_FITTERS_SPECTRALFITTER_H_
#define SCHAAPCOMMON_FITTERS_SPECTRALFITTER_H_
#include <algorithm>
#include <cmath>
#include <cstddef>
#include <stdexcept>
#include <vector>

#include <aocommon/image.h>

#include "../../../oracle.h"

namespace schaapcommon::fitters {

enum class SpectralFittingMode { kNoFitting, kPolynomial, kLogPolynomial, kForcedTerms };

class SpectralFitter {
 public:
  using NumT = float;
  SpectralFitter(SpectralFittingMode mode, size_t n_terms, std::vector<double> frequencies = {},
                 std::vector<NumT> weights = {})
      : mode_(mode), n_terms_(n_terms), frequencies_(std::move(frequencies)),
        weights_(std::move(weights)) {
    if (mode_ == SpectralFittingMode::kLogPolynomial || mode_ == SpectralFittingMode::kForcedTerms)
      throw std::runtime_error("oracle/_ref: only kNoFitting and kPolynomial fitting are restated");
    if (weights_.empty()) weights_.assign(frequencies_.size(), 1.0f);
    if (mode_ == SpectralFittingMode::kPolynomial && !frequencies_.empty()) {
      fitter_ = oracle::SpectralFitter(oracle::SpectralMode::kPolynomial, n_terms_, frequencies_,
                                       weights_);
      BuildTermSolve();
    }
  }
  SpectralFittingMode Mode() const { return mode_; }
  size_t NTerms() const { return n_terms_; }
  const std::vector<double>& Frequencies() const { return frequencies_; }
  const std::vector<NumT>& Weights() const { return weights_; }
  double ReferenceFrequency() const { return frequencies_[frequencies_.size() / 2]; }
  void SetForcedTerms(std::vector<aocommon::Image>&&) {
    throw std::runtime_error("oracle/_ref: forced spectral terms are not restated");
  }

  void FitAndEvaluate(NumT* values, size_t /*x*/, size_t /*y*/,
                      std::vector<NumT>& /*scratch*/) const {
    if (mode_ == SpectralFittingMode::kNoFitting) return;
    fitter_.FitAndEvaluate(values);
  }
  // terms = T * values (T from the same normal equations as the projection)
  void Fit(std::vector<NumT>& terms, const NumT* values, size_t /*x*/, size_t /*y*/) const {
    const size_t n = frequencies_.size();
    if (mode_ == SpectralFittingMode::kNoFitting) {
      terms.assign(values, values + n);
      return;
    }
    terms.assign(n_terms_, 0.0f);
    for (size_t k = 0; k != t_used_; ++k) {
      double s = 0.0;
      for (size_t j = 0; j != n; ++j) s = std::fma(term_solve_[k * n + j], double(values[j]), s);
      terms[k] = float(s);
    }
  }
  NumT Evaluate(const std::vector<NumT>& terms, double frequency) const {
    if (mode_ == SpectralFittingMode::kNoFitting)
      throw std::runtime_error("Evaluate() needs a fitting mode");
    const double x = frequency / ReferenceFrequency() - 1.0;
    double p = 1.0, s = 0.0;
    for (size_t k = 0; k != terms.size(); ++k) {
      s = s + double(terms[k]) * p;
      p = p * x;
    }
    return NumT(s);
  }
  void Evaluate(NumT* values, const std::vector<NumT>& terms) const {
    for (size_t i = 0; i != frequencies_.size(); ++i) values[i] = Evaluate(terms, frequencies_[i]);
  }

 private:
  void BuildTermSolve();
  SpectralFittingMode mode_;
  size_t n_terms_;
  std::vector<double> frequencies_;
  std::vector<NumT> weights_;
  oracle::SpectralFitter fitter_;
  std::vector<double> term_solve_;  // t_used_ x n
  size_t t_used_ = 0;
};

// Weighted normal equations solved column by column for the unit vectors: the t x n operator
// that maps channel values to polynomial terms.
inline void SpectralFitter::BuildTermSolve() {
  const size_t n = frequencies_.size();
  std::vector<size_t> used;
  for (size_t i = 0; i != n; ++i)
    if (weights_[i] > 0.0f) used.push_back(i);
  const size_t t = std::min(n_terms_, used.size());
  t_used_ = t;
  term_solve_.assign(t * n, 0.0);
  if (t == 0) return;
  const double ref = ReferenceFrequency();
  std::vector<double> xp(n * t);
  for (size_t i = 0; i != n; ++i) {
    const double x = frequencies_[i] / ref - 1.0;
    double p = 1.0;
    for (size_t k = 0; k != t; ++k) {
      xp[i * t + k] = p;
      p = p * x;
    }
  }
  std::vector<double> N(t * t, 0.0), B(t * n, 0.0);
  for (size_t a = 0; a != t; ++a) {
    for (size_t b = 0; b != t; ++b) {
      double s = 0.0;
      for (size_t u : used) s = s + double(weights_[u]) * xp[u * t + a] * xp[u * t + b];
      N[a * t + b] = s;
    }
    for (size_t u : used) B[a * n + u] = double(weights_[u]) * xp[u * t + a];
  }
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
      for (size_t c = col; c != t; ++c) N[r * t + c] -= f * N[col * t + c];
      for (size_t c = 0; c != n; ++c) B[r * n + c] -= f * B[col * n + c];
    }
  }
  for (size_t col = t; col-- > 0;) {
    const double d = N[col * t + col];
    for (size_t c = 0; c != n; ++c) {
      double s = B[col * n + c];
      for (size_t k = col + 1; k != t; ++k) s -= N[col * t + k] * B[k * n + c];
      B[col * n + c] = s / d;
    }
  }
  term_solve_ = B;
}

}  // namespace schaapcommon::fitters
#endif
