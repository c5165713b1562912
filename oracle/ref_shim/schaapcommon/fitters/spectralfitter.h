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
    return NumT(oracle::EvaluatePolynomial(terms.data(), terms.size(), x));
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

// The t x n operator that maps channel values to polynomial terms: the oracle leaf's, compiled
// without contraction (this header is compiled with the reference's flags, whose contractions
// would move the last bit of the operator and with it values that sit on float rounding ties).
inline void SpectralFitter::BuildTermSolve() {
  term_solve_ = oracle::PolynomialTermSolve(n_terms_, frequencies_, weights_, t_used_);
}

}  // namespace schaapcommon::fitters
#endif
