// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for schaapcommon/fitters/gaussianfitter.h
// (ours). On the IUWT path the fitted beam shapes only land in ScaleResponse::b_major /
// b_minor / b_pa / convolved_area, which nothing reads, so the fits are
// replaced by no-ops that leave their in/out arguments untouched.
#ifndef SCHAAPCOMMON_FITTERS_GAUSSIANFITTER_H_
#define SCHAAPCOMMON_FITTERS_GAUSSIANFITTER_H_
#include <cstddef>

#include "../math/ellipse.h"

namespace schaapcommon::fitters {
inline math::Ellipse Fit2DGaussianCentred(const float*, size_t, size_t, double beam_estimate,
                                          double = 10.0, bool = false) {
  return math::Ellipse{beam_estimate, beam_estimate, 0.0};
}
inline void Fit2DGaussianFull(const float*, size_t, size_t, double&, double&, double&, double&,
                              double&, double&, double* floor_level = nullptr) {
  if (floor_level) *floor_level = 0.0;
}
inline math::Ellipse DeconvolveGaussian(const math::Ellipse& a, const math::Ellipse&) { return a; }
}  // namespace schaapcommon::fitters
#endif
