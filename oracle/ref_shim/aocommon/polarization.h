// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for aocommon/polarization.h (ours): the enum
// and the set predicates ImageSet::InitializePolFactor uses (image_set.h:296-322).
#ifndef AOCOMMON_POLARIZATION_H_
#define AOCOMMON_POLARIZATION_H_
#include <math.h>

#include <set>
#include <string>
namespace aocommon {
class Polarization {
 public:
  enum PolarizationEnum { StokesI, StokesQ, StokesU, StokesV, RR, RL, LR, LL, XX, XY, YX, YY,
                          FullStokes, Instrumental, DiagonalInstrumental };
  static bool IsStokes(PolarizationEnum p) {
    return p == StokesI || p == StokesQ || p == StokesU || p == StokesV;
  }
  static bool HasDualPolarization(const std::set<PolarizationEnum>& s) {
    return (s.count(XX) && s.count(YY)) || (s.count(RR) && s.count(LL));
  }
  static bool HasFullLinearPolarization(const std::set<PolarizationEnum>& s) {
    return s.count(XX) && s.count(XY) && s.count(YX) && s.count(YY);
  }
  static bool HasFullCircularPolarization(const std::set<PolarizationEnum>& s) {
    return s.count(RR) && s.count(RL) && s.count(LR) && s.count(LL);
  }
  static std::string TypeToShortString(PolarizationEnum p) {
    static const char* names[] = {"I", "Q", "U", "V", "RR", "RL", "LR", "LL", "XX", "XY", "YX",
                                  "YY", "IQUV", "instr", "diag_instr"};
    return names[int(p)];
  }
  static std::string TypeToFullString(PolarizationEnum p) { return TypeToShortString(p); }
};
using PolarizationEnum = Polarization::PolarizationEnum;
}  // namespace aocommon
#endif
