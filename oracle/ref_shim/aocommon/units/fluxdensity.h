// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for aocommon/units/fluxdensity.h (ours);
// only feeds log lines.
#ifndef AOCOMMON_UNITS_FLUXDENSITY_H_
#define AOCOMMON_UNITS_FLUXDENSITY_H_
#include <cmath>
#include <sstream>
#include <string>
namespace aocommon::units {
class FluxDensity {
 public:
  static std::string ToNiceString(double jy) {
    std::ostringstream s;
    const double a = std::fabs(jy);
    if (a >= 1.0 || a == 0.0) s << jy << " Jy";
    else if (a >= 1e-3) s << jy * 1e3 << " mJy";
    else if (a >= 1e-6) s << jy * 1e6 << " µJy";
    else s << jy * 1e9 << " nJy";
    return s.str();
  }
};
}  // namespace aocommon::units
#endif
