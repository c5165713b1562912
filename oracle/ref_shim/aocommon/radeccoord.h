// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for aocommon/radeccoord.h (ours); only the
// sky-model text writer (out of scope) formats coordinates.
#ifndef AOCOMMON_RADECCOORD_H_
#define AOCOMMON_RADECCOORD_H_
#include <cmath>
#include <cstdio>
#include <string>
namespace aocommon {
class RaDecCoord {
 public:
  static std::string RAToString(double ra, char sep) {
    double h = std::fmod(ra * 12.0 / M_PI + 48.0, 24.0);
    const int hh = int(h);
    const double m = (h - hh) * 60.0;
    const int mm = int(m);
    char buf[64];
    std::snprintf(buf, sizeof buf, "%02d%c%02d%c%06.3f", hh, sep, mm, sep, (m - mm) * 60.0);
    return buf;
  }
  static std::string DecToString(double dec, char sep) {
    const double d = std::fabs(dec) * 180.0 / M_PI;
    const int dd = int(d);
    const double m = (d - dd) * 60.0;
    const int mm = int(m);
    char buf[64];
    std::snprintf(buf, sizeof buf, "%c%02d%c%02d%c%06.3f", dec < 0 ? '-' : '+', dd, sep, mm, sep,
                  (m - mm) * 60.0);
    return buf;
  }
};
}  // namespace aocommon
#endif
