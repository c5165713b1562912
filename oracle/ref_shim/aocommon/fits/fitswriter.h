// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for aocommon/fits/fitswriter.h (ours): FITS
// output is debug-only on the compiled path (IuwtDecomposition::Save); writing throws.
#ifndef AOCOMMON_FITS_FITSWRITER_H_
#define AOCOMMON_FITS_FITSWRITER_H_
#include <cstddef>
#include <stdexcept>
#include <string>
namespace aocommon {
class FitsWriter {
 public:
  void SetImageDimensions(size_t, size_t) {}
  void SetImageDimensions(size_t, size_t, double, double) {}
  void WriteMask(const std::string&, const bool*) const {
    throw std::runtime_error("oracle/_ref: FITS output is not available");
  }
  template <typename T>
  void Write(const std::string&, const T*) const {
    throw std::runtime_error("oracle/_ref: FITS output is not available");
  }
};
}  // namespace aocommon
#endif
