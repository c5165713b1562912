// ORACLE — TEST INFRASTRUCTURE ONLY. Stand-in for aocommon/fits/fitsreader.h (ours): there is
// no CFITSIO here, so opening a FITS file throws. Only file-based masks, forced spectra, RMS
// images and the external MORESANE program (all out of scope) read FITS.
#ifndef AOCOMMON_FITS_FITSREADER_H_
#define AOCOMMON_FITS_FITSREADER_H_
#include <cstddef>
#include <stdexcept>
#include <string>
namespace aocommon {
class FitsReader {
 public:
  explicit FitsReader(const std::string& filename, bool = true, bool = false) {
    throw std::runtime_error("oracle/_ref: FITS input is not available (" + filename + ")");
  }
  size_t ImageWidth() const { return 0; }
  size_t ImageHeight() const { return 0; }
  size_t NImages() const { return 0; }
  size_t NFrequencies() const { return 0; }
  template <typename T>
  void Read(T*) {}
  template <typename T>
  void ReadIndex(T*, size_t) {}
};
}  // namespace aocommon
#endif
