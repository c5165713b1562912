// ORACLE — TEST INFRASTRUCTURE ONLY.
// Definitions the reference's algorithm sources link against but whose own sources cannot be
// built here: math/component_optimization.cc needs GSL. The optimisers are only reached when
// component_optimization_algorithm != kClean (off by default, settings.h:370-371; out of
// scope per SURVEY 2 row 20), so they throw.
#include <stdexcept>

#include "math/component_optimization.h"

namespace radler::math {
namespace {
[[noreturn]] void Unavailable() {
  throw std::runtime_error("oracle/_ref: component optimisers need GSL, which is not available");
}
}  // namespace

aocommon::Image LinearComponentSolve(const std::vector<std::pair<size_t, size_t>>&,
                                     const aocommon::Image&, const aocommon::Image&) {
  Unavailable();
}
void LinearComponentSolve(aocommon::Image&, const aocommon::Image&, const aocommon::Image&) {
  Unavailable();
}
aocommon::Image GradientDescent(const std::vector<std::pair<size_t, size_t>>&,
                                const aocommon::Image&, const aocommon::Image&, size_t, size_t,
                                bool) {
  Unavailable();
}
void GradientDescent(aocommon::Image&, const aocommon::Image&, const aocommon::Image&, size_t,
                     size_t, bool) {
  Unavailable();
}
std::vector<aocommon::Image> GradientDescentWithVariablePsf(
    const std::vector<std::vector<std::pair<size_t, size_t>>>&, const aocommon::Image&,
    const std::vector<aocommon::Image>&, size_t, size_t, bool) {
  Unavailable();
}
}  // namespace radler::math

// ---- PythonDeconvolution (embedded interpreter, out of scope) and CasaMaskReader (casacore) --
// radler.cc names both; their own sources need pybind11-embed / casacore. Dummy complete types
// stand in for the pybind11 classes the header only forward-declares.
namespace pybind11 {
class scoped_interpreter {};
class gil_scoped_release {};
class function {};
}  // namespace pybind11

#include "algorithms/python_deconvolution.h"
#include "utils/casa_mask_reader.h"

namespace radler::algorithms {
PythonDeconvolution::PythonDeconvolution(const std::string& filename) : _filename(filename) {
  throw std::runtime_error("oracle/_ref: PythonDeconvolution is not available");
}
PythonDeconvolution::PythonDeconvolution(const PythonDeconvolution& other)
    : DeconvolutionAlgorithm(other), _filename(other._filename) {}
PythonDeconvolution::~PythonDeconvolution() = default;
DeconvolutionResult PythonDeconvolution::ExecuteMajorIteration(ImageSet&, ImageSet&,
                                                               const std::vector<aocommon::Image>&) {
  throw std::runtime_error("oracle/_ref: PythonDeconvolution is not available");
}
}  // namespace radler::algorithms

namespace radler::utils {
CasaMaskReader::CasaMaskReader(const std::string& path)
    : _path(path), _width(0), _height(0), _nPolarizations(0), _nChannels(0) {
  throw std::runtime_error("oracle/_ref: casacore masks are not available");
}
void CasaMaskReader::Read(bool*) {}
}  // namespace radler::utils
