// ORACLE — TEST INFRASTRUCTURE ONLY.
//
// CPU restatement of the Radler minor-loop hot path (reference tree
// ska-sdp-func-radler @ a27db3b, paths below are relative to its cpp/ dir).
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
// reference leg may load this library. The product (ska_sdp_func_radler_b200/)
// never links or calls it.
//
// Parity pinning: the restatement is checked against the reference's own
// known-answer tests (tests/test_oracle_*.py re-encode test_peak_finder.cc,
// test_image_set.cc, test_fft_size_calculations.cc, test_radler.cc point-source
// fixture, test_dijkstra_splitter.cc) and against oracle/_ref, which compiles
// the reference's *own* leaf sources (peak_finder.cc, simple_clean.cc,
// dijkstra_splitter.cc, ...) where they lie under /root/reference.
// Third-party arithmetic that is NOT in the reference tree (aocommon Image ops,
// schaapcommon Convolve / SpectralFitter, FFTW) is restated from call sites;
// for those pieces bit-level parity is UNPINNED.
#ifndef RDL_ORACLE_H_
#define RDL_ORACLE_H_

#include <cmath>
#include <cstddef>
#include <cstdint>
#include <functional>
#include <memory>
#include <optional>
#include <string>
#include <utility>
#include <vector>

namespace oracle {

// ---------------------------------------------------------------------------
// aocommon::Image subset (row-major float, x fastest). Restated from call
// sites, SURVEY Appendix B.
// ---------------------------------------------------------------------------
struct Image {
  size_t w = 0, h = 0;
  std::vector<float> d;
  Image() = default;
  Image(size_t w_, size_t h_, float v = 0.0f) : w(w_), h(h_), d(w_ * h_, v) {}
  size_t Size() const { return d.size(); }
  bool Empty() const { return d.empty(); }
  float* Data() { return d.data(); }
  const float* Data() const { return d.data(); }
  float& operator[](size_t i) { return d[i]; }
  float operator[](size_t i) const { return d[i]; }
};

// Centred crop / zero-pad, start = (outer - inner) / 2 (aocommon::Image::Trim/Untrim).
void Trim(float* dst, size_t dw, size_t dh, const float* src, size_t sw, size_t sh);
void Untrim(float* dst, size_t dw, size_t dh, const float* src, size_t sw, size_t sh);

// ---------------------------------------------------------------------------
// Thread pool stand-in for aocommon::StaticFor / DynamicFor.
// ---------------------------------------------------------------------------
void SetThreadCount(size_t n);
size_t ThreadCount();
// Contiguous chunk per thread (StaticFor::Run(a, b, f(lo, hi))).
void StaticFor(size_t begin, size_t end, const std::function<void(size_t, size_t)>& f);
// One task per index (DynamicFor::Run(a, b, f(i))).
void DynamicFor(size_t begin, size_t end, const std::function<void(size_t)>& f);

// ---------------------------------------------------------------------------
// Spectral fitter (schaapcommon::fitters::SpectralFitter restatement;
// kNoFitting and kPolynomial only).
// ---------------------------------------------------------------------------
enum class SpectralMode : int { kNoFitting = 0, kPolynomial = 1 };

class SpectralFitter {
 public:
  SpectralFitter() = default;
  SpectralFitter(SpectralMode mode, size_t n_terms, const std::vector<double>& freqs,
                 const std::vector<float>& weights);
  SpectralMode Mode() const { return mode_; }
  size_t NChannels() const { return freqs_.size(); }
  // values[0..n) in -> fitted & evaluated values out (FitAndEvaluate).
  void FitAndEvaluate(float* values) const;
  // Row-major n x n projection matrix (values_out = M * values_in).
  const std::vector<double>& Projection() const { return proj_; }

 private:
  SpectralMode mode_ = SpectralMode::kNoFitting;
  size_t n_terms_ = 0;
  std::vector<double> freqs_;
  std::vector<float> weights_;
  std::vector<double> proj_;
};

// Shared deterministic recipe for the polynomial projection matrix (documented
// in DESIGN.md "spectral fit"): normal equations in double, Gaussian elimination
// with partial pivoting, fixed operation order.
std::vector<double> PolynomialProjection(size_t n_terms, const std::vector<double>& freqs,
                                         const std::vector<float>& weights);
// The t_used x n operator of the same fit that maps channel values to polynomial terms, and the
// evaluation sum_k terms[k] x^k by repeated multiplication (SpectralFitter::Fit / Evaluate as
// ImageSet::InterpolateAndStoreModel uses them, image_set.cc:236-283).
std::vector<double> PolynomialTermSolve(size_t n_terms, const std::vector<double>& freqs,
                                        const std::vector<float>& weights, size_t& t_used);
double EvaluatePolynomial(const float* terms, size_t n_terms, double x);

// ---------------------------------------------------------------------------
// ImageSet (image_set.{h,cc}) — only what the algorithms use.
// images ordered [dec_channel][pol]; PsfIndex(i) = i / n_pol.
// ---------------------------------------------------------------------------
struct CubeMeta {
  size_t n_channels = 1;            // deconvolution channels (= PsfCount)
  size_t n_pol = 1;                 // images per channel
  std::vector<float> weights;       // per channel (ImageSet::weights_ is vector<float>)
  std::vector<uint8_t> pol_linked;  // per pol; all 1 when linked_polarizations is empty
  float pol_norm = 1.0f;            // polarization_normalization_factor_
  bool squared_joins = false;
};

class ImageSet {
 public:
  ImageSet() = default;
  ImageSet(const CubeMeta& meta, size_t w, size_t h);
  size_t Size() const { return images.size(); }
  size_t Width() const { return w; }
  size_t Height() const { return h; }
  size_t PsfCount() const { return meta.n_channels; }
  size_t PsfIndex(size_t i) const { return i / meta.n_pol; }
  float* Data(size_t i) { return images[i].Data(); }
  const float* Data(size_t i) const { return images[i].Data(); }
  Image& operator[](size_t i) { return images[i]; }
  const Image& operator[](size_t i) const { return images[i]; }

  void GetLinearIntegrated(Image& dest) const;
  void GetSquareIntegrated(Image& dest, Image& scratch) const;
  void GetIntegratedPsf(Image& dest, const std::vector<Image>& psfs) const;

  CubeMeta meta;
  size_t w = 0, h = 0;
  std::vector<Image> images;

 private:
  void LinearNormal(Image& dest) const;
  void SquareNormal(Image& dest, Image& scratch) const;
  void SquareSquared(Image& dest) const;
};

// ---------------------------------------------------------------------------
// peak_finder (math/peak_finder.{h,cc}: Simple + FindWithMask semantics).
// Returns nullopt when nothing exceeds FLT_MIN.
// ---------------------------------------------------------------------------
std::optional<float> FindPeak(const float* image, size_t width, size_t height, size_t& x,
                              size_t& y, bool allow_negative, size_t start_y, size_t end_y,
                              size_t h_border, size_t v_border, const bool* mask = nullptr);
std::optional<float> FindPeakRatio(const float* image, size_t width, size_t height, size_t& x,
                                   size_t& y, bool allow_negative, size_t start_y,
                                   size_t end_y, float border_ratio,
                                   const bool* mask = nullptr);

// ---------------------------------------------------------------------------
// simple_clean (algorithms/simple_clean.cc) + threaded wrapper.
// ---------------------------------------------------------------------------
void SubtractImageWhole(float* image, const float* psf, size_t width, size_t height, size_t x,
                        size_t y, float factor);  // SubtractImage, :17-59 (drops odd column)
void PartialSubtractImage(float* image, const float* psf, size_t width, size_t height,
                          size_t x, size_t y, float factor, size_t start_y, size_t end_y);
void PartialSubtractImage(float* image, size_t image_width, size_t image_height,
                          const float* psf, size_t psf_width, size_t psf_height, size_t x,
                          size_t y, float factor, size_t start_y, size_t end_y);
void ThreadedSubtractImage(float* image, const Image& psf, size_t x, size_t y, float factor);

// ---------------------------------------------------------------------------
// FFT convolution (schaapcommon::math restatement).
// ---------------------------------------------------------------------------
void PrepareSmallConvolutionKernel(float* dst, size_t w, size_t h, const float* kernel,
                                   size_t n);
void PrepareConvolutionKernel(float* dst, const float* src, size_t w, size_t h);
void Convolve(float* image, const float* kernel, size_t w, size_t h);
// 1-D complex FFT of any length (mixed radix 2/3/4/5/7 + Bluestein); exposed for tests.
void Fft1D(std::vector<std::pair<float, float>>& data, bool inverse);

// utils/fft_size_calculations.h
size_t CalculateGoodFFTSize(size_t minimum_size);
size_t GetConvolutionSize(double convolution_scale, size_t original_size, double padding);

// ---------------------------------------------------------------------------
// Trace of peaks, for bit-exact sequence comparison with the CUDA path.
// ---------------------------------------------------------------------------
struct TraceEntry {
  uint32_t iteration;
  uint32_t scale_index;
  uint32_t x, y;
  float value;  // integrated peak value (signed)
};

// ---------------------------------------------------------------------------
// DeconvolutionAlgorithm (algorithms/deconvolution_algorithm.{h,cc}).
// ---------------------------------------------------------------------------
struct DeconvolutionResult {
  std::optional<float> starting_peak_value;
  float final_peak_value = 0.0f;
  bool another_iteration_required = false;
  bool is_diverging = false;
};

class DeconvolutionAlgorithm {
 public:
  virtual ~DeconvolutionAlgorithm() = default;
  virtual DeconvolutionResult ExecuteMajorIteration(ImageSet& data, ImageSet& model,
                                                    const std::vector<Image>& psfs) = 0;
  // Base-class defaults, deconvolution_algorithm.h:189-202
  float threshold = 0.0f;
  float major_iteration_threshold = 0.0f;
  float minor_loop_gain = 0.1f;
  float major_loop_gain = 1.0f;
  float clean_border_ratio = 0.05f;
  size_t max_iterations = 500;
  float divergence_limit = 4.0f;
  bool allow_negative_components = true;
  bool stop_on_negative_component = false;
  const bool* clean_mask = nullptr;
  Image rms_factor_image;
  size_t iteration_number = 0;
  size_t n_polarizations = 1;
  SpectralFitter fitter;

  void PerformSpectralFit(float* values) const;  // deconvolution_algorithm.cc:29-46

  std::vector<TraceEntry> trace;
  bool record_trace = false;
  size_t trace_limit = 1u << 22;
  void Record(uint32_t scale, size_t x, size_t y, float v) {
    if (record_trace && trace.size() < trace_limit)
      trace.push_back({uint32_t(iteration_number), scale, uint32_t(x), uint32_t(y), v});
  }
};

// ---------------------------------------------------------------------------
// SubMinorLoop (algorithms/subminor_loop.{h,cc}).
// ---------------------------------------------------------------------------
class SubMinorLoop {
 public:
  SubMinorLoop(size_t width, size_t height, size_t padded_width, size_t padded_height)
      : width_(width), height_(height), padded_width_(padded_width),
        padded_height_(padded_height) {}
  float threshold = 0.0f, gain = 0.0f;
  size_t h_border = 0, v_border = 0;
  size_t current_iteration = 0, max_iterations = 0;
  bool allow_negative = true, stop_on_negative = false;
  const bool* mask = nullptr;
  float divergence_limit = 0.0f;
  DeconvolutionAlgorithm* parent = nullptr;
  uint32_t trace_scale = 0;
  Image rms_factor_image;
  float flux_cleaned = 0.0f;

  std::pair<bool, std::optional<float>> Run(ImageSet& convolved_residual,
                                            const std::vector<Image>& twice_convolved_psfs);
  void CorrectResidualDirty(float* scratch_a, float* scratch_b, float* scratch_c,
                            size_t image_index, float* residual,
                            const float* single_convolved_psf) const;
  void GetFullIndividualModel(size_t image_index, float* out) const;
  void UpdateAutoMask(bool* mask_out) const;
  size_t NSelected() const { return positions_.size(); }
  const std::vector<std::pair<size_t, size_t>>& Positions() const { return positions_; }
  const ImageSet& Model() const { return model_; }

 private:
  size_t GetMaxComponent(Image& scratch, float& max_value) const;
  void FindPeakPositions(ImageSet& convolved_residual);
  size_t width_, height_, padded_width_, padded_height_;
  std::vector<std::pair<size_t, size_t>> positions_;
  ImageSet residual_, model_;
  Image compact_rms_;
};

// ---------------------------------------------------------------------------
// GenericClean (algorithms/generic_clean.{h,cc}).
// ---------------------------------------------------------------------------
class GenericClean : public DeconvolutionAlgorithm {
 public:
  explicit GenericClean(bool use_sub_minor_optimization)
      : use_sub_minor_optimization_(use_sub_minor_optimization) {}
  DeconvolutionResult ExecuteMajorIteration(ImageSet& dirty, ImageSet& model,
                                            const std::vector<Image>& psfs) override;
  size_t last_n_selected = 0;

 private:
  std::optional<float> FindPeak(const Image& image, float* scratch, size_t& x, size_t& y);
  float convolution_padding_ = 1.1f;
  bool use_sub_minor_optimization_;
};

// ---------------------------------------------------------------------------
// MultiScale (algorithms/multiscale_algorithm.{h,cc}, multiscale/multiscale_transforms.*).
// ---------------------------------------------------------------------------
enum class MultiscaleShape : int { kTaperedQuadratic = 0, kGaussian = 1 };

Image MakeShapeFunction(float scale_in_pixels, size_t& n, size_t max_n, MultiscaleShape shape);
float KernelPeakValue(double scale_in_pixels, size_t max_n, MultiscaleShape shape);
void AddShapeComponent(float* image, size_t width, size_t height, float scale, size_t x,
                       size_t y, float gain, MultiscaleShape shape);

class MultiScaleTransforms {
 public:
  MultiScaleTransforms(size_t w, size_t h, MultiscaleShape shape)
      : width_(w), height_(h), shape_(shape) {}
  void Transform(std::vector<Image*>& images, Image& scratch, float scale);
  void Transform(Image& image, Image& scratch, float scale) {
    std::vector<Image*> l{&image};
    Transform(l, scratch, scale);
  }
  size_t Width() const { return width_; }
  size_t Height() const { return height_; }

 private:
  size_t width_, height_;
  MultiscaleShape shape_;
};

struct MultiscaleSettings {  // settings.h:465-524
  bool fast_sub_minor_loop = true;
  double sub_minor_loop_gain = 0.2;
  double scale_bias = 0.6;
  size_t max_scales = 0;
  double convolution_padding = 1.1;
  std::vector<double> scale_list;
  MultiscaleShape shape = MultiscaleShape::kTaperedQuadratic;
};

struct ScaleInfo {  // multiscale_algorithm.h:58-77
  float scale = 0.0f, psf_peak = 0.0f, kernel_peak = 0.0f, bias_factor = 0.0f, gain = 0.0f;
  float max_normalized_image_value = 0.0f, max_unnormalized_image_value = 0.0f, rms = 0.0f;
  size_t max_image_value_x = 0, max_image_value_y = 0;
  bool is_active = false;
  size_t n_components_cleaned = 0;
  float total_flux_cleaned = 0.0f;
};

class MultiScaleAlgorithm : public DeconvolutionAlgorithm {
 public:
  MultiScaleAlgorithm(const MultiscaleSettings& settings, double beam_size,
                      double pixel_scale_x, double pixel_scale_y);
  DeconvolutionResult ExecuteMajorIteration(ImageSet& data, ImageSet& model,
                                            const std::vector<Image>& psfs) override;
  void SetAutoMaskMode(bool track, bool use) {
    track_per_scale_masks_ = track;
    use_per_scale_masks_ = use;
  }
  std::vector<ScaleInfo> scale_infos;
  std::vector<std::vector<uint8_t>> scale_masks;  // bool per pixel
  size_t n_minor_iterations = 0;                  // outer (scale-selection) loop count

 private:
  void FindActiveScaleConvolvedMaxima(const ImageSet& image_set, Image& integrated_scratch,
                                      Image& scratch, bool report_rms);
  void ActivateScales(size_t scale_with_last_peak);
  void FindPeakDirect(const Image& image, Image& scratch, size_t scale_index);
  void AddComponentToModel(ImageSet& model, size_t image_index, size_t scale_with_peak,
                           float component_value);
  MultiscaleSettings settings_;
  double beam_size_in_pixels_;
  bool track_per_scale_masks_ = false, use_per_scale_masks_ = false;
};

// ---- IUWT decomposition (iuwt.cc) -----------------------------------------------------------
void IuwtSmooth(float* out, const float* in, float* scratch, size_t width, size_t height, int scale,
                bool mt);
void IuwtDecompose(const float* input, size_t width, size_t height, int n_scales,
                   bool include_largest, float* coefficients);
void IuwtRecompose(const float* coefficients, size_t width, size_t height, int n_scales,
                   bool include_largest, const uint8_t* masks, float* out);


// ---- IUWT structure selection and plane measurements (iuwt_analysis.cc) ---------------------
// masks: n_scales planes of width*height bytes (in/out); return value = voxels newly filled
size_t IuwtFloodfill(const float* coefficients, uint8_t* masks, size_t width, size_t height,
                     int n_scales, const float* thresholds, int min_scale, int end_scale,
                     size_t seed_x, size_t seed_y, int seed_scale, float clean_border,
                     const uint8_t* prior_mask);
size_t IuwtSelectStructures(const float* coefficients, uint8_t* masks, size_t width, size_t height,
                            int n_scales, const float* thresholds, int min_scale, int end_scale,
                            float clean_border, const uint8_t* prior_mask);
size_t IuwtFloodfill2D(const float* image, uint8_t* mask, size_t width, size_t height,
                       float threshold, bool use_abs, size_t seed_x, size_t seed_y);
float IuwtMad(const float* image, size_t n);
float IuwtMaxAbs(const float* image, size_t width, size_t height, float clean_border,
                 bool allow_negative, const uint8_t* mask, size_t& x, size_t& y);

}  // namespace oracle
#endif
