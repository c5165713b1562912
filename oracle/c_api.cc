// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h).
// Flat C entry points so tests/ and bench.py's cpu_baseline leg can drive the
// CPU restatement through ctypes. Not part of the product ABI.
#include <cstring>
#include <memory>
#include <stdexcept>

#include "oracle.h"

using namespace oracle;

extern "C" {

struct orc_cube {
  uint64_t width, height, n_channels, n_pol;
  const float* weights;       // n_channels or null (all 1)
  const uint8_t* pol_linked;  // n_pol or null (all linked)
  float pol_norm;
  int32_t squared_joins;
};

struct orc_settings {
  float threshold, major_iteration_threshold, minor_loop_gain, major_loop_gain;
  float clean_border_ratio, divergence_limit;
  uint64_t max_iterations, iteration_number;
  int32_t allow_negative, stop_on_negative;
  const uint8_t* clean_mask;  // width*height bools or null
  const float* rms_factor;    // width*height or null
  int32_t spectral_mode;      // 0 none, 1 polynomial
  uint64_t n_terms;
  const double* frequencies;  // n_channels
  const float* fit_weights;   // n_channels
  int32_t record_trace;
};

struct orc_result {
  int32_t has_starting_peak;
  float starting_peak_value, final_peak_value;
  int32_t another_iteration_required, is_diverging;
  uint64_t iteration_number;
  uint64_t n_trace;
  uint64_t n_selected;
};

struct orc_ms_config {
  int32_t fast_sub_minor_loop;
  double sub_minor_loop_gain, scale_bias;
  uint64_t max_scales;
  double convolution_padding;
  const double* scale_list;
  uint64_t n_scale_list;
  int32_t shape;
  double beam_size, pixel_scale_x, pixel_scale_y;
};

struct orc_scale_info {
  float scale, psf_peak, kernel_peak, bias_factor, gain;
  float max_normalized_image_value, max_unnormalized_image_value, rms;
  uint64_t max_x, max_y;
  int32_t is_active;
  uint64_t n_components_cleaned;
  float total_flux_cleaned;
};

struct orc_handle {
  std::unique_ptr<DeconvolutionAlgorithm> alg;
  int kind;  // 0 generic, 1 multiscale
  std::string error;
};

static CubeMeta ToMeta(const orc_cube& c) {
  CubeMeta m;
  m.n_channels = c.n_channels;
  m.n_pol = c.n_pol;
  if (c.weights) m.weights.assign(c.weights, c.weights + c.n_channels);
  if (c.pol_linked) m.pol_linked.assign(c.pol_linked, c.pol_linked + c.n_pol);
  m.pol_norm = c.pol_norm;
  m.squared_joins = c.squared_joins != 0;
  return m;
}

static void Apply(DeconvolutionAlgorithm& a, const orc_cube& c, const orc_settings& s) {
  a.threshold = s.threshold;
  a.major_iteration_threshold = s.major_iteration_threshold;
  a.minor_loop_gain = s.minor_loop_gain;
  a.major_loop_gain = s.major_loop_gain;
  a.clean_border_ratio = s.clean_border_ratio;
  a.divergence_limit = s.divergence_limit;
  a.max_iterations = s.max_iterations;
  a.iteration_number = s.iteration_number;
  a.allow_negative_components = s.allow_negative != 0;
  a.stop_on_negative_component = s.stop_on_negative != 0;
  a.clean_mask = reinterpret_cast<const bool*>(s.clean_mask);
  a.rms_factor_image = Image();
  if (s.rms_factor) {
    a.rms_factor_image = Image(c.width, c.height);
    std::memcpy(a.rms_factor_image.Data(), s.rms_factor, c.width * c.height * sizeof(float));
  }
  a.n_polarizations = c.n_pol;
  if (s.spectral_mode == 1) {
    std::vector<double> f(s.frequencies, s.frequencies + c.n_channels);
    std::vector<float> w(c.n_channels, 1.0f);
    if (s.fit_weights) w.assign(s.fit_weights, s.fit_weights + c.n_channels);
    a.fitter = SpectralFitter(SpectralMode::kPolynomial, s.n_terms, f, w);
  } else {
    a.fitter = SpectralFitter();
  }
  a.record_trace = s.record_trace != 0;
  a.trace.clear();
}

void orc_set_threads(uint64_t n) { SetThreadCount(n); }
uint64_t orc_get_threads() { return ThreadCount(); }

orc_handle* orc_create_generic(int32_t use_sub_minor_optimization) {
  auto* h = new orc_handle;
  h->alg = std::make_unique<GenericClean>(use_sub_minor_optimization != 0);
  h->kind = 0;
  return h;
}

orc_handle* orc_create_multiscale(const orc_ms_config* cfg) {
  auto* h = new orc_handle;
  MultiscaleSettings s;
  s.fast_sub_minor_loop = cfg->fast_sub_minor_loop != 0;
  s.sub_minor_loop_gain = cfg->sub_minor_loop_gain;
  s.scale_bias = cfg->scale_bias;
  s.max_scales = cfg->max_scales;
  s.convolution_padding = cfg->convolution_padding;
  if (cfg->scale_list) s.scale_list.assign(cfg->scale_list, cfg->scale_list + cfg->n_scale_list);
  s.shape = MultiscaleShape(cfg->shape);
  h->alg = std::make_unique<MultiScaleAlgorithm>(s, cfg->beam_size, cfg->pixel_scale_x,
                                                 cfg->pixel_scale_y);
  h->kind = 1;
  return h;
}

void orc_destroy(orc_handle* h) { delete h; }

const char* orc_last_error(orc_handle* h) { return h->error.c_str(); }

void orc_ms_set_automask_mode(orc_handle* h, int32_t track, int32_t use) {
  if (h->kind == 1) static_cast<MultiScaleAlgorithm*>(h->alg.get())->SetAutoMaskMode(track, use);
}

uint64_t orc_ms_scale_count(orc_handle* h) {
  return h->kind == 1 ? static_cast<MultiScaleAlgorithm*>(h->alg.get())->scale_infos.size() : 0;
}

uint64_t orc_ms_minor_iterations(orc_handle* h) {
  return h->kind == 1 ? static_cast<MultiScaleAlgorithm*>(h->alg.get())->n_minor_iterations : 0;
}

void orc_ms_scale_info(orc_handle* h, uint64_t i, orc_scale_info* out) {
  const ScaleInfo& s = static_cast<MultiScaleAlgorithm*>(h->alg.get())->scale_infos[i];
  *out = {s.scale, s.psf_peak, s.kernel_peak, s.bias_factor, s.gain,
          s.max_normalized_image_value, s.max_unnormalized_image_value, s.rms,
          s.max_image_value_x, s.max_image_value_y, s.is_active ? 1 : 0,
          s.n_components_cleaned, s.total_flux_cleaned};
}

// copies scale mask i (width*height bytes) into out; returns 0 if absent
int32_t orc_ms_scale_mask(orc_handle* h, uint64_t i, uint8_t* out) {
  auto* ms = static_cast<MultiScaleAlgorithm*>(h->alg.get());
  if (i >= ms->scale_masks.size()) return 0;
  std::memcpy(out, ms->scale_masks[i].data(), ms->scale_masks[i].size());
  return 1;
}

// residual/model: [n_channels*n_pol][height][width] contiguous, updated in place.
// psfs: [n_channels][height][width].
int32_t orc_execute(orc_handle* h, const orc_cube* cube, const orc_settings* settings,
                    float* residual, float* model, const float* psfs, orc_result* out) {
  try {
    const size_t n = cube->width * cube->height;
    const CubeMeta meta = ToMeta(*cube);
    ImageSet data(meta, cube->width, cube->height), mod(meta, cube->width, cube->height);
    for (size_t i = 0; i != data.Size(); ++i) {
      std::memcpy(data.Data(i), residual + i * n, n * sizeof(float));
      std::memcpy(mod.Data(i), model + i * n, n * sizeof(float));
    }
    std::vector<Image> psf_images;
    for (size_t i = 0; i != cube->n_channels; ++i) {
      psf_images.emplace_back(cube->width, cube->height);
      std::memcpy(psf_images.back().Data(), psfs + i * n, n * sizeof(float));
    }
    Apply(*h->alg, *cube, *settings);
    const DeconvolutionResult r = h->alg->ExecuteMajorIteration(data, mod, psf_images);
    for (size_t i = 0; i != data.Size(); ++i) {
      std::memcpy(residual + i * n, data.Data(i), n * sizeof(float));
      std::memcpy(model + i * n, mod.Data(i), n * sizeof(float));
    }
    out->has_starting_peak = r.starting_peak_value.has_value();
    out->starting_peak_value = r.starting_peak_value.value_or(0.0f);
    out->final_peak_value = r.final_peak_value;
    out->another_iteration_required = r.another_iteration_required;
    out->is_diverging = r.is_diverging;
    out->iteration_number = h->alg->iteration_number;
    out->n_trace = h->alg->trace.size();
    out->n_selected =
        h->kind == 0 ? static_cast<GenericClean*>(h->alg.get())->last_n_selected : 0;
    return 0;
  } catch (const std::exception& e) {
    h->error = e.what();
    return 1;
  }
}

uint64_t orc_get_trace(orc_handle* h, TraceEntry* out, uint64_t capacity) {
  const uint64_t n = std::min<uint64_t>(capacity, h->alg->trace.size());
  std::memcpy(out, h->alg->trace.data(), n * sizeof(TraceEntry));
  return n;
}

// ---- leaf functions, for KATs and kernel-level parity ---------------------------
int32_t orc_find_peak(const float* image, uint64_t w, uint64_t h, int32_t allow_negative,
                      uint64_t start_y, uint64_t end_y, uint64_t h_border, uint64_t v_border,
                      const uint8_t* mask, uint64_t* x, uint64_t* y, float* value) {
  size_t px, py;
  auto r = FindPeak(image, w, h, px, py, allow_negative != 0, start_y, end_y, h_border,
                    v_border, reinterpret_cast<const bool*>(mask));
  *x = px;
  *y = py;
  *value = r.value_or(0.0f);
  return r.has_value();
}

void orc_subtract_image(float* image, const float* psf, uint64_t w, uint64_t h, uint64_t x,
                        uint64_t y, float factor) {
  Image p(w, h);
  std::memcpy(p.Data(), psf, w * h * sizeof(float));
  ThreadedSubtractImage(image, p, x, y, factor);
}

// simple_clean::SubtractImage (non-partial; drops the odd last column, quirk A7)
void orc_subtract_image_whole(float* image, const float* psf, uint64_t w, uint64_t h, uint64_t x,
                              uint64_t y, float factor) {
  SubtractImageWhole(image, psf, w, h, x, y, factor);
}

void orc_partial_subtract_image(float* image, uint64_t iw, uint64_t ih, const float* psf,
                                uint64_t pw, uint64_t ph, uint64_t x, uint64_t y, float factor,
                                uint64_t start_y, uint64_t end_y) {
  PartialSubtractImage(image, iw, ih, psf, pw, ph, x, y, factor, start_y, end_y);
}

// mode 0: GetLinearIntegrated, 1: GetSquareIntegrated
void orc_integrate(const orc_cube* cube, const float* images, int32_t mode, float* dest) {
  const size_t n = cube->width * cube->height;
  ImageSet set(ToMeta(*cube), cube->width, cube->height);
  for (size_t i = 0; i != set.Size(); ++i) std::memcpy(set.Data(i), images + i * n, n * 4);
  Image d(cube->width, cube->height), scratch(cube->width, cube->height);
  if (mode == 0)
    set.GetLinearIntegrated(d);
  else
    set.GetSquareIntegrated(d, scratch);
  std::memcpy(dest, d.Data(), n * 4);
}

void orc_integrated_psf(const orc_cube* cube, const float* psfs, float* dest) {
  const size_t n = cube->width * cube->height;
  ImageSet set(ToMeta(*cube), cube->width, cube->height);
  std::vector<Image> p;
  for (size_t i = 0; i != cube->n_channels; ++i) {
    p.emplace_back(cube->width, cube->height);
    std::memcpy(p.back().Data(), psfs + i * n, n * 4);
  }
  Image d(cube->width, cube->height);
  set.GetIntegratedPsf(d, p);
  std::memcpy(dest, d.Data(), n * 4);
}

void orc_spectral_fit(int32_t mode, uint64_t n_terms, uint64_t n, uint64_t n_pol,
                      const double* freqs, const float* weights, float* values) {
  DeconvolutionAlgorithm* a = new GenericClean(false);
  std::vector<double> f(freqs, freqs + n);
  std::vector<float> w(weights, weights + n);
  a->fitter = SpectralFitter(SpectralMode(mode), n_terms, f, w);
  a->n_polarizations = n_pol;
  a->PerformSpectralFit(values);
  delete a;
}

void orc_polynomial_projection(uint64_t n_terms, uint64_t n, const double* freqs,
                               const float* weights, double* out) {
  std::vector<double> f(freqs, freqs + n);
  std::vector<float> w(weights, weights + n);
  std::vector<double> p = PolynomialProjection(n_terms, f, w);
  std::memcpy(out, p.data(), n * n * sizeof(double));
}

void orc_convolve(float* image, const float* kernel, uint64_t w, uint64_t h) {
  Convolve(image, kernel, w, h);
}
void orc_prepare_convolution_kernel(float* dst, const float* src, uint64_t w, uint64_t h) {
  PrepareConvolutionKernel(dst, src, w, h);
}
void orc_prepare_small_convolution_kernel(float* dst, uint64_t w, uint64_t h,
                                          const float* kernel, uint64_t n) {
  PrepareSmallConvolutionKernel(dst, w, h, kernel, n);
}
void orc_fft1d(float* interleaved, uint64_t n, int32_t inverse) {
  std::vector<std::pair<float, float>> d(n);
  for (size_t i = 0; i != n; ++i) d[i] = {interleaved[2 * i], interleaved[2 * i + 1]};
  Fft1D(d, inverse != 0);
  for (size_t i = 0; i != n; ++i) {
    interleaved[2 * i] = d[i].first;
    interleaved[2 * i + 1] = d[i].second;
  }
}
void orc_trim(float* dst, uint64_t dw, uint64_t dh, const float* src, uint64_t sw, uint64_t sh) {
  Trim(dst, dw, dh, src, sw, sh);
}
void orc_untrim(float* dst, uint64_t dw, uint64_t dh, const float* src, uint64_t sw,
                uint64_t sh) {
  Untrim(dst, dw, dh, src, sw, sh);
}
void orc_iuwt_smooth(float* out, const float* in, uint64_t w, uint64_t h, int32_t scale, int32_t mt) {
  std::vector<float> scratch(w * h);
  IuwtSmooth(out, in, scratch.data(), w, h, scale, mt != 0);
}
void orc_iuwt_decompose(const float* input, uint64_t w, uint64_t h, int32_t n_scales,
                        int32_t include_largest, float* coefficients) {
  IuwtDecompose(input, w, h, n_scales, include_largest != 0, coefficients);
}
void orc_iuwt_recompose(const float* coefficients, uint64_t w, uint64_t h, int32_t n_scales,
                        int32_t include_largest, const uint8_t* masks, float* out) {
  IuwtRecompose(coefficients, w, h, n_scales, include_largest != 0, masks, out);
}
uint64_t orc_iuwt_floodfill(const float* coefficients, uint8_t* masks, uint64_t w, uint64_t h,
                            int32_t n_scales, const float* thresholds, int32_t min_scale,
                            int32_t end_scale, uint64_t x, uint64_t y, int32_t scale,
                            float clean_border, const uint8_t* prior_mask) {
  return IuwtFloodfill(coefficients, masks, w, h, n_scales, thresholds, min_scale, end_scale, x, y,
                       scale, clean_border, prior_mask);
}
uint64_t orc_iuwt_select_structures(const float* coefficients, uint8_t* masks, uint64_t w,
                                    uint64_t h, int32_t n_scales, const float* thresholds,
                                    int32_t min_scale, int32_t end_scale, float clean_border,
                                    const uint8_t* prior_mask) {
  return IuwtSelectStructures(coefficients, masks, w, h, n_scales, thresholds, min_scale,
                              end_scale, clean_border, prior_mask);
}
uint64_t orc_iuwt_floodfill_2d(const float* image, uint8_t* mask, uint64_t w, uint64_t h,
                               float threshold, int32_t use_abs, uint64_t x, uint64_t y) {
  return IuwtFloodfill2D(image, mask, w, h, threshold, use_abs != 0, x, y);
}
float orc_iuwt_mad(const float* image, uint64_t n) { return IuwtMad(image, n); }
float orc_iuwt_max_abs(const float* image, uint64_t w, uint64_t h, float clean_border,
                       int32_t allow_negative, const uint8_t* mask, uint64_t* x, uint64_t* y) {
  size_t px, py;
  const float v = IuwtMaxAbs(image, w, h, clean_border, allow_negative != 0, mask, px, py);
  *x = px;
  *y = py;
  return v;
}
uint64_t orc_good_fft_size(uint64_t n) { return CalculateGoodFFTSize(n); }
uint64_t orc_convolution_size(double scale, uint64_t n, double padding) {
  return GetConvolutionSize(scale, n, padding);
}

// returns kernel size n; writes n*n floats if out != null (out must hold >= n*n)
uint64_t orc_make_shape_function(float scale, uint64_t max_n, int32_t shape, float* out) {
  size_t n;
  Image k = MakeShapeFunction(scale, n, max_n, MultiscaleShape(shape));
  if (out) std::memcpy(out, k.Data(), n * n * sizeof(float));
  return n;
}
void orc_ms_transform(float* images, uint64_t n_images, uint64_t w, uint64_t h, float scale,
                      int32_t shape) {
  MultiScaleTransforms ms(w, h, MultiscaleShape(shape));
  std::vector<Image> imgs;
  for (size_t i = 0; i != n_images; ++i) {
    imgs.emplace_back(w, h);
    std::memcpy(imgs.back().Data(), images + i * w * h, w * h * 4);
  }
  std::vector<Image*> l;
  for (Image& i : imgs) l.push_back(&i);
  Image scratch(w, h);
  ms.Transform(l, scratch, scale);
  for (size_t i = 0; i != n_images; ++i) std::memcpy(images + i * w * h, imgs[i].Data(), w * h * 4);
}
void orc_add_shape_component(float* image, uint64_t w, uint64_t h, float scale, uint64_t x,
                             uint64_t y, float gain, int32_t shape) {
  AddShapeComponent(image, w, h, scale, x, y, gain, MultiscaleShape(shape));
}

}  // extern "C"
