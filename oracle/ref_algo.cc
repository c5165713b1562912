// ORACLE — TEST INFRASTRUCTURE ONLY.
// extern "C" driver around the reference's OWN algorithm sources, compiled where they lie under
// /root/reference/cpp (oracle/Makefile target `ref`): image_set.cc, work_table.cc,
// algorithms/{deconvolution_algorithm, generic_clean, subminor_loop, threaded_deconvolution_tools,
// multiscale_algorithm, multiscale/multiscale_transforms, iuwt_deconvolution_algorithm,
// iuwt/*, parallel_deconvolution}.cc, math/{peak_finder, dijkstra_splitter}.cc, ...
// The un-vendored third-party leaves (aocommon Image / loops / logger, schaapcommon Convolve /
// SpectralFitter) are OUR stand-ins under ref_shim/ (see each header). The call sequence below
// is the one Radler::InitializeDeconvolutionAlgorithm (radler.cc:333-395) and Radler::Perform
// (radler.cc:241-273) make: build the algorithm, apply the settings, hand it to
// ParallelDeconvolution, ExecuteMajorIteration on ImageSets.
#include <cstdint>
#include <cstring>
#include <memory>
#include <set>
#include <string>
#include <vector>

#include <schaapcommon/math/convolution.h>

#include "algorithms/generic_clean.h"
#include "algorithms/iuwt_deconvolution.h"
#include "algorithms/multiscale_algorithm.h"
// the frame-sized per-scale masks of ParallelDeconvolution are private; a test reads them
#define private public
#include "algorithms/parallel_deconvolution.h"
#undef private
#include "image_set.h"
#include "radler.h"
#include "settings.h"
#include "work_table.h"
#ifdef RDL_WITH_ADAPTER
// the adapter library (oracle/Makefile target `adapter`): algorithm codes 10 / 11 / 12 put
// radler::algorithms::CudaDeconvolution behind the reference's ParallelDeconvolution
#include "cuda_deconvolution.h"
#endif

using namespace radler;

extern "C" {

struct refa_config {
  int32_t algorithm;  // 0 GenericClean, 1 MultiScaleAlgorithm, 2 IuwtDeconvolution
  int32_t use_sub_minor_optimization;
  float threshold, minor_loop_gain, major_loop_gain, clean_border_ratio, divergence_limit;
  uint64_t max_iterations;
  int32_t allow_negative, stop_on_negative;
  int32_t ms_fast_sub_minor_loop;
  double ms_sub_minor_loop_gain, ms_scale_bias;
  uint64_t ms_max_scales;
  double ms_convolution_padding;
  const double* ms_scale_list;
  uint64_t ms_n_scale_list;
  int32_t ms_shape;
  double beam_size, pixel_scale;
  int32_t track_components;
  int32_t fit_mode;  // 0 none, 1 polynomial
  uint64_t fit_terms;
  uint64_t grid_width, grid_height;
  uint64_t thread_count;
  int32_t squared_joins;
};

struct refa_result {
  int32_t has_start_peak;
  float start_peak, end_peak;
  int32_t another_iteration_required;
  uint64_t iteration_number;  // of the first algorithm (sum over facets in grid mode)
};

}  // extern "C"

namespace {

// Accessor that is never used: the driver fills ImageSets directly.
class UnusedAccessor final : public aocommon::ImageAccessor {
 public:
  void Load(float*) const override { throw std::runtime_error("unused accessor"); }
  void Store(const float*) override { throw std::runtime_error("unused accessor"); }
};

struct Handle {
  Settings settings;
  std::unique_ptr<WorkTable> table;
  std::unique_ptr<algorithms::ParallelDeconvolution> parallel;
  size_t width = 0, height = 0, n_channels = 0, n_pol = 0;
  std::unique_ptr<bool[]> mask;
  std::string error;
};

thread_local std::string g_error;

template <typename F>
int Guard(Handle* h, F&& f) {
  try {
    f();
    return 0;
  } catch (const std::exception& e) {
    (h ? h->error : g_error) = e.what();
    return 1;
  }
}

}  // namespace

extern "C" {

const char* refa_last_error(const void* handle) {
  return handle ? static_cast<const Handle*>(handle)->error.c_str() : g_error.c_str();
}

void* refa_create(const refa_config* c, uint64_t width, uint64_t height, uint64_t n_channels,
                  uint64_t n_pol, const int32_t* polarizations, const double* channel_weights,
                  const double* channel_frequencies) {
  auto h = std::make_unique<Handle>();
  const int rc = Guard(nullptr, [&] {
    Settings& s = h->settings;
    s.trimmed_image_width = width;
    s.trimmed_image_height = height;
    s.channels_out = n_channels;
    s.pixel_scale.x = s.pixel_scale.y = c->pixel_scale;
    s.thread_count = c->thread_count ? c->thread_count : 1;
    s.absolute_threshold = c->threshold;
    s.minor_loop_gain = c->minor_loop_gain;
    s.major_loop_gain = c->major_loop_gain;
    s.border_ratio = c->clean_border_ratio;
    s.divergence_limit = c->divergence_limit;
    s.minor_iteration_count = c->max_iterations;
    s.allow_negative_components = c->allow_negative != 0;
    s.stop_on_negative_components = c->stop_on_negative != 0;
    s.squared_joins = c->squared_joins != 0;
    s.save_source_list = c->track_components != 0;
    s.generic.use_sub_minor_optimization = c->use_sub_minor_optimization != 0;
    s.multiscale.fast_sub_minor_loop = c->ms_fast_sub_minor_loop != 0;
    s.multiscale.sub_minor_loop_gain = c->ms_sub_minor_loop_gain;
    s.multiscale.scale_bias = c->ms_scale_bias;
    s.multiscale.max_scales = c->ms_max_scales;
    s.multiscale.convolution_padding = c->ms_convolution_padding;
    s.multiscale.scale_list.assign(c->ms_scale_list, c->ms_scale_list + c->ms_n_scale_list);
    s.multiscale.shape =
        c->ms_shape == 1 ? MultiscaleShape::kGaussianShape : MultiscaleShape::kTaperedQuadraticShape;
    s.parallel.grid_width = c->grid_width ? c->grid_width : 1;
    s.parallel.grid_height = c->grid_height ? c->grid_height : 1;
    s.parallel.max_threads = s.thread_count;
    s.spectral_fitting.mode = c->fit_mode == 1
                                  ? schaapcommon::fitters::SpectralFittingMode::kPolynomial
                                  : schaapcommon::fitters::SpectralFittingMode::kNoFitting;
    s.spectral_fitting.terms = c->fit_terms;
    const int32_t algorithm_code = c->algorithm % 10;
    const bool use_adapter = c->algorithm >= 10;
    s.algorithm_type = algorithm_code == 0   ? AlgorithmType::kGenericClean
                       : algorithm_code == 1 ? AlgorithmType::kMultiscale
                                             : AlgorithmType::kIuwt;
    h->width = width;
    h->height = height;
    h->n_channels = n_channels;
    h->n_pol = n_pol;

    // one original channel per deconvolution channel; entries ordered [channel][pol]
    h->table = std::make_unique<WorkTable>(std::vector<PsfOffset>{}, n_channels, n_channels);
    for (size_t ch = 0; ch != n_channels; ++ch) {
      for (size_t p = 0; p != n_pol; ++p) {
        auto e = std::make_unique<WorkTableEntry>();
        const double f = channel_frequencies ? channel_frequencies[ch] : 1.0e8 + 1.0e6 * ch;
        e->band_start_frequency = f;
        e->band_end_frequency = f;
        e->polarization = polarizations ? aocommon::PolarizationEnum(polarizations[p])
                                        : aocommon::PolarizationEnum::StokesI;
        e->original_channel_index = ch;
        e->image_weight = channel_weights ? channel_weights[ch] : 1.0;
        e->model_accessor = std::make_unique<UnusedAccessor>();
        e->residual_accessor = std::make_unique<UnusedAccessor>();
        if (p == 0) e->psf_accessors.emplace_back(std::make_unique<UnusedAccessor>());
        h->table->AddEntry(std::move(e));
      }
    }

    // Radler::InitializeDeconvolutionAlgorithm, radler.cc:348-388
    std::unique_ptr<algorithms::DeconvolutionAlgorithm> algorithm;
    if (use_adapter) {
#ifdef RDL_WITH_ADAPTER
      // what INTEGRATION.md adds to Radler::InitializeDeconvolutionAlgorithm (radler.cc:348-375)
      rdl_config rc;
      rdl_config_default(&rc);
      rc.algorithm = algorithm_code == 0 ? RDL_GENERIC_CLEAN
                                         : (algorithm_code == 1 ? RDL_MULTISCALE : RDL_IUWT);
      rc.use_sub_minor_optimization = s.generic.use_sub_minor_optimization;
      rc.ms_fast_sub_minor_loop = s.multiscale.fast_sub_minor_loop;
      rc.ms_sub_minor_loop_gain = s.multiscale.sub_minor_loop_gain;
      rc.ms_scale_bias = s.multiscale.scale_bias;
      rc.ms_max_scales = s.multiscale.max_scales;
      rc.ms_convolution_padding = s.multiscale.convolution_padding;
      rc.ms_scale_list = s.multiscale.scale_list.data();
      rc.ms_n_scale_list = s.multiscale.scale_list.size();
      rc.ms_shape = c->ms_shape;
      rc.beam_size = c->beam_size;
      rc.pixel_scale_x = rc.pixel_scale_y = s.pixel_scale.x;
      algorithm = std::make_unique<algorithms::CudaDeconvolution>(rc);
#else
      throw std::runtime_error("this library was built without the CUDA adapter");
#endif
    } else
    switch (s.algorithm_type) {
      case AlgorithmType::kGenericClean:
        algorithm =
            std::make_unique<algorithms::GenericClean>(s.generic.use_sub_minor_optimization);
        break;
      case AlgorithmType::kMultiscale:
        algorithm = std::make_unique<algorithms::MultiScaleAlgorithm>(
            s.multiscale, c->beam_size, s.pixel_scale.x, s.pixel_scale.y, s.save_source_list);
        break;
      default:
        algorithm = std::make_unique<algorithms::IuwtDeconvolution>();
        break;
    }
    algorithm->SetMaxIterations(s.minor_iteration_count);
    algorithm->SetThreshold(s.absolute_threshold);
    algorithm->SetMinorLoopGain(s.minor_loop_gain);
    algorithm->SetMajorLoopGain(s.major_loop_gain);
    algorithm->SetCleanBorderRatio(s.border_ratio);
    algorithm->SetDivergenceLimit(s.divergence_limit);
    algorithm->SetAllowNegativeComponents(s.allow_negative_components);
    algorithm->SetStopOnNegativeComponents(s.stop_on_negative_components);
    // Radler::CreateSpectralFitter, radler.cc:318-331
    std::vector<double> frequencies;
    std::vector<float> weights;
    if (s.spectral_fitting.mode != schaapcommon::fitters::SpectralFittingMode::kNoFitting)
      ImageSet::CalculateDeconvolutionFrequencies(*h->table, frequencies, weights);
    algorithm->SetSpectralFitter(
        std::make_unique<schaapcommon::fitters::SpectralFitter>(
            s.spectral_fitting.mode, s.spectral_fitting.terms, std::move(frequencies),
            std::move(weights)),
        n_pol);
    h->parallel = std::make_unique<algorithms::ParallelDeconvolution>(h->settings);
    h->parallel->SetAlgorithm(std::move(algorithm));
  });
  return rc == 0 ? h.release() : nullptr;
}

void refa_destroy(void* handle) { delete static_cast<Handle*>(handle); }

int refa_set_threshold(void* handle, double v) {
  Handle* h = static_cast<Handle*>(handle);
  return Guard(h, [&] { h->parallel->SetThreshold(v); });
}
int refa_set_minor_loop_gain(void* handle, double v) {
  Handle* h = static_cast<Handle*>(handle);
  return Guard(h, [&] { h->parallel->SetMinorLoopGain(v); });
}
int refa_set_auto_mask_mode(void* handle, int32_t track, int32_t use) {
  Handle* h = static_cast<Handle*>(handle);
  return Guard(h, [&] { h->parallel->SetAutoMaskMode(track != 0, use != 0); });
}
int refa_set_clean_mask(void* handle, const uint8_t* mask) {
  Handle* h = static_cast<Handle*>(handle);
  return Guard(h, [&] {
    if (!mask) {
      h->mask.reset();
      h->parallel->SetCleanMask(nullptr);
      return;
    }
    const size_t n = h->width * h->height;
    h->mask.reset(new bool[n]);
    for (size_t i = 0; i != n; ++i) h->mask[i] = mask[i] != 0;
    h->parallel->SetCleanMask(h->mask.get());
  });
}
int refa_set_rms_factor_image(void* handle, const float* image) {
  Handle* h = static_cast<Handle*>(handle);
  return Guard(h, [&] {
    aocommon::Image img;
    if (image) {
      img = aocommon::Image(h->width, h->height);
      std::memcpy(img.Data(), image, h->width * h->height * sizeof(float));
    }
    h->parallel->SetRmsFactorImage(std::move(img));
  });
}
int refa_set_iteration_number(void* handle, uint64_t n) {
  Handle* h = static_cast<Handle*>(handle);
  return Guard(h, [&] { h->parallel->FirstAlgorithm().SetIterationNumber(n); });
}
int refa_set_max_iterations(void* handle, uint64_t n) {
  Handle* h = static_cast<Handle*>(handle);
  return Guard(h, [&] { h->parallel->FirstAlgorithm().SetMaxIterations(n); });
}
int refa_set_major_iteration_threshold(void* handle, float v) {
  Handle* h = static_cast<Handle*>(handle);
  return Guard(h, [&] { h->parallel->FirstAlgorithm().SetMajorIterationThreshold(v); });
}

// residual / model: n_channels*n_pol planes of width*height floats, updated in place.
// psfs: n_dd * n_channels planes of psf_width*psf_height floats ([dd][channel]).
int refa_execute(void* handle, float* residual, float* model, const float* psfs, uint64_t n_dd,
                 uint64_t psf_width, uint64_t psf_height, const int64_t* psf_offset_x,
                 const int64_t* psf_offset_y, double major_loop_gain, refa_result* out) {
  Handle* h = static_cast<Handle*>(handle);
  return Guard(h, [&] {
    const size_t n = h->width * h->height, n_images = h->n_channels * h->n_pol;
    const std::set<aocommon::PolarizationEnum> linked;  // empty = all linked
    ImageSet data(*h->table, h->settings.squared_joins, linked, h->width, h->height);
    ImageSet mdl(*h->table, h->settings.squared_joins, linked, h->width, h->height);
    for (size_t i = 0; i != n_images; ++i) {
      std::memcpy(data.Data(i), residual + i * n, n * sizeof(float));
      std::memcpy(mdl.Data(i), model + i * n, n * sizeof(float));
    }
    const size_t pn = psf_width * psf_height;
    std::vector<std::vector<aocommon::Image>> psf_images(n_dd);
    std::vector<PsfOffset> offsets;
    for (size_t d = 0; d != n_dd; ++d) {
      for (size_t c = 0; c != h->n_channels; ++c) {
        aocommon::Image img(psf_width, psf_height);
        std::memcpy(img.Data(), psfs + (d * h->n_channels + c) * pn, pn * sizeof(float));
        psf_images[d].push_back(std::move(img));
      }
      if (psf_offset_x && psf_offset_y) offsets.emplace_back(psf_offset_x[d], psf_offset_y[d]);
    }
    const algorithms::ParallelDeconvolutionResult r =
        h->parallel->ExecuteMajorIteration(data, mdl, psf_images, offsets, major_loop_gain);
    for (size_t i = 0; i != n_images; ++i) {
      std::memcpy(residual + i * n, data.Data(i), n * sizeof(float));
      std::memcpy(model + i * n, mdl.Data(i), n * sizeof(float));
    }
    out->has_start_peak = r.start_peak.HasValue();
    out->start_peak = r.start_peak.ValueOr(0.0f);
    out->end_peak = r.end_peak.ValueOr(0.0f);
    out->another_iteration_required = r.another_iteration_required;
    out->iteration_number = h->parallel->FirstAlgorithm().IterationNumber();
  });
}

// Convolution statistics since the last reset (oracle/ref_shim/schaapcommon/math/convolution.h):
// up to `capacity` (width, height, count) triples, the seconds spent inside Convolve; returns the
// number of distinct sizes.
uint64_t refa_convolve_stats(uint64_t* triples, uint64_t capacity, double* seconds, int32_t reset) {
  auto& s = schaapcommon::math::ConvolveStats::Get();
  const std::lock_guard<std::mutex> lock(s.mutex);
  uint64_t n = 0;
  for (const auto& kv : s.calls) {
    if (triples && n < capacity) {
      triples[3 * n] = kv.first.first;
      triples[3 * n + 1] = kv.first.second;
      triples[3 * n + 2] = kv.second;
    }
    ++n;
  }
  if (seconds) *seconds = s.seconds;
  if (reset) {
    s.calls.clear();
    s.seconds = 0.0;
  }
  return n;
}

uint64_t refa_iteration_number(void* handle) {
  return static_cast<Handle*>(handle)->parallel->FirstAlgorithm().IterationNumber();
}

uint64_t refa_scale_count(void* handle) {
  Handle* h = static_cast<Handle*>(handle);
  if (h->settings.algorithm_type != AlgorithmType::kMultiscale) return 0;
  return static_cast<algorithms::MultiScaleAlgorithm&>(h->parallel->FirstAlgorithm()).ScaleCount();
}
float refa_scale_size(void* handle, uint64_t i) {
  Handle* h = static_cast<Handle*>(handle);
  return static_cast<algorithms::MultiScaleAlgorithm&>(h->parallel->FirstAlgorithm()).ScaleSize(i);
}
uint64_t refa_scale_mask_count(void* handle) {
  Handle* h = static_cast<Handle*>(handle);
  return static_cast<algorithms::MultiScaleAlgorithm&>(h->parallel->FirstAlgorithm())
      .GetScaleMaskCount();
}
int refa_scale_mask(void* handle, uint64_t i, uint8_t* out) {
  Handle* h = static_cast<Handle*>(handle);
  return Guard(h, [&] {
    auto& m = static_cast<algorithms::MultiScaleAlgorithm&>(h->parallel->FirstAlgorithm())
                  .GetScaleMask(i);
    for (size_t k = 0; k != m.size(); ++k) out[k] = m[k] ? 1 : 0;
  });
}

// ParallelDeconvolution::scale_masks_ (frame-sized, kept between major iterations of a facet run)
uint64_t refa_parallel_scale_mask_count(void* handle) {
  return static_cast<Handle*>(handle)->parallel->scale_masks_.size();
}
int refa_parallel_scale_mask(void* handle, uint64_t i, uint64_t width, uint64_t height,
                             uint8_t* out) {
  Handle* h = static_cast<Handle*>(handle);
  return Guard(h, [&] {
    h->parallel->scale_masks_.at(i).GetBox(reinterpret_cast<bool*>(out), 0, 0, width, height);
  });
}

// The component list Radler would hand out (ParallelDeconvolution::GetComponentList, :184-207):
// the single algorithm's own list, or the list merged over the facets; after MergeDuplicates.
}  // extern "C"
namespace {
ComponentList& ComponentsOf(Handle* h) {
  if (h->parallel->algorithms_.size() == 1)
    return static_cast<algorithms::MultiScaleAlgorithm&>(h->parallel->FirstAlgorithm())
        .GetComponentList();
  return *h->parallel->component_list_;
}
}  // namespace
extern "C" {
uint64_t refa_component_count(void* handle, uint64_t scale) {
  ComponentList& list = ComponentsOf(static_cast<Handle*>(handle));
  list.MergeDuplicates();
  return scale < list.NScales() ? list.ComponentCount(scale) : 0;
}
uint64_t refa_component_scales(void* handle) {
  return ComponentsOf(static_cast<Handle*>(handle)).NScales();
}
int refa_component_get(void* handle, uint64_t scale, uint64_t i, uint64_t* x, uint64_t* y,
                       float* values) {
  Handle* h = static_cast<Handle*>(handle);
  return Guard(h, [&] {
    size_t px = 0, py = 0;
    ComponentsOf(h).GetComponent(scale, i, px, py, values);
    *x = px;
    *y = py;
  });
}

}  // extern "C"

// ---- Radler::Perform's steps either side of ExecuteMajorIteration ----------------------------
namespace {
// residual accessor over a caller's plane
class MemoryAccessor final : public aocommon::ImageAccessor {
 public:
  MemoryAccessor(const float* data, size_t w, size_t h) : data_(data), w_(w), h_(h) {}
  size_t Width() const override { return w_; }
  size_t Height() const override { return h_; }
  void Load(float* data) const override { std::memcpy(data, data_, w_ * h_ * sizeof(float)); }
  void Store(const float*) override { throw std::runtime_error("read-only accessor"); }

 private:
  const float* data_;
  size_t w_, h_;
};
}  // namespace

extern "C" {
// ImageSet::LoadAndAverage(true) (image_set.cc:105-140) on a work table of n_original channels x
// n_pol polarisations averaged to n_deconvolution channels; inputs[channel * n_pol + pol],
// weights per entry in the same order; outputs[deconvolution channel * n_pol + pol].
int refa_load_and_average(const float* const* inputs, const double* weights, uint64_t n_original,
                          uint64_t n_pol, uint64_t n_deconvolution, uint64_t width,
                          uint64_t height, float* const* outputs) {
  return Guard(nullptr, [&] {
    static const aocommon::PolarizationEnum kPols[4] = {
        aocommon::PolarizationEnum::StokesI, aocommon::PolarizationEnum::StokesQ,
        aocommon::PolarizationEnum::StokesU, aocommon::PolarizationEnum::StokesV};
    WorkTable table(std::vector<PsfOffset>{}, n_original, n_deconvolution);
    for (size_t ch = 0; ch != n_original; ++ch) {
      for (size_t p = 0; p != n_pol; ++p) {
        auto e = std::make_unique<WorkTableEntry>();
        e->band_start_frequency = e->band_end_frequency = 1.0e8 + 1.0e6 * ch;
        e->polarization = kPols[p % 4];
        e->original_channel_index = ch;
        e->image_weight = weights[ch * n_pol + p];
        e->model_accessor = std::make_unique<UnusedAccessor>();
        e->residual_accessor =
            std::make_unique<MemoryAccessor>(inputs[ch * n_pol + p], width, height);
        if (p == 0) e->psf_accessors.emplace_back(std::make_unique<UnusedAccessor>());
        table.AddEntry(std::move(e));
      }
    }
    ImageSet set(table, false, {}, width, height);
    set.LoadAndAverage(true);
    for (size_t i = 0; i != set.Size(); ++i)
      std::memcpy(outputs[i], set[i].Data(), width * height * sizeof(float));
  });
}

// ImageSet::InterpolateAndStoreModel (image_set.cc:208-283): models[deconvolution channel *
// n_pol + pol] -> outputs[original channel * n_pol + pol] through a polynomial fitter over the
// deconvolution frequencies / weights (the schaapcommon stand-in of ref_shim).
int refa_interpolate_model(const float* const* models, uint64_t n_deconvolution, uint64_t n_pol,
                           const double* fit_frequencies, const float* fit_weights,
                           uint64_t n_terms, const double* original_frequencies,
                           uint64_t n_original, uint64_t width, uint64_t height,
                           float* const* outputs) {
  return Guard(nullptr, [&] {
    class StoreAccessor final : public aocommon::ImageAccessor {
     public:
      StoreAccessor(float* data, size_t w, size_t h) : data_(data), w_(w), h_(h) {}
      size_t Width() const override { return w_; }
      size_t Height() const override { return h_; }
      void Load(float*) const override { throw std::runtime_error("write-only accessor"); }
      void Store(const float* data) override { std::memcpy(data_, data, w_ * h_ * sizeof(float)); }

     private:
      float* data_;
      size_t w_, h_;
    };
    static const aocommon::PolarizationEnum kPols[4] = {
        aocommon::PolarizationEnum::StokesI, aocommon::PolarizationEnum::StokesQ,
        aocommon::PolarizationEnum::StokesU, aocommon::PolarizationEnum::StokesV};
    WorkTable table(std::vector<PsfOffset>{}, n_original, n_deconvolution);
    for (size_t ch = 0; ch != n_original; ++ch) {
      for (size_t p = 0; p != n_pol; ++p) {
        auto e = std::make_unique<WorkTableEntry>();
        e->band_start_frequency = e->band_end_frequency = original_frequencies[ch];
        e->polarization = kPols[p % 4];
        e->original_channel_index = ch;
        e->image_weight = 1.0;
        e->model_accessor = std::make_unique<StoreAccessor>(outputs[ch * n_pol + p], width, height);
        e->residual_accessor = std::make_unique<UnusedAccessor>();
        if (p == 0) e->psf_accessors.emplace_back(std::make_unique<UnusedAccessor>());
        table.AddEntry(std::move(e));
      }
    }
    ImageSet set(table, false, {}, width, height);
    for (size_t i = 0; i != set.Size(); ++i)
      std::memcpy(set.Data(i), models[i], width * height * sizeof(float));
    const schaapcommon::fitters::SpectralFitter fitter(
        schaapcommon::fitters::SpectralFittingMode::kPolynomial, n_terms,
        std::vector<double>(fit_frequencies, fit_frequencies + n_deconvolution),
        std::vector<float>(fit_weights, fit_weights + n_deconvolution));
    set.InterpolateAndStoreModel(fitter);
  });
}

// ImageSet::LoadAndAveragePsfs (image_set.cc:142-205): one PSF per original channel ->
// outputs[deconvolution channel].
int refa_load_and_average_psfs(const float* const* inputs, const double* weights,
                               uint64_t n_original, uint64_t n_deconvolution, uint64_t width,
                               uint64_t height, float* const* outputs) {
  return Guard(nullptr, [&] {
    WorkTable table(std::vector<PsfOffset>{}, n_original, n_deconvolution);
    for (size_t ch = 0; ch != n_original; ++ch) {
      auto e = std::make_unique<WorkTableEntry>();
      e->band_start_frequency = e->band_end_frequency = 1.0e8 + 1.0e6 * ch;
      e->polarization = aocommon::PolarizationEnum::StokesI;
      e->original_channel_index = ch;
      e->image_weight = weights[ch];
      e->model_accessor = std::make_unique<UnusedAccessor>();
      e->residual_accessor = std::make_unique<UnusedAccessor>();
      e->psf_accessors.emplace_back(std::make_unique<MemoryAccessor>(inputs[ch], width, height));
      table.AddEntry(std::move(e));
    }
    const ImageSet set(table, false, {}, width, height);
    const std::vector<std::vector<aocommon::Image>> psfs = set.LoadAndAveragePsfs();
    for (size_t i = 0; i != psfs[0].size(); ++i)
      std::memcpy(outputs[i], psfs[0][i].Data(), width * height * sizeof(float));
  });
}

// Radler::Perform itself (radler.cc:130-316), major iterations of GenericClean with or without
// the sub-minor loop: n_original channels (one polarisation) averaged to n_deconvolution channels,
// the threshold from the MAD noise estimate (auto_threshold_sigma), a polynomial spectral fit,
// residuals assigned and the model interpolated back to the original channels. residuals /
// models are read and overwritten through load-and-store accessors, like wsclean's. With
// auto_mask_sigma > 0 and several major iterations the two-phase auto-masking runs (:171-185,
// :221-227, :252-268).
int refa_radler_perform(uint64_t width, uint64_t height, uint64_t n_original,
                        uint64_t n_deconvolution, const double* frequencies,
                        const double* weights, float* const* residuals, float* const* models,
                        const float* const* psfs, double auto_threshold_sigma,
                        double minor_loop_gain, double major_loop_gain,
                        uint64_t minor_iteration_count, uint64_t fit_terms,
                        double auto_mask_sigma, uint64_t n_major_iterations,
                        int32_t use_sub_minor_optimization, int32_t algorithm,
                        double beam_size, double pixel_scale,
                        int32_t* another_iteration_required, uint64_t* iteration_number) {
  return Guard(nullptr, [&] {
    class BufferAccessor final : public aocommon::ImageAccessor {
     public:
      BufferAccessor(float* data, size_t w, size_t h) : data_(data), w_(w), h_(h) {}
      size_t Width() const override { return w_; }
      size_t Height() const override { return h_; }
      void Load(float* data) const override { std::memcpy(data, data_, w_ * h_ * sizeof(float)); }
      void Store(const float* data) override { std::memcpy(data_, data, w_ * h_ * sizeof(float)); }

     private:
      float* data_;
      size_t w_, h_;
    };
    Settings s;
    s.trimmed_image_width = width;
    s.trimmed_image_height = height;
    s.channels_out = n_original;
    s.algorithm_type = algorithm == 1   ? AlgorithmType::kMultiscale
                       : algorithm == 2 ? AlgorithmType::kIuwt
                                        : AlgorithmType::kGenericClean;
    s.pixel_scale.x = s.pixel_scale.y = pixel_scale;
    s.generic.use_sub_minor_optimization = use_sub_minor_optimization != 0;
    s.auto_threshold_sigma = auto_threshold_sigma;
    if (auto_mask_sigma > 0.0) s.auto_mask_sigma = auto_mask_sigma;
    s.minor_loop_gain = minor_loop_gain;
    s.major_loop_gain = major_loop_gain;
    s.minor_iteration_count = minor_iteration_count;
    s.border_ratio = 0.0;
    if (fit_terms) {
      s.spectral_fitting.mode = schaapcommon::fitters::SpectralFittingMode::kPolynomial;
      s.spectral_fitting.terms = fit_terms;
    }
    auto table = std::make_unique<WorkTable>(std::vector<PsfOffset>{}, n_original, n_deconvolution);
    for (size_t ch = 0; ch != n_original; ++ch) {
      auto e = std::make_unique<WorkTableEntry>();
      e->band_start_frequency = e->band_end_frequency = frequencies[ch];
      e->polarization = aocommon::PolarizationEnum::StokesI;
      e->original_channel_index = ch;
      e->image_weight = weights[ch];
      e->model_accessor = std::make_unique<BufferAccessor>(models[ch], width, height);
      e->residual_accessor = std::make_unique<BufferAccessor>(residuals[ch], width, height);
      e->psf_accessors.emplace_back(
          std::make_unique<BufferAccessor>(const_cast<float*>(psfs[ch]), width, height));
      table->AddEntry(std::move(e));
    }
    Radler radler(s, std::move(table), beam_size);
    // successive major iterations on the stored residuals (no re-imaging in between); the
    // flag and the iteration count after each one
    for (uint64_t major = 0; major != n_major_iterations; ++major) {
      bool another = false;
      radler.Perform(another, major + 1);
      another_iteration_required[major] = another ? 1 : 0;
      iteration_number[major] = radler.IterationNumber();
    }
  });
}

// integrated.MedianAndStdDevFromMAD() (radler.cc:165-166): the aocommon stand-in of ref_shim
void refa_median_and_stddev_from_mad(const float* image, uint64_t n, float* median, float* stddev) {
  aocommon::Image img(n, 1);
  std::memcpy(img.Data(), image, n * sizeof(float));
  const std::pair<float, float> r = img.MedianAndStdDevFromMAD();
  *median = r.first;
  *stddev = r.second;
}
}  // extern "C"
