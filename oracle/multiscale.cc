// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h).
// MultiScaleTransforms (algorithms/multiscale/multiscale_transforms.{h,cc}),
// ThreadedDeconvolutionTools::FindMultiScalePeak (threaded_deconvolution_tools.cc)
// and MultiScaleAlgorithm (algorithms/multiscale_algorithm.cc:29-748) restated.
// The component-optimiser paths (:750-930) are out of scope (default kClean).
#include "oracle.h"

#include <algorithm>
#include <cmath>
#include <map>
#include <set>
#include <stdexcept>

namespace oracle {

// ---- shape functions, multiscale_transforms.h:91-195 -------------------------
static float HannWindowFunction(float x, size_t n) {
  return (x * 2 <= n + 1) ? float(0.5 * (1.0 + std::cos(2.0 * M_PI * x / double(n + 1)))) : 0.0f;
}
static float ShapeFunction(float x) {
  // `1.0 - x * x` with float x: the reference build (GCC, -ffp-contract=fast) evaluates the
  // product exactly inside a double fnmadd; verified against oracle/_ref bit for bit.
  return (x < 1.0) ? float(std::fma(-double(x), double(x), 1.0)) : 0.0f;
}

static Image MakeTaperedQuadratic(double scale, size_t& n) {
  n = size_t(std::ceil(scale * 0.5) * 2.0) + 1;
  Image out(n, n);
  if (scale == 0.0) {
    out[0] = 1.0f;
    return out;
  }
  float sum = 0.0f;
  float* p = out.Data();
  for (int y = 0; y != int(n); ++y) {
    const float dy = y - 0.5 * (n - 1);
    const float dydy = dy * dy;
    for (int x = 0; x != int(n); ++x) {
      const float dx = x - 0.5 * (n - 1);
      const float r = std::sqrt(dx * dx + dydy);
      *p = HannWindowFunction(r, n) * ShapeFunction(r / scale);
      sum += *p;
      ++p;
    }
  }
  const float norm = 1.0 / sum;
  for (float& v : out.d) v *= norm;
  return out;
}

static Image MakeGaussian(double scale, size_t& n, size_t max_n) {
  float sigma = float(scale) * float(3.0 / 16.0);
  n = int(std::ceil(sigma * 12.0 / 2.0)) * 2 + 1;
  if (n > max_n) {
    n = max_n;
    if ((n % 2) == 0 && n > 0) --n;
  }
  if (n < 1) n = 1;
  if (sigma == 0.0) {
    sigma = 1.0;
    n = 1;
  }
  Image out(n, n);
  const float mu = int(n / 2);
  const float two_sigma_squared = 2.0 * sigma * sigma;
  float sum = 0.0f;
  std::vector<float> gaus(n);
  for (int i = 0; i != int(n); ++i) {
    const float v = float(i) - mu;
    gaus[i] = std::exp(-v * v / two_sigma_squared);
  }
  float* p = out.Data();
  for (int y = 0; y != int(n); ++y)
    for (int x = 0; x != int(n); ++x) {
      *p = gaus[x] * gaus[y];
      sum += *p;
      ++p;
    }
  const float norm = 1.0 / sum;
  for (float& v : out.d) v *= norm;
  return out;
}

Image MakeShapeFunction(float scale, size_t& n, size_t max_n, MultiscaleShape shape) {
  if (shape == MultiscaleShape::kGaussian) return MakeGaussian(scale, n, max_n);
  return MakeTaperedQuadratic(scale, n);
}

float KernelPeakValue(double scale, size_t max_n, MultiscaleShape shape) {  // :56-60
  size_t n;
  Image k = MakeShapeFunction(scale, n, max_n, shape);
  return k[n / 2 + (n / 2) * n];
}

void AddShapeComponent(float* image, size_t width, size_t height, float scale, size_t x,
                       size_t y, float gain, MultiscaleShape shape) {  // :62-89
  size_t n;
  Image kernel = MakeShapeFunction(scale, n, std::min(width, height), shape);
  const int left = x > n / 2 ? int(x - n / 2) : 0;
  const int top = y > n / 2 ? int(y - n / 2) : 0;
  const size_t right = std::min(x + (n + 1) / 2, width);
  const size_t bottom = std::min(y + (n + 1) / 2, height);
  for (size_t yi = top; yi != bottom; ++yi) {
    float* row = &image[yi * width];
    const float* kp = &kernel.Data()[(yi + n / 2 - y) * n + left + n / 2 - x];
    for (size_t xi = left; xi != right; ++xi) {
      row[xi] = std::fmaf(*kp, gain, row[xi]);
      ++kp;
    }
  }
}

void MultiScaleTransforms::Transform(std::vector<Image*>& images, Image& scratch, float scale) {
  // multiscale_transforms.cc:11-23 — circular, unpadded
  size_t kernel_size;
  Image shape = MakeShapeFunction(scale, kernel_size, std::min(width_, height_), shape_);
  std::fill(scratch.d.begin(), scratch.d.begin() + width_ * height_, 0.0f);
  PrepareSmallConvolutionKernel(scratch.Data(), width_, height_, shape.Data(), kernel_size);
  for (Image* image : images) Convolve(image->Data(), scratch.Data(), width_, height_);
}

// ---- MultiScaleAlgorithm ---------------------------------------------------------
MultiScaleAlgorithm::MultiScaleAlgorithm(const MultiscaleSettings& settings, double beam_size,
                                         double pixel_scale_x, double pixel_scale_y)
    : settings_(settings),
      beam_size_in_pixels_(beam_size / std::max(pixel_scale_x, pixel_scale_y)) {
  if (beam_size_in_pixels_ <= 0.0) beam_size_in_pixels_ = 1;
}

static void InitializeScales(std::vector<ScaleInfo>& scales, double beam_size_in_pixels,
                             size_t min_width_height, MultiscaleShape shape, size_t max_scales,
                             const std::vector<double>& scale_list) {  // :90-131
  if (scale_list.empty()) {
    if (scales.empty()) {
      size_t scale_index = 0;
      double scale = beam_size_in_pixels * 2.0;
      do {
        ScaleInfo& e = scales.emplace_back();
        e.scale = scale_index == 0 ? 0.0f : float(scale);
        e.kernel_peak = KernelPeakValue(scale, min_width_height, shape);
        scale *= 2.0;
        ++scale_index;
      } while (scale < min_width_height * 0.5 && (max_scales == 0 || scale_index < max_scales));
    } else {
      while (!scales.empty() && scales.back().scale >= min_width_height * 0.5)
        scales.pop_back();
    }
  } else if (scales.empty()) {
    std::multiset<double> sorted(scale_list.begin(), scale_list.end());
    for (double s : sorted) {
      ScaleInfo& e = scales.emplace_back();
      e.scale = float(s);
      e.kernel_peak = KernelPeakValue(e.scale, min_width_height, shape);
    }
  }
}

static void ConvolvePsfs(std::vector<Image>& convolved, const Image& psf, Image& scratch,
                         bool is_integrated, std::vector<ScaleInfo>& scales,
                         double beam_size_in_pixels, double scale_bias, double minor_loop_gain,
                         MultiscaleShape shape) {  // :29-88
  MultiScaleTransforms ms(psf.w, psf.h, shape);
  convolved.assign(scales.size(), Image());
  const double first_auto_scale_size = beam_size_in_pixels * 2.0;
  for (size_t si = 0; si != scales.size(); ++si) {
    ScaleInfo& e = scales[si];
    convolved[si] = psf;
    if (e.scale != 0.0f) ms.Transform(convolved[si], scratch, e.scale);
    if (is_integrated) {
      e.psf_peak = convolved[si][psf.w / 2 + (psf.h / 2) * psf.w];
      double exp_term;
      if (e.scale == 0.0f || scales.size() < 2)
        exp_term = 0.0;
      else
        exp_term = std::log2(e.scale / first_auto_scale_size);
      e.bias_factor = float(std::pow(scale_bias, -exp_term));
      e.gain = float(minor_loop_gain / e.psf_peak);
      e.is_active = true;
    }
  }
}

static std::optional<size_t> SelectMaximumScale(const std::vector<ScaleInfo>& scales) {
  // :133-151 — std::map keeps the first inserted among equal keys
  std::map<float, size_t> peak_to_scale;
  for (size_t i = 0; i != scales.size(); ++i)
    if (scales[i].is_active) {
      const float max_val =
          std::fabs(scales[i].max_unnormalized_image_value * scales[i].bias_factor);
      peak_to_scale.insert({max_val, i});
    }
  if (peak_to_scale.empty()) return {};
  return peak_to_scale.rbegin()->second;
}

static float Rms(const Image& image, size_t n) {  // threaded_deconvolution_tools.h:40-44
  float result = 0.0f;
  for (size_t i = 0; i != n; ++i) result = std::fmaf(image[i], image[i], result);
  return std::sqrt(result / float(n));
}

void MultiScaleAlgorithm::FindPeakDirect(const Image& image, Image& scratch, size_t si) {
  // :700-748
  ScaleInfo& info = scale_infos[si];
  const size_t hb = size_t(std::round(image.w * clean_border_ratio));
  const size_t vb = size_t(std::round(image.h * clean_border_ratio));
  const float* actual = image.Data();
  if (!rms_factor_image.Empty()) {
    for (size_t i = 0; i != image.Size(); ++i) scratch[i] = image[i] * rms_factor_image[i];
    actual = scratch.Data();
  }
  const bool* m = use_per_scale_masks_ ? reinterpret_cast<const bool*>(scale_masks[si].data())
                                       : clean_mask;
  std::optional<float> max_value =
      FindPeak(actual, image.w, image.h, info.max_image_value_x, info.max_image_value_y,
               allow_negative_components, 0, image.h, hb, vb, m);
  if (max_value) {
    info.max_unnormalized_image_value = *max_value;
    if (rms_factor_image.Empty())
      info.max_normalized_image_value = *max_value;
    else
      info.max_normalized_image_value =
          *max_value / rms_factor_image[info.max_image_value_x + info.max_image_value_y * image.w];
  } else {
    info.max_unnormalized_image_value = 0.0f;
    info.max_normalized_image_value = 0.0f;
  }
}

void MultiScaleAlgorithm::FindActiveScaleConvolvedMaxima(const ImageSet& image_set,
                                                         Image& integrated_scratch,
                                                         Image& scratch, bool report_rms) {
  // :578-634 + threaded_deconvolution_tools.cc:30-107
  const size_t width = image_set.Width(), height = image_set.Height();
  image_set.GetLinearIntegrated(integrated_scratch);
  std::vector<size_t> transform_indices;
  for (size_t si = 0; si != scale_infos.size(); ++si) {
    ScaleInfo& e = scale_infos[si];
    if (!e.is_active) continue;
    if (e.scale == 0.0f) {
      FindPeakDirect(integrated_scratch, scratch, si);
      if (report_rms) e.rms = Rms(integrated_scratch, width * height);
    } else {
      transform_indices.push_back(si);
    }
  }
  struct PeakData {
    std::optional<float> normalized, unnormalized;
    float rms;
    size_t x, y;
  };
  std::vector<PeakData> results(transform_indices.size());
  // The reference runs scales in a DynamicFor; our FFT is itself threaded, so
  // the scales are done one after another (results are identical).
  for (size_t t = 0; t != transform_indices.size(); ++t) {
    const size_t si = transform_indices[t];
    const float scale = scale_infos[si].scale;
    Image copy = integrated_scratch;
    Image local_scratch(width, height);
    MultiScaleTransforms ms(width, height, settings_.shape);
    ms.Transform(copy, local_scratch, scale);
    const size_t border_scale = size_t(std::ceil(scale * 0.5));
    const size_t xb = std::max<size_t>(size_t(std::round(width * clean_border_ratio)), border_scale);
    const size_t yb = std::max<size_t>(size_t(std::round(height * clean_border_ratio)), border_scale);
    PeakData& r = results[t];
    r.rms = report_rms ? Rms(copy, width * height) : -1.0f;
    const bool* m = use_per_scale_masks_ ? reinterpret_cast<const bool*>(scale_masks[si].data())
                                         : clean_mask;
    if (rms_factor_image.Empty()) {
      r.unnormalized = FindPeak(copy.Data(), width, height, r.x, r.y, allow_negative_components,
                                0, height, xb, yb, m);
      r.normalized = r.unnormalized;
    } else {
      for (size_t i = 0; i != rms_factor_image.Size(); ++i)
        local_scratch[i] = copy[i] * rms_factor_image[i];
      r.unnormalized = FindPeak(local_scratch.Data(), width, height, r.x, r.y,
                                allow_negative_components, 0, height, xb, yb, m);
      if (r.unnormalized)
        r.normalized = *r.unnormalized / rms_factor_image[r.x + r.y * width];
    }
  }
  for (size_t t = 0; t != results.size(); ++t) {
    ScaleInfo& e = scale_infos[transform_indices[t]];
    e.max_normalized_image_value = results[t].normalized.value_or(0.0f);
    e.max_unnormalized_image_value = results[t].unnormalized.value_or(0.0f);
    e.max_image_value_x = results[t].x;
    e.max_image_value_y = results[t].y;
    if (report_rms) e.rms = results[t].rms;
  }
}

void MultiScaleAlgorithm::ActivateScales(size_t last) {  // :636-656
  for (size_t i = 0; i != scale_infos.size(); ++i) {
    const bool do_activate =
        i == last ||
        std::fabs(scale_infos[i].max_unnormalized_image_value) * scale_infos[i].bias_factor >
            std::fabs(scale_infos[last].max_unnormalized_image_value) *
                (1.0 - minor_loop_gain) * scale_infos[last].bias_factor;
    scale_infos[i].is_active = do_activate;
  }
}

void MultiScaleAlgorithm::AddComponentToModel(ImageSet& model, size_t image_index, size_t swp,
                                              float component_value) {  // :676-698
  const size_t x = scale_infos[swp].max_image_value_x, y = scale_infos[swp].max_image_value_y;
  float* data = model.Data(image_index);
  if (scale_infos[swp].scale == 0.0f)
    data[x + model.Width() * y] += component_value;
  else
    AddShapeComponent(data, model.Width(), model.Height(), scale_infos[swp].scale, x, y,
                      component_value, settings_.shape);
  scale_infos[swp].n_components_cleaned++;
  scale_infos[swp].total_flux_cleaned += component_value;
  if (track_per_scale_masks_) scale_masks[swp][x + model.Width() * y] = 1;
}

DeconvolutionResult MultiScaleAlgorithm::ExecuteMajorIteration(ImageSet& data, ImageSet& model,
                                                               const std::vector<Image>& psfs) {
  // multiscale_algorithm.cc:183-576
  const size_t width = data.Width(), height = data.Height();
  if (stop_on_negative_component) allow_negative_components = true;
  InitializeScales(scale_infos, beam_size_in_pixels_, std::min(width, height), settings_.shape,
                   settings_.max_scales, settings_.scale_list);
  if (track_per_scale_masks_) {
    for (const auto& m : scale_masks)
      if (m.size() != width * height)
        throw std::runtime_error("Invalid automask size in multiscale algorithm");
    while (scale_masks.size() < scale_infos.size()) scale_masks.emplace_back(width * height, 0);
  }
  if (!rms_factor_image.Empty() && (rms_factor_image.w != width || rms_factor_image.h != height))
    throw std::runtime_error("Error in RMS factor image dimensions!");

  bool has_hit_threshold_in_sub_loop = false;
  size_t threshold_countdown = std::max(size_t{8}, scale_infos.size() * 3 / 2);

  const size_t scratch_w =
      GetConvolutionSize(scale_infos.back().scale, width, settings_.convolution_padding);
  const size_t scratch_h =
      GetConvolutionSize(scale_infos.back().scale, height, settings_.convolution_padding);
  Image scratch(scratch_w, scratch_h), scratch_b(scratch_w, scratch_h);
  Image integrated_scratch(width, height);
  std::vector<std::vector<Image>> convolved_psfs(data.PsfCount());
  data.GetIntegratedPsf(integrated_scratch, psfs);
  ConvolvePsfs(convolved_psfs[0], integrated_scratch, scratch, true, scale_infos,
               beam_size_in_pixels_, settings_.scale_bias, minor_loop_gain, settings_.shape);
  if (data.PsfCount() > 1)
    for (size_t i = 0; i != data.PsfCount(); ++i)
      ConvolvePsfs(convolved_psfs[i], psfs[i], scratch, false, scale_infos,
                   beam_size_in_pixels_, settings_.scale_bias, minor_loop_gain, settings_.shape);

  MultiScaleTransforms ms_transforms(width, height, settings_.shape);
  FindActiveScaleConvolvedMaxima(data, integrated_scratch, scratch, true);
  DeconvolutionResult result;
  std::optional<size_t> opt_scale = SelectMaximumScale(scale_infos);
  if (!opt_scale) {
    result.another_iteration_required = false;
    return result;
  }
  size_t swp = *opt_scale;

  bool is_final_threshold = false;
  const float initial_peak_value =
      std::fabs(scale_infos[swp].max_unnormalized_image_value * scale_infos[swp].bias_factor);
  float mgain_threshold = initial_peak_value * (1.0 - major_loop_gain);
  mgain_threshold = std::max(mgain_threshold, major_iteration_threshold);
  float first_threshold = mgain_threshold;
  if (threshold > first_threshold) {
    first_threshold = threshold;
    is_final_threshold = true;
  }

  ImageSet individual_convolved(data.meta, width, height);
  bool diverging = false;

  while (iteration_number < max_iterations &&
         std::fabs(scale_infos[swp].max_unnormalized_image_value * scale_infos[swp].bias_factor) >
             first_threshold &&
         (!stop_on_negative_component || scale_infos[swp].max_unnormalized_image_value >= 0.0f) &&
         threshold_countdown > 0 && !diverging) {
    ++n_minor_iterations;
    Record(uint32_t(swp), scale_infos[swp].max_image_value_x, scale_infos[swp].max_image_value_y,
           scale_infos[swp].max_unnormalized_image_value);
    std::vector<Image> twice_convolved(data.PsfCount());
    std::vector<Image*> transform_list;
    for (size_t i = 0; i != data.PsfCount(); ++i) {
      twice_convolved[i] = convolved_psfs[i][swp];
      transform_list.push_back(&twice_convolved[i]);
    }
    for (size_t i = 0; i != data.Size(); ++i) {
      individual_convolved[i] = data[i];
      transform_list.push_back(&individual_convolved[i]);
    }
    if (scale_infos[swp].scale != 0.0f)
      ms_transforms.Transform(transform_list, scratch, scale_infos[swp].scale);

    const float sub_iteration_gain_threshold =
        std::fabs(scale_infos[swp].max_unnormalized_image_value * scale_infos[swp].bias_factor) *
        (1.0 - settings_.sub_minor_loop_gain);
    float first_sub_iteration_threshold = sub_iteration_gain_threshold;
    if (first_threshold > first_sub_iteration_threshold) {
      first_sub_iteration_threshold = first_threshold;
      has_hit_threshold_in_sub_loop = true;
      threshold_countdown--;
    }
    (void)has_hit_threshold_in_sub_loop;

    if (settings_.fast_sub_minor_loop) {
      const size_t sub_minor_start_iteration = iteration_number;
      const size_t conv_w =
          GetConvolutionSize(scale_infos[swp].scale, width, settings_.convolution_padding);
      const size_t conv_h =
          GetConvolutionSize(scale_infos[swp].scale, height, settings_.convolution_padding);
      SubMinorLoop sub(width, height, conv_w, conv_h);
      sub.current_iteration = iteration_number;
      sub.max_iterations = max_iterations;
      sub.threshold = first_sub_iteration_threshold / scale_infos[swp].bias_factor;
      sub.gain = scale_infos[swp].gain;
      sub.divergence_limit = divergence_limit;
      sub.allow_negative = allow_negative_components;
      sub.stop_on_negative = stop_on_negative_component;
      const size_t scale_border = size_t(std::ceil(scale_infos[swp].scale * 0.5));
      sub.h_border = std::max<size_t>(size_t(std::round(width * clean_border_ratio)), scale_border);
      sub.v_border = std::max<size_t>(size_t(std::round(height * clean_border_ratio)), scale_border);
      if (!rms_factor_image.Empty()) sub.rms_factor_image = rms_factor_image;
      if (use_per_scale_masks_)
        sub.mask = reinterpret_cast<const bool*>(scale_masks[swp].data());
      else if (clean_mask)
        sub.mask = clean_mask;
      sub.parent = this;
      sub.trace_scale = uint32_t(swp);

      std::optional<float> peak_value;
      std::tie(diverging, peak_value) = sub.Run(individual_convolved, twice_convolved);
      if (divergence_limit != 0.0f && peak_value)
        diverging =
            diverging || std::fabs(*peak_value) > initial_peak_value * divergence_limit;
      if (!peak_value) break;

      iteration_number = sub.current_iteration;
      scale_infos[swp].n_components_cleaned += iteration_number - sub_minor_start_iteration;
      scale_infos[swp].total_flux_cleaned += sub.flux_cleaned;

      for (size_t i = 0; i != data.Size(); ++i) {
        const Image& psf = convolved_psfs[data.PsfIndex(i)][swp];
        sub.CorrectResidualDirty(scratch.Data(), scratch_b.Data(), integrated_scratch.Data(), i,
                                 data.Data(i), psf.Data());
        // GetFullIndividualModel writes W*H floats into `scratch`, which the
        // reference then moves into Transform as a W x H image (:442-455).
        Image full_model(width, height);
        sub.GetFullIndividualModel(i, full_model.Data());
        if (i == 0 && track_per_scale_masks_)
          sub.UpdateAutoMask(reinterpret_cast<bool*>(scale_masks[swp].data()));
        if (scale_infos[swp].scale != 0.0f)
          ms_transforms.Transform(full_model, integrated_scratch, scale_infos[swp].scale);
        float* m = model.Data(i);
        for (size_t k = 0; k != width * height; ++k) m[k] += full_model[k];
      }
    } else {
      const ScaleInfo& max_info = scale_infos[swp];
      std::vector<float> component_values(data.Size());
      while (iteration_number < max_iterations &&
             std::fabs(max_info.max_unnormalized_image_value * max_info.bias_factor) >
                 first_sub_iteration_threshold &&
             (!stop_on_negative_component || max_info.max_unnormalized_image_value >= 0.0f) &&
             !diverging) {
        const size_t x = max_info.max_image_value_x, y = max_info.max_image_value_y;
        for (size_t i = 0; i != data.Size(); ++i)
          component_values[i] = individual_convolved[i][x + y * width];
        PerformSpectralFit(component_values.data());
        for (size_t i = 0; i != data.Size(); ++i) {
          component_values[i] = component_values[i] * max_info.gain;
          const Image& psf = convolved_psfs[data.PsfIndex(i)][swp];
          ThreadedSubtractImage(data.Data(i), psf, x, y, component_values[i]);
          ThreadedSubtractImage(individual_convolved.Data(i), twice_convolved[data.PsfIndex(i)],
                                x, y, component_values[i]);
          AddComponentToModel(model, i, swp, component_values[i]);
        }
        individual_convolved.GetLinearIntegrated(integrated_scratch);
        FindPeakDirect(integrated_scratch, scratch, swp);
        const float abs_peak_value = std::fabs(scale_infos[swp].max_unnormalized_image_value *
                                               scale_infos[swp].bias_factor);
        if (divergence_limit != 0.0f)
          diverging = abs_peak_value > initial_peak_value * divergence_limit;
        ++iteration_number;
        Record(uint32_t(swp), max_info.max_image_value_x, max_info.max_image_value_y,
               max_info.max_unnormalized_image_value);
      }
    }

    ActivateScales(swp);
    FindActiveScaleConvolvedMaxima(data, integrated_scratch, scratch, false);
    opt_scale = SelectMaximumScale(scale_infos);
    if (!opt_scale) {
      result.another_iteration_required = false;
      return result;
    }
    swp = *opt_scale;
  }

  const bool max_iter_reached = iteration_number >= max_iterations;
  const bool negative_reached =
      stop_on_negative_component && scale_infos[swp].max_unnormalized_image_value < 0.0f;
  result.is_diverging = diverging;
  result.another_iteration_required =
      !max_iter_reached && !is_final_threshold && !negative_reached && !diverging;
  result.final_peak_value =
      scale_infos[swp].max_unnormalized_image_value * scale_infos[swp].bias_factor;
  return result;
}

}  // namespace oracle
