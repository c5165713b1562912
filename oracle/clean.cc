// ORACLE — TEST INFRASTRUCTURE ONLY (see oracle.h).
// SubMinorLoop (algorithms/subminor_loop.cc) and GenericClean
// (algorithms/generic_clean.cc) restated.
#include "oracle.h"

#include <algorithm>
#include <cmath>

namespace oracle {

// ---- SubMinorModel::GetMaxComponent, subminor_loop.cc:13-36 ----------------
size_t SubMinorLoop::GetMaxComponent(Image& scratch, float& max_value) const {
  residual_.GetLinearIntegrated(scratch);
  const size_t n = positions_.size();
  if (!compact_rms_.Empty())
    for (size_t i = 0; i != n; ++i) scratch[i] *= compact_rms_[i];
  size_t max_component = 0;
  max_value = scratch[0];
  for (size_t i = 0; i != n; ++i) {
    const float value = allow_negative ? std::fabs(scratch[i]) : scratch[i];
    if (value > max_value) {
      max_component = i;
      max_value = value;
    }
  }
  max_value = scratch[max_component];
  return max_component;
}

// ---- SubMinorLoop::findPeakPositions, subminor_loop.cc:143-184 --------------
void SubMinorLoop::FindPeakPositions(ImageSet& convolved_residual) {
  Image integrated(width_, height_);
  convolved_residual.GetLinearIntegrated(integrated);
  if (!rms_factor_image.Empty())
    for (size_t i = 0; i != integrated.Size(); ++i) integrated[i] *= rms_factor_image[i];
  const size_t xi_start = h_border,
               xi_end = std::max<long>(long(xi_start), long(width_) - long(h_border)),
               yi_start = v_border,
               yi_end = std::max<long>(long(yi_start), long(height_) - long(v_border));
  positions_.clear();
  for (size_t y = yi_start; y != yi_end; ++y) {
    const float* row = integrated.Data() + y * width_;
    const bool* mrow = mask ? mask + y * width_ : nullptr;
    for (size_t x = xi_start; x != xi_end; ++x) {
      const float value = allow_negative ? std::fabs(row[x]) : row[x];
      if (value >= threshold && (!mrow || mrow[x])) positions_.emplace_back(x, y);
    }
  }
}

// ---- SubMinorLoop::Run, subminor_loop.cc:38-117 ------------------------------
std::pair<bool, std::optional<float>> SubMinorLoop::Run(
    ImageSet& convolved_residual, const std::vector<Image>& twice_convolved_psfs) {
  FindPeakPositions(convolved_residual);
  const size_t n_sel = positions_.size();
  // MakeSets (:119-132): gather the selected pixels of every image.
  residual_ = ImageSet(convolved_residual.meta, n_sel, 1);
  model_ = ImageSet(convolved_residual.meta, n_sel, 1);
  for (size_t i = 0; i != convolved_residual.Size(); ++i) {
    const float* src = convolved_residual.Data(i);
    float* dst = residual_.Data(i);
    for (size_t px = 0; px != n_sel; ++px)
      dst[px] = src[positions_[px].second * width_ + positions_[px].first];
  }
  compact_rms_ = Image();
  if (!rms_factor_image.Empty()) {
    compact_rms_ = Image(n_sel, 1);
    for (size_t px = 0; px != n_sel; ++px)
      compact_rms_[px] = rms_factor_image[positions_[px].second * width_ + positions_[px].first];
  }
  if (n_sel == 0) return {false, std::nullopt};

  Image scratch(n_sel, 1);
  float max_value;
  size_t max_component = GetMaxComponent(scratch, max_value);
  const float max_value_at_start = std::fabs(max_value);
  bool diverging = false;
  const size_t n_images = residual_.Size();
  std::vector<float> component_values(n_images);
  while (std::fabs(max_value) > threshold && current_iteration < max_iterations &&
         (!stop_on_negative || max_value >= 0.0f) && !diverging) {
    for (size_t i = 0; i != n_images; ++i)
      component_values[i] = residual_[i][max_component] * gain;
    flux_cleaned += max_value * gain;
    const size_t x = positions_[max_component].first, y = positions_[max_component].second;
    if (parent) {
      parent->iteration_number = current_iteration;
      parent->Record(trace_scale, x, y, max_value);
      parent->PerformSpectralFit(component_values.data());
    }
    for (size_t i = 0; i != n_images; ++i) model_[i][max_component] += component_values[i];

    for (size_t i = 0; i != n_images; ++i) {
      float* image = residual_.Data(i);
      const Image& psf = twice_convolved_psfs[residual_.PsfIndex(i)];
      const float psf_factor = component_values[i];
      for (size_t px = 0; px != n_sel; ++px) {
        const int psf_x = int(positions_[px].first) - int(x) + int(width_ / 2);
        const int psf_y = int(positions_[px].second) - int(y) + int(height_ / 2);
        if (psf_x >= 0 && psf_x < int(width_) && psf_y >= 0 && psf_y < int(height_))
          image[px] = std::fmaf(-psf[size_t(psf_x) + size_t(psf_y) * width_], psf_factor, image[px]);
      }
    }
    max_component = GetMaxComponent(scratch, max_value);
    if (divergence_limit != 0.0f)
      diverging = std::fabs(max_value) > max_value_at_start * divergence_limit;
    ++current_iteration;
  }
  return {diverging, max_value};
}

void SubMinorLoop::GetFullIndividualModel(size_t image_index, float* out) const {  // :186-193
  std::fill(out, out + width_ * height_, 0.0f);
  const float* data = model_[image_index].Data();
  for (size_t px = 0; px != positions_.size(); ++px)
    out[positions_[px].first + positions_[px].second * width_] = data[px];
}

void SubMinorLoop::CorrectResidualDirty(float* scratch_a, float* scratch_b, float* scratch_c,
                                        size_t image_index, float* residual,
                                        const float* single_convolved_psf) const {  // :195-218
  Untrim(scratch_a, padded_width_, padded_height_, single_convolved_psf, width_, height_);
  PrepareConvolutionKernel(scratch_b, scratch_a, padded_width_, padded_height_);
  GetFullIndividualModel(image_index, scratch_c);
  Untrim(scratch_a, padded_width_, padded_height_, scratch_c, width_, height_);
  Convolve(scratch_a, scratch_b, padded_width_, padded_height_);
  Trim(scratch_c, width_, height_, scratch_a, padded_width_, padded_height_);
  for (size_t i = 0; i != width_ * height_; ++i) residual[i] -= scratch_c[i];
}

void SubMinorLoop::UpdateAutoMask(bool* mask_out) const {  // :220-228
  for (size_t i = 0; i != model_.Size(); ++i)
    for (size_t px = 0; px != positions_.size(); ++px)
      if (model_[i][px] != 0.0f)
        mask_out[positions_[px].first + positions_[px].second * width_] = true;
}

// ---- GenericClean -------------------------------------------------------------
std::optional<float> GenericClean::FindPeak(const Image& image, float* scratch, size_t& x,
                                            size_t& y) {  // generic_clean.cc:255-277
  const float* actual = image.Data();
  if (!rms_factor_image.Empty()) {
    for (size_t i = 0; i != image.Size(); ++i) scratch[i] = image[i] * rms_factor_image[i];
    actual = scratch;
  }
  return FindPeakRatio(actual, image.w, image.h, x, y, allow_negative_components, 0, image.h,
                       clean_border_ratio, clean_mask);
}

DeconvolutionResult GenericClean::ExecuteMajorIteration(ImageSet& dirty, ImageSet& model,
                                                        const std::vector<Image>& psfs) {
  // generic_clean.cc:56-253
  const size_t width = dirty.Width(), height = dirty.Height();
  const size_t iteration_counter_at_start = iteration_number;
  if (stop_on_negative_component) allow_negative_components = true;
  size_t conv_w = size_t(std::ceil(convolution_padding_ * float(width)));
  size_t conv_h = size_t(std::ceil(convolution_padding_ * float(height)));
  if (conv_w % 2 != 0) ++conv_w;
  if (conv_h % 2 != 0) ++conv_h;

  Image integrated(width, height);
  Image scratch_a(conv_w, conv_h), scratch_b(conv_w, conv_h);
  dirty.GetLinearIntegrated(integrated);
  size_t cx = 0, cy = 0;
  std::optional<float> max_value = FindPeak(integrated, scratch_a.Data(), cx, cy);
  DeconvolutionResult result;
  result.starting_peak_value = max_value;
  result.final_peak_value = max_value.value_or(0.0f);
  if (!max_value) return result;
  if (iteration_number >= max_iterations) return result;

  const float initial_max_value = std::fabs(*max_value);
  float first_threshold = threshold;
  const float major_iter_threshold =
      std::max(major_iteration_threshold, initial_max_value * (1.0f - major_loop_gain));
  if (major_iter_threshold > first_threshold) first_threshold = major_iter_threshold;

  bool diverging = false;
  if (use_sub_minor_optimization_) {
    SubMinorLoop loop(width, height, conv_w, conv_h);
    loop.current_iteration = iteration_number;
    loop.max_iterations = max_iterations;
    loop.threshold = first_threshold;
    loop.gain = minor_loop_gain;
    loop.allow_negative = allow_negative_components;
    loop.stop_on_negative = stop_on_negative_component;
    loop.parent = this;
    loop.divergence_limit = divergence_limit;
    if (!rms_factor_image.Empty()) loop.rms_factor_image = rms_factor_image;
    loop.mask = clean_mask;
    loop.h_border = size_t(std::round(width * clean_border_ratio));
    loop.v_border = size_t(std::round(height * clean_border_ratio));
    std::tie(diverging, max_value) = loop.Run(dirty, psfs);
    iteration_number = loop.current_iteration;
    last_n_selected = loop.NSelected();
    for (size_t i = 0; i != dirty.Size(); ++i) {
      const Image& psf = psfs[dirty.PsfIndex(i)];
      loop.CorrectResidualDirty(scratch_a.Data(), scratch_b.Data(), integrated.Data(), i,
                                dirty.Data(i), psf.Data());
      loop.GetFullIndividualModel(i, scratch_a.Data());
      float* m = model.Data(i);
      for (size_t k = 0; k != width * height; ++k) m[k] += scratch_a[k];
    }
    if (!max_value) max_value = FindPeak(integrated, scratch_a.Data(), cx, cy);
  } else {
    size_t peak_index = cx + cy * width;
    std::vector<float> peak_values(dirty.Size());
    while (max_value && std::fabs(*max_value) > first_threshold &&
           iteration_number < max_iterations &&
           !(*max_value < 0.0f && stop_on_negative_component) && !diverging) {
      Record(0, cx, cy, *max_value);
      for (size_t i = 0; i != dirty.Size(); ++i) peak_values[i] = dirty[i][peak_index];
      PerformSpectralFit(peak_values.data());
      for (size_t i = 0; i != dirty.Size(); ++i) {
        peak_values[i] *= minor_loop_gain;
        model.Data(i)[peak_index] += peak_values[i];
        ThreadedSubtractImage(dirty.Data(i), psfs[dirty.PsfIndex(i)], cx, cy, peak_values[i]);
      }
      dirty.GetSquareIntegrated(integrated, scratch_a);
      max_value = FindPeak(integrated, scratch_a.Data(), cx, cy);
      peak_index = cx + cy * width;
      if (max_value && divergence_limit != 0.0f)
        diverging = std::fabs(*max_value) > initial_max_value * divergence_limit;
      ++iteration_number;
    }
  }
  if (diverging) {
    if (max_value) result.final_peak_value = *max_value;
    result.another_iteration_required = false;
    result.is_diverging = true;
  } else if (max_value) {
    const bool final_threshold_reached =
        std::fabs(*max_value) <= threshold || *max_value == 0.0f;
    const bool negative_reached = *max_value < 0.0f && stop_on_negative_component;
    const bool mgain_reached = std::fabs(*max_value) <= major_iter_threshold;
    const bool did_work = (iteration_number - iteration_counter_at_start) != 0;
    result.another_iteration_required =
        mgain_reached && did_work && !negative_reached && !final_threshold_reached;
    result.final_peak_value = *max_value;
  } else {
    result.another_iteration_required = false;
  }
  return result;
}

}  // namespace oracle
