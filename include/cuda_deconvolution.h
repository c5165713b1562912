// radler::algorithms::CudaDeconvolution -- the adapter a Radler maintainer adds to the reference
// tree (e.g. as cpp/algorithms/cuda_deconvolution.h): a DeconvolutionAlgorithm whose
// ExecuteMajorIteration runs in libradler_b200.so. It compiles against the reference's UNMODIFIED
// headers (cpp/algorithms/deconvolution_algorithm.h:60-210, cpp/image_set.h) and is what
// tests/test_gpu_adapter.py hands to the reference's own ParallelDeconvolution::SetAlgorithm
// (parallel_deconvolution.cc:227-242) through oracle/_ref/libradler_ref_adapter.so.
//
// Everything the C ABI needs is read through public interfaces: settings from the base class,
// channel weights from ImageSet::CalculateDeconvolutionFrequencies (the function ImageSet fills
// its own weights_ with, image_set.cc:62), the polarisation normalisation by the rule of
// ImageSet::InitializePolFactor (image_set.h:296-324).
#ifndef RADLER_ALGORITHMS_CUDA_DECONVOLUTION_H_
#define RADLER_ALGORITHMS_CUDA_DECONVOLUTION_H_

#include <memory>
#include <set>
#include <stdexcept>
#include <string>
#include <vector>

#include <aocommon/image.h>
#include <aocommon/polarization.h>

#include "algorithms/deconvolution_algorithm.h"
#include "image_set.h"
#include "radler_b200.h"

namespace radler::algorithms {

class CudaDeconvolution final : public DeconvolutionAlgorithm {
 public:
  explicit CudaDeconvolution(const rdl_config& config) {
    Check(rdl_create(&config, &handle_), nullptr);
    rdl_set_psfs_constant(handle_, 1);  // Radler passes the same psf_images every major iteration
  }
  CudaDeconvolution(const CudaDeconvolution& other) : DeconvolutionAlgorithm(other) {
    Check(rdl_clone(other.handle_, &handle_), other.handle_);  // Clone(): deep copy incl. state
    rdl_set_psfs_constant(handle_, 1);
  }
  ~CudaDeconvolution() override { rdl_destroy(handle_); }

  std::unique_ptr<DeconvolutionAlgorithm> Clone() const final {
    return std::make_unique<CudaDeconvolution>(*this);
  }

  DeconvolutionResult ExecuteMajorIteration(ImageSet& data, ImageSet& model,
                                            const std::vector<aocommon::Image>& psfs) final {
    // base-class state -> handle (the values the CPU algorithms read from the base class)
    rdl_set_threshold(handle_, Threshold());
    rdl_set_major_iteration_threshold(handle_, MajorIterationThreshold());
    rdl_set_minor_loop_gain(handle_, MinorLoopGain());
    rdl_set_major_loop_gain(handle_, MajorLoopGain());
    rdl_set_clean_border_ratio(handle_, CleanBorderRatio());
    rdl_set_max_iterations(handle_, MaxIterations());
    rdl_set_iteration_number(handle_, IterationNumber());
    rdl_set_divergence_limit(handle_, DivergenceLimit());
    rdl_set_allow_negative_components(handle_, AllowNegativeComponents());
    rdl_set_stop_on_negative_components(handle_, StopOnNegativeComponents());
    // the reference holds a non-owning mask pointer whose content may change between calls
    // (SubImage::mask, the auto-mask): copied at every call
    Check(rdl_set_clean_mask(handle_, reinterpret_cast<const uint8_t*>(CleanMask()), data.Width(),
                             data.Height()),
          handle_);
    Check(rdl_set_rms_factor_image(handle_,
                                   RmsFactorImage().Empty() ? nullptr : RmsFactorImage().Data(),
                                   data.Width(), data.Height()),
          handle_);
    const size_t n_channels = data.PsfCount(), n_pol = data.Size() / data.PsfCount();
    const schaapcommon::fitters::SpectralFitter& fitter = Fitter();
    if (fitter.Mode() == schaapcommon::fitters::SpectralFittingMode::kNoFitting) {
      Check(rdl_set_spectral_fitter(handle_, RDL_FIT_NONE, 0, nullptr, nullptr, 0, n_pol), handle_);
    } else {
      std::vector<float> fit_weights(fitter.Weights().begin(), fitter.Weights().end());
      Check(rdl_set_spectral_fitter(handle_, int32_t(fitter.Mode()), fitter.NTerms(),
                                    fitter.Frequencies().data(), fit_weights.data(),
                                    fitter.Frequencies().size(), n_pol),
            handle_);
    }

    std::vector<double> frequencies;
    std::vector<float> weights;
    ImageSet::CalculateDeconvolutionFrequencies(data.Table(), frequencies, weights);
    std::vector<uint8_t> linked(n_pol, 1);
    float pol_norm = 1.0f;
    PolarizationLayout(data, linked, pol_norm);

    rdl_cube_desc cube{};
    cube.width = data.Width();
    cube.height = data.Height();
    cube.n_channels = n_channels;
    cube.n_pol = n_pol;
    cube.channel_weights = weights.data();
    cube.pol_linked = linked.data();
    cube.pol_norm = pol_norm;
    cube.squared_joins = data.SquareJoinedChannels();
    cube.memory_space = RDL_MEM_HOST;  // aocommon::Image buffers
    std::vector<float*> residual(data.Size()), model_planes(model.Size());
    std::vector<const float*> psf_planes(psfs.size());
    for (size_t i = 0; i != data.Size(); ++i) {
      residual[i] = data.Data(i);
      model_planes[i] = model.Data(i);
    }
    for (size_t i = 0; i != psfs.size(); ++i) psf_planes[i] = psfs[i].Data();

    rdl_result out{};
    Check(rdl_execute_major_iteration(handle_, &cube, residual.data(), model_planes.data(),
                                      psf_planes.data(), nullptr, &out),
          handle_);
    SetIterationNumber(out.iteration_number);
    DeconvolutionResult result;
    if (out.has_starting_peak) result.starting_peak_value = out.starting_peak_value;
    result.final_peak_value = out.final_peak_value;
    result.another_iteration_required = out.another_iteration_required;
    result.is_diverging = out.is_diverging;
    return result;
  }

  // MultiScaleAlgorithm's extra surface (multiscale_algorithm.h:41-56)
  void SetAutoMaskMode(bool track, bool use) { rdl_set_auto_mask_mode(handle_, track, use); }
  size_t ScaleCount() const { return rdl_scale_count(handle_); }
  rdl_handle* Handle() const { return handle_; }

 private:
  // ImageSet::InitializePolFactor (image_set.h:296-324) on the public table
  static void PolarizationLayout(const ImageSet& data, std::vector<uint8_t>& linked, float& norm) {
    const WorkTable::Group& first = data.Table().OriginalGroups().front();
    const std::set<aocommon::PolarizationEnum>& links = data.LinkedPolarizations();
    std::set<aocommon::PolarizationEnum> pols;
    bool all_stokes_without_i = true;
    size_t p = 0;
    for (const WorkTableEntry* entry : first) {
      const bool is_linked = links.empty() || links.count(entry->polarization) != 0;
      if (p < linked.size()) linked[p] = is_linked ? 1 : 0;
      ++p;
      if (is_linked) {
        if (!aocommon::Polarization::IsStokes(entry->polarization) ||
            entry->polarization == aocommon::Polarization::StokesI)
          all_stokes_without_i = false;
        pols.insert(entry->polarization);
      }
    }
    const bool is_dual = pols.size() == 2 && aocommon::Polarization::HasDualPolarization(pols);
    const bool is_full = pols.size() == 4 && (aocommon::Polarization::HasFullLinearPolarization(pols) ||
                                              aocommon::Polarization::HasFullCircularPolarization(pols));
    if (all_stokes_without_i)
      norm = 1.0 / pols.size();
    else if (is_dual || is_full)
      norm = 0.5;
    else
      norm = 1.0;
  }

  static void Check(rdl_status s, const rdl_handle* h) {
    if (s != RDL_OK) throw std::runtime_error(rdl_last_error(h));  // the reference's convention
  }
  rdl_handle* handle_ = nullptr;
};

}  // namespace radler::algorithms
#endif
