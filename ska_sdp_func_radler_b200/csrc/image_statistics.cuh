// The steps of Radler::Perform either side of ExecuteMajorIteration on the
// device: see image_statistics.cu.
#ifndef RDL_IMAGE_STATISTICS_CUH_
#define RDL_IMAGE_STATISTICS_CUH_

#include <cstddef>
#include <cstdint>

#include "handle.h"

namespace rdl {

struct MadStatistics {
  float median = 0.0f, stddev = 0.0f;
  uint64_t n_finite = 0;
};

// aocommon::Image::MedianAndStdDevFromMAD as Radler::Perform calls it on the integrated residual
// (radler.cc:162-169): median of the finite pixels, and 1.4826 x the median of |pixel - median|.
MadStatistics ImageMedianAndStdDevFromMad(rdl_handle& h, const float* image, size_t n,
                                          cudaStream_t stream);

// ImageSet::LoadAndAverage (image_set.cc:105-140) with the input planes already in device
// memory: output[output_index[i]] accumulates input[i] x weight[i] in the order given (inputs of
// weight zero are left out, as the reference does), then every output is scaled by
// 1 / (sum of its weights). `inputs` / `outputs` are HOST arrays of device pointers.
// psf_rule: ImageSet::LoadAndAveragePsfs (image_set.cc:142-205) instead -- the same grouping
// with the products and the final scale in double, zero-weight inputs included.
void ImageLoadAndAverage(rdl_handle& h, const float* const* inputs, const double* weights,
                         const uint32_t* output_index, size_t n_inputs, float* const* outputs,
                         size_t n_outputs, size_t n_pixels, bool psf_rule, cudaStream_t stream);

// The auto-mask Radler::Perform derives from the model images when the second auto-masking phase
// starts (radler.cc:254-266): a pixel is set where any model plane is finite and non-zero.
// `models` is a HOST array of device pointers.
void ImageMaskFromModel(rdl_handle& h, const float* const* models, size_t n_models, uint8_t* mask,
                        size_t n_pixels, cudaStream_t stream);

// ImageSet::InterpolateAndStoreModel (image_set.cc:208-283) for one polarisation when there are
// fewer deconvolution channels than original ones: per pixel, fit the handle's polynomial
// spectral fitter to the deconvolution-channel model planes and evaluate it at every output
// frequency. `models` / `outputs` are HOST arrays of device pointers.
void ImageInterpolateModel(rdl_handle& h, const float* const* models, size_t n_models,
                           const double* output_frequencies, size_t n_outputs,
                           float* const* outputs, size_t n_pixels, cudaStream_t stream);

}  // namespace rdl
#endif
