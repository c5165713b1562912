// Local-RMS maps (cpp/math/rms_image.cc of the reference) on the device: see rms_image.cu.
#ifndef RDL_RMS_IMAGE_CUH_
#define RDL_RMS_IMAGE_CUH_

#include <cstddef>

#include "handle.h"

namespace rdl {

// SlidingMinimum (maximum = false) / SlidingMaximum (rms_image.cc:39-84); out of place.
void RmsSlidingMinimum(rdl_handle& h, const float* in, float* out, size_t width, size_t height,
                       size_t window_size, bool maximum, cudaStream_t stream);
// MakeRmsFactorImage (rms_image.cc:97-123), in place; returns the lowest RMS (the stddev).
double RmsMakeFactorImage(rdl_handle& h, float* image, size_t n, double strength, cudaStream_t stream);
// Make / MakeWithNegativityLimit (rms_image.cc:16-37, 86-95). Angles in radians, beam and pixel
// scales in the same unit.
void RmsMake(rdl_handle& h, const float* in, float* out, size_t width, size_t height, double window_size,
             double beam_major, double beam_minor, double beam_pa, double pixel_scale_l,
             double pixel_scale_m, bool negativity_limit, cudaStream_t stream);

}  // namespace rdl
#endif
