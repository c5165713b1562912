// MultiScaleAlgorithm on the device (algorithms/multiscale_algorithm.{h,cc}).
#ifndef RDL_MULTISCALE_CUH_
#define RDL_MULTISCALE_CUH_

#include <memory>

#include "handle.h"

namespace rdl {
std::shared_ptr<MultiScaleState> CloneMultiScaleState(const MultiScaleState& src,
                                                      cudaStream_t stream);
// drops what the state keeps of the PSFs (rdl_set_psfs_constant runs): call when PSF pixels change
void MultiScaleInvalidatePsfCache(MultiScaleState& s);
uint64_t MultiScaleScaleCount(const MultiScaleState& s);
rdl_scale_info MultiScaleScaleInfo(const MultiScaleState& s, uint64_t i);
void MultiScaleCopyScaleMask(const MultiScaleState& s, uint64_t i, uint8_t* out_host);
// The per-scale masks of an algorithm (MultiScaleAlgorithm::GetScaleMask / SetScaleMaskCount,
// multiscale_algorithm.h:41-56), created on first use; mask_pixels = pixels of one mask.
std::vector<DeviceBuffer<uint8_t>>& MultiScaleMasks(rdl_handle& h, size_t mask_pixels);
void MultiScaleTransform(rdl_handle& h, float* const* images, uint64_t n_images, uint64_t width,
                         uint64_t height, float scale, int32_t shape, cudaStream_t stream);
}  // namespace rdl
#endif
