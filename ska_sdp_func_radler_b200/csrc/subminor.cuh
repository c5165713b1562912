// Sub-minor (Clark-like) loop on the device (algorithms/subminor_loop.{h,cc}).
#ifndef RDL_SUBMINOR_CUH_
#define RDL_SUBMINOR_CUH_

#include "handle.h"

namespace rdl {

// The knobs SubMinorLoop exposes (subminor_loop.h) as GenericClean / MultiScale set them.
struct SubMinorSetup {
  size_t width = 0, height = 0;
  float threshold = 0.0f, gain = 0.0f;
  size_t h_border = 0, v_border = 0;
  uint64_t current_iteration = 0, max_iterations = 0;
  bool allow_negative = true, stop_on_negative = false;
  const uint8_t* mask = nullptr;     // device, W*H
  const float* rms_factor = nullptr; // device, W*H
  float divergence_limit = 0.0f;
  uint32_t trace_scale = 0;
};

struct SubMinorOutcome {
  bool diverging = false;
  bool has_value = false;  // false when no pixel was selected (Run returns an unset optional)
  float max_value = 0.0f;
  uint64_t iteration = 0;  // CurrentIteration() after the run
  float flux_cleaned = 0.0f;
  size_t n_selected = 0;
};

// SubMinorLoop::Run (subminor_loop.cc:38-117): selection (findPeakPositions), gather
// (MakeSets) and the component loop as one persistent cooperative kernel. `residual` are the
// (convolved) residual images, `psf_per_image[i]` the (twice convolved) PSF of image i.
SubMinorOutcome RunSubMinorLoop(rdl_handle& h, const IntegSpec& linear_spec,
                                const ConstPtrTable& residual, const ConstPtrTable& psf_per_image,
                                int n_images, int n_pol, const SubMinorSetup& setup,
                                cudaStream_t stream);

// GetFullIndividualModel (subminor_loop.cc:186-193): full = 0; full[pos] = model_i.
void SubMinorScatterModel(rdl_handle& h, int image_index, float* full, size_t n_pixels,
                          cudaStream_t stream);
// model_full[pos] += model_i  (what `model += GetFullIndividualModel(i)` amounts to)
void SubMinorAddModelTo(rdl_handle& h, int image_index, float* model_full, cudaStream_t stream);
// UpdateAutoMask (subminor_loop.cc:220-228)
void SubMinorUpdateAutoMask(rdl_handle& h, int n_images, uint8_t* mask, cudaStream_t stream);
// UpdateComponentList (subminor_loop.cc:230-246): appends to h.components
void SubMinorUpdateComponentList(rdl_handle& h, int n_images, size_t scale_index,
                                 cudaStream_t stream);
// appends the device trace of the last run to h.trace_host
void SubMinorCollectTrace(rdl_handle& h, cudaStream_t stream);

rdl_result RunGenericCleanSubMinor(rdl_handle& h, DeviceCube& cube, cudaStream_t stream);
}  // namespace rdl
#endif
