// IUWT decomposition kernels (iuwt_decomposition.{h,cc} of the reference, device-side).
#ifndef RDL_IUWT_CUH_
#define RDL_IUWT_CUH_

#include <cstddef>
#include <cstdint>

#include "handle.h"

namespace rdl {

// out = B3 a-trous smoothing of in at `scale` >= 1 (taps at (2^scale - 1) * {-2..2}).
// simple = false: IuwtDecomposition::convolveMT (decomposition flavour);
// simple = true: IuwtDecomposition::convolve (Recompose / Convolve flavour).
void IuwtSmooth(rdl_handle& h, const float* in, float* out, size_t width, size_t height, int scale,
                bool simple, cudaStream_t stream);
// IuwtDecomposition::DecomposeMt: coefficients = n_scales (+1 when include_largest) planes.
void IuwtDecompose(rdl_handle& h, const float* input, size_t width, size_t height, int n_scales,
                   bool include_largest, float* coefficients, cudaStream_t stream);
// IuwtDecomposition::Decompose(x, x, false) -- input aliased to the scratch image -- exactly as
// the reference's code behaves (see iuwt.cu). n_scales planes; x is clobbered.
void IuwtDecomposeAliased(rdl_handle& h, float* x, size_t width, size_t height, int n_scales,
                          float* coefficients, cudaStream_t stream);
// ApplyMask (masks != nullptr: n_scales byte planes) + Recompose.
void IuwtRecompose(rdl_handle& h, const float* coefficients, size_t width, size_t height,
                   int n_scales, bool include_largest, const uint8_t* masks, float* out,
                   cudaStream_t stream);


// ---- structure selection and plane measurements (iuwt_analysis.cu) --------------------------
// h.iuwt_scalars layout in 64-bit words: [0] voxel counter, [2] radix-select prefix, [3] rank,
// [4] arg-max key, [5..7] arg-max result, [8..8+kIuwtFillBatch/2) per-round change counters,
// [32..) 2048-bin histogram of 32-bit counts.
constexpr int kIuwtFillBatch = 8;
constexpr int kIuwtChangedWord = 8;
constexpr int kIuwtHistWord = 32;
constexpr int kIuwtScalarWords = kIuwtHistWord + 1024;

// image_analysis::SelectStructures (image_analysis.cc:208-239): marks every voxel of scales
// [min_scale, end_scale) that exceeds its scale's threshold (ExceedsThreshold, :9-15), is not
// yet marked, lies in the prior mask (nullptr = everywhere) and inside the clean border.
// masks: n_scales planes of width*height bytes, in/out. Returns the number of voxels added.
uint64_t IuwtSelectStructures(rdl_handle& h, const float* coefficients, uint8_t* masks,
                              size_t width, size_t height, int n_scales, const float* thresholds,
                              int min_scale, int end_scale, float clean_border,
                              const uint8_t* prior_mask, cudaStream_t stream);
// image_analysis::Floodfill / MaskedFloodfill (:86-206): the 6-connected structure around one
// seed voxel. Returns the number of voxels filled (the seed included).
uint64_t IuwtFloodfill(rdl_handle& h, const float* coefficients, uint8_t* masks, size_t width,
                       size_t height, int n_scales, const float* thresholds, int min_scale,
                       int end_scale, size_t seed_x, size_t seed_y, int seed_scale,
                       float clean_border, const uint8_t* prior_mask, cudaStream_t stream);
// image_analysis::FloodFill2D (:241-333): use_abs selects the fabs(v) > threshold flavour.
uint64_t IuwtFloodfill2D(rdl_handle& h, const float* image, uint8_t* mask, size_t width,
                         size_t height, float threshold, bool use_abs, size_t seed_x,
                         size_t seed_y, cudaStream_t stream);
// IuwtDeconvolutionAlgorithm::Mad (iuwt_deconvolution_algorithm.cc:104-110):
// element n/2 of sorted |v|, divided by 0.674559.
float IuwtMad(rdl_handle& h, const float* image, size_t n, cudaStream_t stream);
// GetMaxAbsWithoutMask / GetMaxAbsWithMask (:112-167): first row-major maximum of |v| (or v)
// inside the clean border; x = width, y = height and lowest() when nothing qualifies.
float IuwtMaxAbs(rdl_handle& h, const float* image, size_t width, size_t height,
                 float clean_border, bool allow_negative, const uint8_t* mask, size_t& x,
                 size_t& y, cudaStream_t stream);

}  // namespace rdl
#endif
