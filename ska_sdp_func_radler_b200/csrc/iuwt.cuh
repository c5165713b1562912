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
// ApplyMask (masks != nullptr: n_scales byte planes) + Recompose.
void IuwtRecompose(rdl_handle& h, const float* coefficients, size_t width, size_t height,
                   int n_scales, bool include_largest, const uint8_t* masks, float* out,
                   cudaStream_t stream);

}  // namespace rdl
#endif
