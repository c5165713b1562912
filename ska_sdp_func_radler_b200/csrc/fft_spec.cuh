// Radix-specialised instances of the register-radix FFT kernels (fft_spec.cu) for the frame sizes
// of the BASELINE configs and their paddings. The generic kernels in fft.cu dispatch the radix of
// every pass at run time, so ptxas has to give each of them the registers of the largest radix
// (128 at 25 % occupancy); an instance that knows its radices takes what it needs.
#ifndef RDL_FFT_SPEC_CUH_
#define RDL_FFT_SPEC_CUH_

#include <cuda_runtime.h>

#include "fft_fast.cuh"

namespace rdl {
namespace fast {

// Each returns false when there is no instance for the plan's radices (the caller then launches
// the generic kernel), true after launching.
bool LaunchRowForwardSpec(const RowPlan& p, const RowIo& io, int grid, int threads, size_t smem,
                          cudaStream_t stream);
bool LaunchRowInverseSpec(const RowPlan& p, const RowIo& io, int grid, int threads, size_t smem,
                          cudaStream_t stream);
bool LaunchColumnStridedSpec(bool inverse, const ColPlan& p, const ColIo& io, int tile0, dim3 grid,
                             int threads, cudaStream_t stream);
bool LaunchColumnBlockSpec(bool convolve, const ColPlan& p, const ColIo& io, int tile0, dim3 grid,
                           int threads, size_t smem, cudaStream_t stream);

// The whole column transform of a convolution in one kernel (WholePlan): is there an instance
// for these radices / launch it over `tiles` tiles of p.tw columns.
bool HasWholeColumnSpec(int ra, int rb, int rc);
bool LaunchWholeColumnSpec(bool convolve, const WholePlan& p, const ColIo& io, int tiles, int threads,
                           size_t smem, cudaStream_t stream);

}  // namespace fast
}  // namespace rdl
#endif
