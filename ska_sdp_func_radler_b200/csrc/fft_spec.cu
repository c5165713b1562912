// Radix-specialised instances of the register-radix FFT kernels: see fft_spec.cuh.
// The bodies are the ones of fft_fast.cuh; only the radix dispatch is resolved at compile time.
#include "fft_spec.cuh"

#include <mutex>

#include "common.cuh"

namespace rdl {
namespace fast {
namespace {

// T threads: enough for the pair pass (ra * rb / 2 + 1 tasks per row) to finish in ONE round. With
// the plan's 128 threads a 4096-point row has 129 pair tasks: one thread ran a second round
// alone while the CTA waited at the barrier (ncu source page of RowInverseSpecKernel<16,16,16>:
// 17 % of the samples on that barrier, profiles/r02_ncu_rowinverse_8192_source.txt). Worth 2-3 %
// per convolution where 160 threads do it (129-145 pair tasks); 192 threads for the 163-169 pair
// tasks of 9072 / 9408 cost more occupancy than the second round (measured +9 %), so those stay
// at 128.
template <int RA, int RB, int RC, int T, int MINB>
__global__ void __launch_bounds__(T, MINB)
    RowForwardSpecKernel(const __grid_constant__ RowPlan p, const __grid_constant__ RowIo io) {
  extern __shared__ __align__(16) float2 smem[];
  const int row0 = blockIdx.x * p.rows_per_cta;
  const int n_rows = MinInt(p.rows_per_cta, io.sh - row0);
  RowForwardPass1<RA>(p, io, smem, row0, n_rows, threadIdx.x, blockDim.x);
  __syncthreads();
  if (RB > 1) RowForwardPass2<RB>(p, smem, n_rows, threadIdx.x, blockDim.x);
  __syncthreads();
  RowForwardPass3<RC>(p, io, smem, row0, n_rows, threadIdx.x, blockDim.x);
}

template <int RA, int RB, int RC, int T, int MINB>
__global__ void __launch_bounds__(T, MINB)
    RowInverseSpecKernel(const __grid_constant__ RowPlan p, const __grid_constant__ RowIo io) {
  extern __shared__ __align__(16) float2 smem[];
  const int row0 = blockIdx.x * p.rows_per_cta;
  const int n_rows = MinInt(p.rows_per_cta, io.sh - row0);
  RowInversePass1<RC>(p, io, smem, row0, n_rows, threadIdx.x, blockDim.x);
  __syncthreads();
  if (RB > 1) RowInversePass2<RB>(p, smem, n_rows, threadIdx.x, blockDim.x);
  __syncthreads();
  RowInversePass3<RA>(p, io, smem, row0, n_rows, threadIdx.x, blockDim.x);
}

template <int R2, bool INV, int MINB>
__global__ void __launch_bounds__(kTile * kStridedRows, MINB)
    ColumnStridedSpecKernel(const __grid_constant__ ColPlan p, const __grid_constant__ ColIo io,
                            int tile0) {
  const int c = threadIdx.x % kTile, r = blockIdx.y * kStridedRows + threadIdx.x / kTile;
  if (r >= p.H1) return;
  ColumnStridedBody<R2, INV>(p, io, (tile0 + blockIdx.x) * kTile, r, c);
}

template <int RA, int RB, bool CONVOLVE, int MINB>
__global__ void __launch_bounds__(256, MINB)
    ColumnBlockSpecKernel(const __grid_constant__ ColPlan p, const __grid_constant__ ColIo io,
                          int tile0) {
  extern __shared__ __align__(16) float2 smem[];
  const size_t row0 = size_t(blockIdx.y) * p.H1;
  const int c0 = (tile0 + blockIdx.x) * kTile;
  ColumnBlockPassA<RA>(p, io, smem, row0, c0, threadIdx.x, blockDim.x);
  __syncthreads();
  ColumnBlockPassB<RB, CONVOLVE>(p, io, smem, row0, c0, threadIdx.x, blockDim.x);
  if (!CONVOLVE) return;
  __syncthreads();
  ColumnBlockPassC<RA>(p, io, smem, row0, c0, threadIdx.x, blockDim.x);
}

// one CTA per tile of `tw` whole columns: fft_fast.cuh, "whole columns in shared memory"
template <int RA, int RB, int RC, bool CONVOLVE>
__global__ void __launch_bounds__(512, 1)
    WholeColumnSpecKernel(const __grid_constant__ WholePlan p, const __grid_constant__ ColIo io) {
  extern __shared__ __align__(16) float2 smem[];
  const int tile = blockIdx.x;
  WholeColumnPhase<0, RA, RB, RC, CONVOLVE>(p, io, smem, tile, threadIdx.x, blockDim.x);
  __syncthreads();
  WholeColumnPhase<1, RA, RB, RC, CONVOLVE>(p, io, smem, tile, threadIdx.x, blockDim.x);
  __syncthreads();
  WholeColumnPhase<2, RA, RB, RC, CONVOLVE>(p, io, smem, tile, threadIdx.x, blockDim.x);
  if (!CONVOLVE) return;
  __syncthreads();
  WholeColumnPhase<3, RA, RB, RC, CONVOLVE>(p, io, smem, tile, threadIdx.x, blockDim.x);
  __syncthreads();
  WholeColumnPhase<4, RA, RB, RC, CONVOLVE>(p, io, smem, tile, threadIdx.x, blockDim.x);
}

template <typename K>
void AllowSharedMemory(K kernel, size_t smem) {
  if (smem > size_t(48) * 1024)
    RDL_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, int(smem)));
}

// (ra, rb, rc, threads, CTAs per SM the register budget is cut for)
#define RDL_ROW_SPECS(X)                                                                            \
  X(16, 16, 16, 160, 5) /* 8192 */ X(16, 16, 8, 160, 5) /* 4096 */ X(18, 16, 16, 160, 4) /* 9216 */ \
  X(21, 16, 14, 128, 4) /* 9408 */ X(14, 14, 12, 128, 6) /* 4704 */ X(16, 16, 9, 160, 5) /* 4608 */ \
  X(18, 15, 9, 160, 4) /* 4860 */ X(18, 18, 7, 128, 5) /* 4536 */ X(18, 18, 14, 128, 5) /* 9072 */  \
  X(16, 16, 4, 128, 6) /* 2048 */
// (H2, CTAs per SM)
#define RDL_STRIDED_SPECS(X) X(16, 6) X(18, 6) X(21, 5) X(24, 4) X(28, 4) X(32, 4)
// (ra, rb, CTAs per SM)
#define RDL_BLOCK_SPECS(X) X(16, 16, 4) X(24, 16, 3) X(21, 16, 3) X(16, 14, 4) X(18, 15, 4) X(18, 14, 4) \
  X(21, 18, 3) X(16, 8, 4)

// (ra, rb, rc) of the column height
#define RDL_WHOLE_SPECS(X)                                                                        \
  X(16, 16, 8) /* 2048 */ X(16, 16, 9) /* 2304 */ X(16, 16, 16) /* 4096 */ X(18, 18, 14) /* 4536 */ \
  X(18, 16, 16) /* 4608 */ X(21, 16, 14) /* 4704 */ X(18, 18, 15) /* 4860 */ X(32, 32, 8) /* 8192 */ \
  X(24, 21, 18) /* 9072 */ X(24, 24, 16) /* 9216 */ X(28, 28, 12) /* 9408 */ X(27, 24, 15) /* 9720 */

}  // namespace

bool HasWholeColumnSpec(int ra, int rb, int rc) {
#define X(A, B, C) \
  if (ra == A && rb == B && rc == C) return true;
  RDL_WHOLE_SPECS(X)
#undef X
  return false;
}

bool LaunchWholeColumnSpec(bool convolve, const WholePlan& p, const ColIo& io, int tiles, int threads,
                           size_t smem, cudaStream_t stream) {
#define X(A, B, C)                                                                           \
  if (p.ra == A && p.rb == B && p.rc == C) {                                                 \
    if (convolve) {                                                                          \
      AllowSharedMemory(WholeColumnSpecKernel<A, B, C, true>, smem);                         \
      WholeColumnSpecKernel<A, B, C, true><<<tiles, threads, smem, stream>>>(p, io);         \
    } else {                                                                                 \
      AllowSharedMemory(WholeColumnSpecKernel<A, B, C, false>, smem);                        \
      WholeColumnSpecKernel<A, B, C, false><<<tiles, threads, smem, stream>>>(p, io);        \
    }                                                                                        \
    return true;                                                                             \
  }
  RDL_WHOLE_SPECS(X)
#undef X
  return false;
}

bool LaunchRowForwardSpec(const RowPlan& p, const RowIo& io, int grid, int threads, size_t smem,
                          cudaStream_t stream) {
#define X(A, B, C, T, M)                                                          \
  if (p.ra == A && p.rb == B && p.rc == C) {                                      \
    AllowSharedMemory(RowForwardSpecKernel<A, B, C, T, M>, smem);                 \
    RowForwardSpecKernel<A, B, C, T, M>                                           \
        <<<grid, p.rows_per_cta == 1 ? T : threads, smem, stream>>>(p, io);       \
    return true;                                                                  \
  }
  RDL_ROW_SPECS(X)
#undef X
  return false;
}

bool LaunchRowInverseSpec(const RowPlan& p, const RowIo& io, int grid, int threads, size_t smem,
                          cudaStream_t stream) {
#define X(A, B, C, T, M)                                                          \
  if (p.ra == A && p.rb == B && p.rc == C) {                                      \
    AllowSharedMemory(RowInverseSpecKernel<A, B, C, T, M>, smem);                 \
    RowInverseSpecKernel<A, B, C, T, M>                                           \
        <<<grid, p.rows_per_cta == 1 ? T : threads, smem, stream>>>(p, io);       \
    return true;                                                                  \
  }
  RDL_ROW_SPECS(X)
#undef X
  return false;
}

bool LaunchColumnStridedSpec(bool inverse, const ColPlan& p, const ColIo& io, int tile0, dim3 grid,
                             int threads, cudaStream_t stream) {
#define X(R, M)                                                                                 \
  if (p.H2 == R) {                                                                              \
    if (inverse)                                                                                \
      ColumnStridedSpecKernel<R, true, M><<<grid, threads, 0, stream>>>(p, io, tile0);          \
    else                                                                                        \
      ColumnStridedSpecKernel<R, false, M><<<grid, threads, 0, stream>>>(p, io, tile0);         \
    return true;                                                                                \
  }
  RDL_STRIDED_SPECS(X)
#undef X
  return false;
}

bool LaunchColumnBlockSpec(bool convolve, const ColPlan& p, const ColIo& io, int tile0, dim3 grid,
                           int threads, size_t smem, cudaStream_t stream) {
#define X(A, B, M)                                                                              \
  if (p.ra == A && p.rb == B) {                                                                 \
    if (convolve) {                                                                             \
      AllowSharedMemory(ColumnBlockSpecKernel<A, B, true, M>, smem);                            \
      ColumnBlockSpecKernel<A, B, true, M><<<grid, threads, smem, stream>>>(p, io, tile0);      \
    } else {                                                                                    \
      AllowSharedMemory(ColumnBlockSpecKernel<A, B, false, M>, smem);                           \
      ColumnBlockSpecKernel<A, B, false, M><<<grid, threads, smem, stream>>>(p, io, tile0);     \
    }                                                                                           \
    return true;                                                                                \
  }
  RDL_BLOCK_SPECS(X)
#undef X
  return false;
}

}  // namespace fast
}  // namespace rdl
