// IUWT (isotropic undecimated wavelet transform) decomposition on the device — the B200-native
// counterpart of radler::algorithms::iuwt::IuwtDecomposition
//   cpp/algorithms/iuwt/iuwt_decomposition.cc:9-53    DecomposeMt
//   cpp/algorithms/iuwt/iuwt_decomposition.cc:55-237  convolveMT (horizontal + vertical B3 pass)
//   cpp/algorithms/iuwt/iuwt_decomposition.h:121-169  Recompose / ApplyMask, :249-267 convolve
//
// One smoothing is ONE kernel: the reference's row pass -> scratch image -> column pass
// (16 bytes/pixel of traffic) becomes a shared-memory tile; the horizontally smoothed rows a
// tile needs never leave the SM. Because the a-trous taps sit at multiples of d = 2^scale - 1,
// a CTA owns a COMB of rows {r + k d}: TY output rows need only TY + 4 smoothed rows whatever
// the scale (a block of contiguous rows would need TY + 4 d). The difference (w = i0 - i2,
// iuwt_decomposition.cc:25,39) and the recomposition add / ApplyMask (iuwt_decomposition.h:141,
// 162-169) are fused as epilogues, so a decomposition moves 20 bytes/pixel/scale instead of the
// reference's 44 and no image is ever copied (i0 = i1 is a pointer swap).
//
// Arithmetic parity: taps {1,4,6,4,1}/16; only the 6/16 product is inexact. The reference's
// -O3 build (GCC, -ffp-contract=fast) fuses the first product of each tap expression with the
// second one left as a rounded product, and every later term as fma(v, h, acc); the "simple"
// flavour accumulates fma(v, h, acc) onto zero in tap order. Both are written explicitly
// (this file is compiled with --fmad=false) and checked bit-exactly against the reference's
// own sources (tests/test_gpu_iuwt.py).
#include "iuwt.cuh"

#include <algorithm>

namespace rdl {
namespace {

constexpr int kTX = 128;      // tile width (columns, contiguous)
constexpr int kTY = 28;       // comb rows per tile -> 32 smoothed rows in shared memory
constexpr int kThreads = 256;

__device__ __forceinline__ float TapWeight(int k) {
  return k == 2 ? 0.375f : ((k == 1 || k == 3) ? 0.25f : 0.0625f);
}

// v[k], k = 0..4 = taps at offsets (k - 2) d; bit k of `present` set when the tap is inside the
// image. ORDER_MT_H: tap order h2,h1,h0,h3,h4 (convolveHorizontalFast), else h0..h4.
template <bool SIMPLE, bool ORDER_MT_H>
__device__ __forceinline__ float Taps(const float (&v)[5], unsigned present) {
  if (present == 31u) {  // interior fast path, same operation order as the general one
    if (SIMPLE) {
      float acc = __fmaf_rn(v[0], 0.0625f, 0.0f);
      acc = __fmaf_rn(v[1], 0.25f, acc);
      acc = __fmaf_rn(v[2], 0.375f, acc);
      acc = __fmaf_rn(v[3], 0.25f, acc);
      return __fmaf_rn(v[4], 0.0625f, acc);
    }
    if (ORDER_MT_H) {
      float acc = __fmaf_rn(v[2], 0.375f, __fmul_rn(v[1], 0.25f));
      acc = __fmaf_rn(v[0], 0.0625f, acc);
      acc = __fmaf_rn(v[3], 0.25f, acc);
      return __fmaf_rn(v[4], 0.0625f, acc);
    }
    float acc = __fmaf_rn(v[0], 0.0625f, __fmul_rn(v[1], 0.25f));
    acc = __fmaf_rn(v[2], 0.375f, acc);
    acc = __fmaf_rn(v[3], 0.25f, acc);
    return __fmaf_rn(v[4], 0.0625f, acc);
  }
  float tv[5], th[5];
  int n = 0;
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    const int k = ORDER_MT_H ? (i == 0 ? 2 : (i == 1 ? 1 : (i == 2 ? 0 : i))) : i;
    if (present & (1u << k)) {
      // compact without dynamic indexing: shift into slot n
#pragma unroll
      for (int s = 0; s < 5; ++s)
        if (s == n) {
          tv[s] = v[k];
          th[s] = TapWeight(k);
        }
      ++n;
    }
  }
  float acc;
  if (SIMPLE) {
    acc = 0.0f;
#pragma unroll
    for (int i = 0; i < 5; ++i)
      if (i < n) acc = __fmaf_rn(tv[i], th[i], acc);
    return acc;
  }
  if (n == 1) return __fmul_rn(tv[0], th[0]);
  acc = __fmaf_rn(tv[0], th[0], __fmul_rn(tv[1], th[1]));
#pragma unroll
  for (int i = 2; i < 5; ++i)
    if (i < n) acc = __fmaf_rn(tv[i], th[i], acc);
  return acc;
}

struct SmoothParams {
  const float* in;
  float* out;
  const float* other;   // epilogue operand (EPI 1: minuend i0; EPI 2: coefficient plane)
  const uint8_t* mask;  // EPI 2: optional ApplyMask plane
  int width, height, d;
  int blocks_per_residue;  // comb blocks per residue class
  int rows_per_block;      // comb rows a CTA produces (<= kTY), balanced over the blocks
};

// EPI 0: out = B(in); 1: out = other - B(in); 2: out = B(in) + (mask ? other : 0)
template <bool SIMPLE, int EPI>
__global__ void __launch_bounds__(kThreads) IuwtSmoothKernel(const __grid_constant__ SmoothParams p) {
  __shared__ float hs[kTY + 4][kTX];
  const int W = p.width, H = p.height, d = p.d;
  const int x0 = blockIdx.x * kTX;
  const int r = blockIdx.y / p.blocks_per_residue;           // residue class of the comb
  const int kb = blockIdx.y - r * p.blocks_per_residue;      // comb block
  const int k0 = kb * p.rows_per_block;                      // first comb index of the tile
  // phase 1: horizontally smoothed comb rows k0-2 .. k0+ty+1; four outputs (20 loads) in flight
  // per thread before the first tap sum
  const int ty = p.rows_per_block;
  const int n1 = (ty + 4) * kTX;
  for (int i0 = threadIdx.x; i0 < n1; i0 += 4 * kThreads) {
    float v[4][5];
    unsigned present[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * kThreads;
      present[u] = 0;
      const int row = i / kTX, xl = i - row * kTX;
      const int x = x0 + xl;
      const long y = long(r) + long(k0 + row - 2) * d;
      if (i < n1 && y >= 0 && y < H && x < W) {
        const float* src = p.in + size_t(y) * W;
#pragma unroll
        for (int k = 0; k < 5; ++k) {
          const int xx = x + (k - 2) * d;
          const bool ok = xx >= 0 && xx < W;
          v[u][k] = ok ? __ldg(src + xx) : 0.0f;
          present[u] |= ok ? (1u << k) : 0u;
        }
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * kThreads;
      if (i < n1) (&hs[0][0])[i] = present[u] ? Taps<SIMPLE, !SIMPLE>(v[u], present[u]) : 0.0f;
    }
  }
  __syncthreads();
  // phase 2: vertical pass over the comb + epilogue
  for (int i = threadIdx.x; i < ty * kTX; i += kThreads) {
    const int row = i / kTX, xl = i - row * kTX;
    const int x = x0 + xl;
    const long y = long(r) + long(k0 + row) * d;
    if (y >= H || x >= W) continue;
    float v[5];
    unsigned present = 0;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      const long yy = y + long(k - 2) * d;
      const bool ok = yy >= 0 && yy < H;
      v[k] = hs[row + k][xl];
      present |= ok ? (1u << k) : 0u;
    }
    float s = Taps<SIMPLE, false>(v, present);
    const size_t idx = size_t(y) * W + x;
    if (EPI == 1) s = __fsub_rn(__ldg(p.other + idx), s);
    if (EPI == 2) {
      const float c = (p.mask == nullptr || p.mask[idx]) ? __ldg(p.other + idx) : 0.0f;
      s = __fadd_rn(s, c);
    }
    p.out[idx] = s;
  }
}

// out = mask ? in : 0 (ApplyMask of the first recomposed plane) or a plain copy
__global__ void MaskCopyKernel(float* out, const float* in, const uint8_t* mask, size_t n) {
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    out[i] = (mask == nullptr || mask[i]) ? in[i] : 0.0f;
}

// One pass of the separable smoothing on its own (convolveHorizontalFast / convolveVerticalPartialFast,
// iuwt_decomposition.cc:78-237): only the aliased decomposition below needs the passes apart.
template <bool VERTICAL>
__global__ void IuwtSinglePassKernel(const float* __restrict__ in, float* __restrict__ out, int W, int H,
                                     int d) {
  const size_t n = size_t(W) * H;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    const int y = int(i / W), x = int(i - size_t(y) * W);
    float v[5];
    unsigned present = 0;
#pragma unroll
    for (int k = 0; k < 5; ++k) {
      const long xx = VERTICAL ? x : long(x) + long(k - 2) * d;
      const long yy = VERTICAL ? long(y) + long(k - 2) * d : y;
      const bool ok = xx >= 0 && xx < W && yy >= 0 && yy < H;
      v[k] = ok ? __ldg(in + size_t(yy) * W + xx) : 0.0f;
      present |= ok ? (1u << k) : 0u;
    }
    out[i] = Taps<false, !VERTICAL>(v, present);
  }
}

// convolveHorizontalFast at step 1 with output == input, which is what
// `iuwt.Decompose(scratch.Data(), scratch.Data(), false)` does to its first smoothing
// (iuwt_deconvolution_algorithm.cc:344,385,685,780 -> iuwt_decomposition.cc:19,58-61): pixel x is
// written before pixel x + 1 is computed, so the two LEFT taps read values that were already
// smoothed -- a recursive filter along the row. One lane per row, 32-column tiles through shared
// memory so that global accesses stay coalesced.
__global__ void __launch_bounds__(128) IuwtAliasedRowPassKernel(float* img, int W, int H) {
  __shared__ float tile[4][32][35];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row0 = (blockIdx.x * 4 + warp) * 32;
  if (row0 >= H) return;
  float (*t)[35] = tile[warp];
  float n1 = 0.0f, n2 = 0.0f;  // the new values of pixels x - 1 and x - 2
  const int row = row0 + lane;
  for (int x0 = 0; x0 < W; x0 += 32) {
    for (int r = 0; r < 32; ++r) {
      const int y = row0 + r;
      if (y < H) {
        if (x0 + lane < W) t[r][lane] = img[size_t(y) * W + x0 + lane];
        if (lane < 2 && x0 + 32 + lane < W) t[r][32 + lane] = img[size_t(y) * W + x0 + 32 + lane];
      }
    }
    __syncwarp();
    if (row < H) {
      for (int c = 0; c < 32 && x0 + c < W; ++c) {
        const int x = x0 + c;
        float v[5];
        unsigned present = 4u;
        v[2] = t[lane][c];
        v[1] = n1;
        v[0] = n2;
        v[3] = x + 1 < W ? t[lane][c + 1] : 0.0f;
        v[4] = x + 2 < W ? t[lane][c + 2] : 0.0f;
        present |= (x >= 1 ? 2u : 0u) | (x >= 2 ? 1u : 0u) | (x + 1 < W ? 8u : 0u) | (x + 2 < W ? 16u : 0u);
        const float o = Taps<false, true>(v, present);
        t[lane][c] = o;
        n2 = n1;
        n1 = o;
      }
    }
    __syncwarp();
    for (int r = 0; r < 32; ++r) {
      const int y = row0 + r;
      if (y < H && x0 + lane < W) img[size_t(y) * W + x0 + lane] = t[r][lane];
    }
    __syncwarp();
  }
}

template <bool SIMPLE, int EPI>
void LaunchSmooth(rdl_handle& h, const float* in, float* out, const float* other,
                  const uint8_t* mask, size_t width, size_t height, int scale, cudaStream_t stream) {
  SmoothParams p{};
  p.in = in;
  p.out = out;
  p.other = other;
  p.mask = mask;
  p.width = int(width);
  p.height = int(height);
  p.d = (1 << scale) - 1;
  const int residues = std::min<long>(p.d, long(height));
  const int comb_rows = (int(height) + p.d - 1) / p.d;  // rows of the longest comb
  p.blocks_per_residue = (comb_rows + kTY - 1) / kTY;
  p.rows_per_block = (comb_rows + p.blocks_per_residue - 1) / p.blocks_per_residue;
  dim3 grid((unsigned(width) + kTX - 1) / kTX, unsigned(residues) * p.blocks_per_residue);
  IuwtSmoothKernel<SIMPLE, EPI><<<grid, kThreads, 0, stream>>>(p);
  RDL_CUDA(cudaGetLastError());
  ++h.launches;
}

void CheckScale(size_t width, size_t height, int scale) {
  // the reference's loops assume 2 d <= size - 2 d (iuwt_decomposition.cc:90-95): callers keep
  // scale <= EndScale(min(W,H)) = log2(size) - 3 (iuwt_decomposition.h:182-184)
  const size_t d = (size_t(1) << scale) - 1;
  if (scale < 1 || scale > 20 || 4 * d > std::min(width, height))
    throw std::invalid_argument("IUWT scale " + std::to_string(scale) + " too large for a " +
                                std::to_string(width) + " x " + std::to_string(height) + " image");
}

}  // namespace

void IuwtSmooth(rdl_handle& h, const float* in, float* out, size_t width, size_t height, int scale,
                bool simple, cudaStream_t stream) {
  CheckScale(width, height, scale);
  if (in == out) throw std::invalid_argument("IUWT smoothing is out of place");
  if (simple)
    LaunchSmooth<true, 0>(h, in, out, nullptr, nullptr, width, height, scale, stream);
  else
    LaunchSmooth<false, 0>(h, in, out, nullptr, nullptr, width, height, scale, stream);
}

namespace {
// scales first_scale .. n_scales - 1 of DecomposeMt, starting from the smooth image i0; `next`
// tells which of the two ping-pong planes is free
void DecomposeFrom(rdl_handle& h, const float* i0, int first_scale, int next, size_t width,
                   size_t height, int n_scales, bool include_largest, float* coefficients,
                   cudaStream_t stream);
}  // namespace

void IuwtDecompose(rdl_handle& h, const float* input, size_t width, size_t height, int n_scales,
                   bool include_largest, float* coefficients, cudaStream_t stream) {
  if (n_scales < 1) throw std::invalid_argument("IUWT needs at least one scale");
  CheckScale(width, height, n_scales);
  const size_t a = width * height;
  h.iuwt_a.Reserve(a);
  h.iuwt_b.Reserve(a);
  DecomposeFrom(h, input, 0, 0, width, height, n_scales, include_largest, coefficients, stream);
}

// IuwtDecomposition::Decompose(x, x, false): input and scratch are the SAME buffer. What the
// reference's code then computes (and what is reproduced here) is, for scale 0,
//   x  <- recursive row pass of x          (convolveHorizontalPartial(scratch = x, image = x))
//   i1 <- column pass of x
//   x  <- row pass of i1                   (scratch of the second convolveMT)
//   w0 <- x - B(i1)                        (differenceMT reads `input`, which is x)
// and the ordinary decomposition of i1 for the scales above. x is left holding scratch data.
void IuwtDecomposeAliased(rdl_handle& h, float* x, size_t width, size_t height, int n_scales,
                          float* coefficients, cudaStream_t stream) {
  if (n_scales < 1) throw std::invalid_argument("IUWT needs at least one scale");
  CheckScale(width, height, n_scales);
  const size_t a = width * height;
  h.iuwt_a.Reserve(a);
  h.iuwt_b.Reserve(a);
  float* i1 = h.iuwt_a.get();
  const int W = int(width), H = int(height);
  IuwtAliasedRowPassKernel<<<(H + 127) / 128, 128, 0, stream>>>(x, W, H);
  const int grid = int(std::min<size_t>((a + 255) / 256, size_t(kSmCount) * 16));
  IuwtSinglePassKernel<true><<<grid, 256, 0, stream>>>(x, i1, W, H, 1);
  IuwtSinglePassKernel<false><<<grid, 256, 0, stream>>>(i1, x, W, H, 1);
  RDL_CUDA(cudaGetLastError());
  h.launches += 3;
  LaunchSmooth<false, 1>(h, i1, coefficients, x, nullptr, width, height, 1, stream);
  if (n_scales > 1)
    DecomposeFrom(h, i1, 1, 1, width, height, n_scales, false, coefficients, stream);
}

namespace {
void DecomposeFrom(rdl_handle& h, const float* i0, int first_scale, int next, size_t width,
                   size_t height, int n_scales, bool include_largest, float* coefficients,
                   cudaStream_t stream) {
  const size_t a = width * height;
  float* bufs[2] = {h.iuwt_a.get(), h.iuwt_b.get()};
  for (int s = first_scale; s != n_scales; ++s) {
    const bool last = s + 1 == n_scales;
    // i1 = B(i0): the smooth remainder goes straight to its plane on the last scale
    float* i1 = (last && include_largest) ? coefficients + size_t(n_scales) * a : bufs[next];
    LaunchSmooth<false, 0>(h, i0, i1, nullptr, nullptr, width, height, s + 1, stream);
    // w_s = i0 - B(i1)
    LaunchSmooth<false, 1>(h, i1, coefficients + size_t(s) * a, i0, nullptr, width, height, s + 1,
                           stream);
    i0 = i1;  // "i0 = i1" of the reference without the copy
    next ^= 1;
  }
}
}  // namespace

void IuwtRecompose(rdl_handle& h, const float* coefficients, size_t width, size_t height,
                   int n_scales, bool include_largest, const uint8_t* masks, float* out,
                   cudaStream_t stream) {
  if (n_scales < 1) throw std::invalid_argument("IUWT needs at least one scale");
  CheckScale(width, height, n_scales);
  const size_t a = width * height;
  h.iuwt_a.Reserve(a);
  const bool start_with_largest = include_largest && masks == nullptr;
  // smoothing steps still to run after the first plane is in place
  int steps = start_with_largest ? n_scales : n_scales - 1;
  float* bufs[2] = {out, h.iuwt_a.get()};
  int cur = steps % 2;  // so that the last step lands in `out`
  const int first_scale = start_with_largest ? n_scales : n_scales - 1;
  const float* first = coefficients + size_t(first_scale) * a;
  const uint8_t* first_mask = (masks && !start_with_largest) ? masks + size_t(first_scale) * a : nullptr;
  MaskCopyKernel<<<kSmCount * 8, 256, 0, stream>>>(bufs[cur], first, first_mask, a);
  RDL_CUDA(cudaGetLastError());
  ++h.launches;
  for (int s = first_scale - 1; s != -1; --s) {
    LaunchSmooth<true, 2>(h, bufs[cur], bufs[cur ^ 1], coefficients + size_t(s) * a,
                          masks ? masks + size_t(s) * a : nullptr, width, height, s + 1, stream);
    cur ^= 1;
  }
}

}  // namespace rdl
