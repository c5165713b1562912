// Shared device/host definitions of the B200-native Radler minor-loop library.
#ifndef RDL_COMMON_CUH_
#define RDL_COMMON_CUH_

#include <cuda_runtime.h>

#include <cfloat>
#include <cstdint>
#include <stdexcept>
#include <string>

namespace rdl {

constexpr int kMaxImages = 64;  // n_channels * n_pol per cube
constexpr int kSmCount = 148;   // B200

struct CudaError : std::runtime_error {
  using std::runtime_error::runtime_error;
};

#define RDL_CUDA(expr)                                                                   \
  do {                                                                                   \
    cudaError_t err__ = (expr);                                                          \
    if (err__ != cudaSuccess)                                                            \
      throw ::rdl::CudaError(std::string(#expr) + ": " + cudaGetErrorString(err__));     \
  } while (0)

// Pointer table passed by value to kernels.
struct PtrTable {
  float* p[kMaxImages];
};
struct ConstPtrTable {
  const float* p[kMaxImages];
};

// How an ImageSet is collapsed to one image (image_set.cc:309-462). Built on the
// host from the cube descriptor; the arithmetic order below is the reference's.
enum IntegMode : int {
  kIntegCopy = 0,          // single image: dest = image
  kIntegLinear = 1,        // dest = img0*w0; dest = fma(img_k, w_k, dest); dest *= factor
  kIntegSquareNormal = 2,  // per channel q = sqrt(sum_pol v^2); weighted like linear
  kIntegSquareSingle = 3,  // one channel, several pols: sqrt(sum_pol v^2) * factor
  kIntegSquared = 4,       // squared joins: sqrt(sum w v^2) * factor
  kIntegZero = 5           // weight sum is zero
};

struct IntegSpec {
  int mode;
  int n_terms;     // linear/squared: number of (image, weight) terms; square-normal: channels
  int n_pol_used;  // square modes: linked polarizations per channel
  float factor;
  int img[kMaxImages];  // image index per term (square-normal: n_terms * n_pol_used entries)
  float w[kMaxImages];  // weight per term / per channel
};

// ---- peak key --------------------------------------------------------------------
// peak_finder.cc:19-56,97-131: strict '>' against FLT_MIN, first (lowest) row-major
// index wins ties, NaN never wins, the signed value at the winner is returned.
// 64-bit key, larger is better:  [ |v| bits (31) | ~index (32) | sign (1) ].
__host__ __device__ inline unsigned long long MakePeakKey(float v, bool allow_negative,
                                                          uint32_t index) {
  const float a = allow_negative ? fabsf(v) : v;
  if (!(a > FLT_MIN)) return 0ull;
#ifdef __CUDA_ARCH__
  const uint32_t bits = __float_as_uint(a);
#else
  uint32_t bits;
  memcpy(&bits, &a, 4);
#endif
  return ((unsigned long long)bits << 33) | ((unsigned long long)(~index) << 1) |
         (v < 0.0f ? 1ull : 0ull);
}

__host__ __device__ inline bool DecodePeakKey(unsigned long long key, float& value,
                                              uint32_t& index) {
  if (key == 0ull) return false;
  const uint32_t bits = uint32_t(key >> 33);
#ifdef __CUDA_ARCH__
  float a = __uint_as_float(bits);
#else
  float a;
  memcpy(&a, &bits, 4);
#endif
  value = (key & 1ull) ? -a : a;
  index = ~uint32_t(key >> 1);
  return true;
}

#ifdef __CUDACC__
// Integrated value of pixel i; explicit fmaf mirrors the contraction of the
// reference's -O3 build (SURVEY 7.3 iv), __fmul_rn keeps products un-fused.
__device__ __forceinline__ float IntegratePixel(const IntegSpec& s, const ConstPtrTable& t,
                                                size_t i) {
  switch (s.mode) {
    case kIntegCopy:
      return t.p[s.img[0]][i];
    case kIntegLinear: {
      float acc = __fmul_rn(t.p[s.img[0]][i], s.w[0]);
      for (int k = 1; k < s.n_terms; ++k) acc = __fmaf_rn(t.p[s.img[k]][i], s.w[k], acc);
      return __fmul_rn(acc, s.factor);
    }
    case kIntegSquareNormal: {
      float acc = 0.0f;
      for (int c = 0; c < s.n_terms; ++c) {
        float q;
        if (s.n_pol_used == 1) {
          q = t.p[s.img[c]][i];
        } else {
          const float v0 = t.p[s.img[c * s.n_pol_used]][i];
          q = __fmul_rn(v0, v0);
          for (int p = 1; p < s.n_pol_used; ++p) {
            const float v = t.p[s.img[c * s.n_pol_used + p]][i];
            q = __fmaf_rn(v, v, q);
          }
          q = __fsqrt_rn(q);
        }
        acc = c == 0 ? __fmul_rn(q, s.w[0]) : __fmaf_rn(q, s.w[c], acc);
      }
      return __fmul_rn(acc, s.factor);
    }
    case kIntegSquareSingle: {
      const float v0 = t.p[s.img[0]][i];
      float q = __fmul_rn(v0, v0);
      for (int p = 1; p < s.n_pol_used; ++p) {
        const float v = t.p[s.img[p]][i];
        q = __fmaf_rn(v, v, q);
      }
      return __fmul_rn(__fsqrt_rn(q), s.factor);
    }
    case kIntegSquared: {
      const float v0 = t.p[s.img[0]][i];
      float acc = __fmul_rn(v0, __fmul_rn(v0, s.w[0]));
      for (int k = 1; k < s.n_terms; ++k) {
        const float v = t.p[s.img[k]][i];
        acc = __fmaf_rn(__fmul_rn(v, v), s.w[k], acc);
      }
      return __fmul_rn(__fsqrt_rn(acc), s.factor);
    }
    default:
      return 0.0f;
  }
}

__device__ __forceinline__ unsigned long long WarpMaxU64(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long other = __shfl_xor_sync(0xffffffffu, v, o);
    v = other > v ? other : v;
  }
  return v;
}

// Block-wide max; result valid in thread 0. `smem` holds >= 32 entries.
__device__ __forceinline__ unsigned long long BlockMaxU64(unsigned long long v,
                                                          unsigned long long* smem) {
  v = WarpMaxU64(v);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (lane == 0) smem[warp] = v;
  __syncthreads();
  if (warp == 0) {
    const int n_warps = (blockDim.x + 31) >> 5;
    v = lane < n_warps ? smem[lane] : 0ull;
    v = WarpMaxU64(v);
  }
  return v;
}
#endif  // __CUDACC__

}  // namespace rdl
#endif
