// Hand-written FFT convolution engine for sm_100a — the device-side restatement of
// schaapcommon::math::Convolve / PrepareConvolutionKernel / PrepareSmallConvolutionKernel as
// Radler calls them (multiscale_transforms.cc:11-23, subminor_loop.cc:195-218).
//
//   out = irfft2( rfft2(in) * rfft2(kernel) ) / (W*H)          (circular, in place)
//
// Every 1-D transform is done entirely in shared memory:
//   RowForwardKernel      real rows -> half spectrum (R2C through a length-W/2 complex FFT)
//   ColumnStridedKernel   four-step column transform, step 1: H2-point transforms over rows at
//                         stride H1 + step twiddles, on 16-column (128-byte) tiles
//   ColumnBlockKernel     step 2 on H1 contiguous rows: forward FFT, * kernel spectrum, inverse
//                         FFT in one pass (the whole column transform when H <= 512)
//   ColumnStridedKernel   step 1 undone (inverse)
//   RowInverseKernel      half spectrum -> real rows, fused with the caller's epilogue
//                      (store / residual -= result / model += result) and with the
//                      reference's Untrim/Trim (centred zero-pad and crop).
// The 1-D transforms are in-place decimation-in-frequency forward and decimation-in-time
// inverse passes (radix 8/4/2/3/5/7, generic prime radix <= 127), so no bit-reversal pass is
// ever needed: the column spectra stay in digit-reversed order (the kernel spectrum is
// produced by the same forward pass), only the row split step looks frequencies up through a
// small permutation table. Algorithmic HBM traffic: 28 bytes per padded pixel;
// the four-step column transform moves 44.
// Results are not bit-comparable with FFTW or the CPU checker; parity is by tolerance.
//
// Embedded frames. GenericClean pads to even(ceil(1.1f * N)) (generic_clean.cc:63-66) and facets
// have data-dependent sizes, so a frame length can contain a large prime (8192 -> 9012 = 4 * 3 *
// 751) for which FFTW falls back to Rader/Bluestein. Here such a frame is NOT transformed
// directly: the circular convolution at P is, term for term, the sum
//   out[x] = sum_x' src[x'] * k[(x - x') mod P],
// and with sources and destinations confined to `content` pixels only displacements |d| < content
// occur. Writing the kernel out over those displacements, K[d mod Q] = k[d mod P], gives the
// identical sum as a circular convolution at any Q >= 2 * content - 1, so Q is chosen 2-3-5-7
// smooth and the native engine is used. The result is the reference's circular convolution at P
// (including its wrap-around), evaluated at FFT round-off like every other convolution here.
#include "fft.cuh"
#include "fft_fast_plan.h"
#include "fft_spec.cuh"

#include <algorithm>
#include <cmath>
#include <cstdlib>
#include <map>
#include <vector>

namespace rdl {

size_t CalculateGoodFFTSize(size_t minimum_size) {  // fft_size_calculations.h:15-29
  size_t best = 2 * minimum_size;
  for (size_t f2 = 2; f2 < best; f2 *= 2)
    for (size_t f23 = f2; f23 < best; f23 *= 3)
      for (size_t f235 = f23; f235 < best; f235 *= 5)
        for (size_t f2357 = f235; f2357 < best; f2357 *= 7)
          if (f2357 >= minimum_size) best = f2357;
  return best;
}

size_t GetConvolutionSize(double scale, size_t original_size, double padding) {  // :39-50
  return CalculateGoodFFTSize(size_t(std::ceil(padding * (scale * 1.5 + double(original_size)))));
}

namespace {

constexpr int kMaxStages = 24;
constexpr int kMaxGenericRadix = 127;
constexpr size_t kMaxSmemBytes = 227 * 1024;

struct Plan {
  int n;
  int n_stages;
  int radix[kMaxStages];
  int len[kMaxStages];  // block length the stage works on (n, n/r0, n/(r0 r1), ...)
  const float2* tw;     // exp(-2 pi i t / n), t in [0, n)
};

// One padding slot per 32 complex values keeps the strided late stages off the same banks.
// one pad slot per 32 elements (unit-stride butterflies) and one more per 512: the digit-reversed
// gathers of the split phase walk with stride n/8, which the first term alone maps to one bank
__host__ __device__ __forceinline__ int Pad(int i) { return i + (i >> 5) + (i >> 9); }
inline int PaddedLength(int n) { return (Pad(n - 1) + 1) | 1; }

__device__ __forceinline__ float2 operator+(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 operator-(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
__device__ __forceinline__ float2 operator*(float s, float2 a) { return make_float2(s * a.x, s * a.y); }
__device__ __forceinline__ float2 CMul(float2 a, float2 b) {
  return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
__device__ __forceinline__ float2 Conj(float2 a) { return make_float2(a.x, -a.y); }
// forward: * (-i); inverse: * (+i)
template <bool INV>
__device__ __forceinline__ float2 MulI(float2 a) {
  return INV ? make_float2(-a.y, a.x) : make_float2(a.y, -a.x);
}

template <int R, bool INV>
struct Dft;

template <bool INV>
struct Dft<2, INV> {
  __device__ __forceinline__ static void Run(float2 (&v)[2]) {
    const float2 t = v[0];
    v[0] = t + v[1];
    v[1] = t - v[1];
  }
};

template <bool INV>
struct Dft<4, INV> {
  __device__ __forceinline__ static void Run(float2 (&v)[4]) {
    const float2 t0 = v[0] + v[2], t1 = v[0] - v[2], t2 = v[1] + v[3];
    const float2 t3 = MulI<INV>(v[1] - v[3]);
    v[0] = t0 + t2;
    v[1] = t1 + t3;
    v[2] = t0 - t2;
    v[3] = t1 - t3;
  }
};

template <bool INV>
struct Dft<8, INV> {
  __device__ __forceinline__ static void Run(float2 (&v)[8]) {
    float2 e[4] = {v[0], v[2], v[4], v[6]};
    float2 o[4] = {v[1], v[3], v[5], v[7]};
    Dft<4, INV>::Run(e);
    Dft<4, INV>::Run(o);
    constexpr float c = 0.70710678118654752440f;
    const float2 o1 = INV ? make_float2(c * (o[1].x - o[1].y), c * (o[1].x + o[1].y))
                          : make_float2(c * (o[1].x + o[1].y), c * (o[1].y - o[1].x));
    const float2 o2 = MulI<INV>(o[2]);
    const float2 o3 = INV ? make_float2(c * (-o[3].x - o[3].y), c * (o[3].x - o[3].y))
                          : make_float2(c * (o[3].y - o[3].x), c * (-o[3].y - o[3].x));
    v[0] = e[0] + o[0];
    v[4] = e[0] - o[0];
    v[1] = e[1] + o1;
    v[5] = e[1] - o1;
    v[2] = e[2] + o2;
    v[6] = e[2] - o2;
    v[3] = e[3] + o3;
    v[7] = e[3] - o3;
  }
};

template <int P>
struct Trig;
template <>
struct Trig<3> {
  __device__ __forceinline__ static float C(int t) { return t == 0 ? 1.0f : -0.5f; }
  __device__ __forceinline__ static float S(int t) {
    return t == 0 ? 0.0f : (t == 1 ? 0.86602540378443864676f : -0.86602540378443864676f);
  }
};
template <>
struct Trig<5> {
  __device__ __forceinline__ static float C(int t) {
    return t == 0 ? 1.0f : ((t == 1 || t == 4) ? 0.30901699437494742410f : -0.80901699437494742410f);
  }
  __device__ __forceinline__ static float S(int t) {
    return t == 0 ? 0.0f
                  : (t == 1 ? 0.95105651629515357212f
                            : (t == 2 ? 0.58778525229247312917f
                                      : (t == 3 ? -0.58778525229247312917f
                                                : -0.95105651629515357212f)));
  }
};
template <>
struct Trig<7> {
  __device__ __forceinline__ static float C(int t) {
    return t == 0 ? 1.0f
                  : ((t == 1 || t == 6) ? 0.62348980185873353053f
                                        : ((t == 2 || t == 5) ? -0.22252093395631440429f
                                                              : -0.90096886790241912624f));
  }
  __device__ __forceinline__ static float S(int t) {
    const float s1 = 0.78183148246802980871f, s2 = 0.97492791218182360702f,
                s3 = 0.43388373911755812048f;
    return t == 0 ? 0.0f
                  : (t == 1 ? s1 : (t == 2 ? s2 : (t == 3 ? s3 : (t == 4 ? -s3 : (t == 5 ? -s2 : -s1)))));
  }
};

template <int P, bool INV>
struct DftOdd {
  __device__ __forceinline__ static void Run(float2 (&v)[P]) {
    constexpr int H = (P - 1) / 2;
    float2 a[H], b[H];
#pragma unroll
    for (int m = 1; m <= H; ++m) {
      a[m - 1] = v[m] + v[P - m];
      b[m - 1] = v[m] - v[P - m];
    }
    const float2 x0 = v[0];
    float2 sum = x0;
#pragma unroll
    for (int m = 0; m < H; ++m) sum = sum + a[m];
#pragma unroll
    for (int q = 1; q <= H; ++q) {
      float2 c = x0, s = make_float2(0.0f, 0.0f);
#pragma unroll
      for (int m = 1; m <= H; ++m) {
        const int t = (q * m) % P;
        c = c + Trig<P>::C(t) * a[m - 1];
        s = s + Trig<P>::S(t) * b[m - 1];
      }
      // forward: X_q = c - i s, X_{P-q} = c + i s; inverse: swapped
      const float2 lo = make_float2(c.x + s.y, c.y - s.x), hi = make_float2(c.x - s.y, c.y + s.x);
      v[q] = INV ? hi : lo;
      v[P - q] = INV ? lo : hi;
    }
    v[0] = sum;
  }
};
template <bool INV>
struct Dft<3, INV> : DftOdd<3, INV> {};
template <bool INV>
struct Dft<5, INV> : DftOdd<5, INV> {};
template <bool INV>
struct Dft<7, INV> : DftOdd<7, INV> {};

// One in-place radix-R pass over n_seq sequences of length n held in shared memory.
// DIF (forward): r-point DFT then twiddle; DIT (inverse): conjugate twiddle then inverse DFT.
template <int R, bool INV>
__device__ __forceinline__ void Stage(float2* s, int n_seq, int seq_stride, int n, int L,
                                      const float2* __restrict__ tw) {
  const int m = L / R;
  const int tw_step = n / L;
  const int n_bf = n / R;
  for (int t = threadIdx.x; t < n_bf; t += blockDim.x) {
    const int b = t / m, j = t - b * m;
    const int base = b * L + j;
    float2 w[R];
    if (m > 1) {
      // every power straight from the double-rounded table: chaining products of w^j costs
      // about 4x in accuracy, which the reference's 2e-6 point-source test does not allow
#pragma unroll
      for (int q = 1; q < R; ++q) {
        w[q] = __ldg(tw + j * q * tw_step);
        if (INV) w[q].y = -w[q].y;
      }
    }
    for (int c = 0; c < n_seq; ++c) {
      float2* p = s + c * seq_stride;
      float2 v[R];
#pragma unroll
      for (int q = 0; q < R; ++q) v[q] = p[Pad(base + q * m)];
      if (INV && m > 1) {
#pragma unroll
        for (int q = 1; q < R; ++q) v[q] = CMul(v[q], w[q]);
      }
      Dft<R, INV>::Run(v);
      if (!INV && m > 1) {
#pragma unroll
        for (int q = 1; q < R; ++q) v[q] = CMul(v[q], w[q]);
      }
#pragma unroll
      for (int q = 0; q < R; ++q) p[Pad(base + q * m)] = v[q];
    }
  }
}

// Any prime radix p <= kMaxGenericRadix by direct O(p^2) evaluation (local-memory staging).
// Only reached for the non-smooth padded sizes of GenericClean path A
// (generic_clean.cc:63-66: even(ceil(1.1f*N)), e.g. 1128 = 2^3 * 3 * 47).
template <bool INV>
__device__ void GenericStage(float2* s, int n_seq, int seq_stride, int n, int L, int p,
                             const float2* __restrict__ tw) {
  const int m = L / p;
  const int tw_step = n / L;
  const int root_step = n / p;
  const int n_bf = n / p;
  for (int t = threadIdx.x; t < n_bf; t += blockDim.x) {
    const int b = t / m, j = t - b * m;
    const int base = b * L + j;
    for (int c = 0; c < n_seq; ++c) {
      float2* ptr = s + c * seq_stride;
      float2 x[kMaxGenericRadix];
      for (int q = 0; q < p; ++q) {
        float2 v = ptr[Pad(base + q * m)];
        if (INV && m > 1) v = CMul(v, Conj(__ldg(tw + j * q * tw_step)));
        x[q] = v;
      }
      for (int q = 0; q < p; ++q) {
        float2 acc = x[0];
        int idx = 0;
        for (int k = 1; k < p; ++k) {
          idx += q;
          if (idx >= p) idx -= p;
          float2 r = __ldg(tw + idx * root_step);
          if (INV) r.y = -r.y;
          acc = acc + CMul(x[k], r);
        }
        if (!INV && m > 1) acc = CMul(acc, __ldg(tw + j * q * tw_step));
        ptr[Pad(base + q * m)] = acc;
      }
    }
  }
}

// GENERIC = false compiles the generic-prime stage (and its 1 KB local array) out of the kernel.
template <bool INV, bool GENERIC>
__device__ __forceinline__ void RunStage(float2* s, int n_seq, int seq_stride, const Plan& plan,
                                         int st) {
  const int n = plan.n, L = plan.len[st];
  switch (plan.radix[st]) {
    case 8: Stage<8, INV>(s, n_seq, seq_stride, n, L, plan.tw); break;
    case 4: Stage<4, INV>(s, n_seq, seq_stride, n, L, plan.tw); break;
    case 2: Stage<2, INV>(s, n_seq, seq_stride, n, L, plan.tw); break;
    case 3: Stage<3, INV>(s, n_seq, seq_stride, n, L, plan.tw); break;
    case 5: Stage<5, INV>(s, n_seq, seq_stride, n, L, plan.tw); break;
    case 7: Stage<7, INV>(s, n_seq, seq_stride, n, L, plan.tw); break;
    default:
      if (GENERIC) GenericStage<INV>(s, n_seq, seq_stride, n, L, plan.radix[st], plan.tw);
      break;
  }
}

// Forward: natural order in, digit-reversed out. Inverse: digit-reversed in, natural out.
// Both leave the block synchronised.
template <bool INV, bool GENERIC>
__device__ __forceinline__ void RunFft(float2* s, int n_seq, int seq_stride, const Plan& plan) {
  if (!INV) {
    for (int st = 0; st < plan.n_stages; ++st) {
      RunStage<false, GENERIC>(s, n_seq, seq_stride, plan, st);
      __syncthreads();
    }
  } else {
    for (int st = plan.n_stages - 1; st >= 0; --st) {
      RunStage<true, GENERIC>(s, n_seq, seq_stride, plan, st);
      __syncthreads();
    }
  }
}

// ---- column transforms on [len][TC] tiles ----------------------------------------------
// A tile holds TC adjacent spectrum columns (TC * 8 bytes = whole 128-byte lines of every
// spectrum row) with the column index fastest, so global loads/stores are full lines and the
// TC lanes that share a butterfly index touch consecutive shared-memory words.
template <int R, bool INV, int TC>
__device__ __forceinline__ void TileStage(float2* s, int n, int L, const float2* __restrict__ tw) {
  const int m = L / R;
  const int tw_step = n / L;
  const int work = (n / R) * TC;
  for (int i = threadIdx.x; i < work; i += blockDim.x) {
    const int c = i % TC, t = i / TC;
    const int b = t / m, j = t - b * m;
    float2* p = s + (b * L + j) * TC + c;
    float2 v[R];
#pragma unroll
    for (int q = 0; q < R; ++q) v[q] = p[q * m * TC];
    if (INV && m > 1) {
#pragma unroll
      for (int q = 1; q < R; ++q) v[q] = CMul(v[q], Conj(tw[j * q * tw_step]));
    }
    Dft<R, INV>::Run(v);
    if (!INV && m > 1) {
#pragma unroll
      for (int q = 1; q < R; ++q) v[q] = CMul(v[q], tw[j * q * tw_step]);
    }
#pragma unroll
    for (int q = 0; q < R; ++q) p[q * m * TC] = v[q];
  }
}

template <bool INV, int TC>
__device__ void TileGenericStage(float2* s, int n, int L, int p, const float2* __restrict__ tw) {
  const int m = L / p;
  const int tw_step = n / L;
  const int root_step = n / p;
  const int work = (n / p) * TC;
  for (int i = threadIdx.x; i < work; i += blockDim.x) {
    const int c = i % TC, t = i / TC;
    const int b = t / m, j = t - b * m;
    float2* ptr = s + (b * L + j) * TC + c;
    float2 x[kMaxGenericRadix];
    for (int q = 0; q < p; ++q) {
      float2 v = ptr[q * m * TC];
      if (INV && m > 1) v = CMul(v, Conj(tw[j * q * tw_step]));
      x[q] = v;
    }
    for (int q = 0; q < p; ++q) {
      float2 acc = x[0];
      int idx = 0;
      for (int k = 1; k < p; ++k) {
        idx += q;
        if (idx >= p) idx -= p;
        float2 r = tw[idx * root_step];
        if (INV) r.y = -r.y;
        acc = acc + CMul(x[k], r);
      }
      if (!INV && m > 1) acc = CMul(acc, tw[j * q * tw_step]);
      ptr[q * m * TC] = acc;
    }
  }
}

// Forward: natural order in, digit-reversed out; inverse the other way. Leaves the block
// synchronised.
// `tw`: the plan's twiddle table, staged in shared memory by the caller (n entries).
template <bool INV, int TC, bool GENERIC>
__device__ __forceinline__ void RunTileFft(float2* s, const Plan& plan, const float2* tw) {
  for (int k = 0; k < plan.n_stages; ++k) {
    const int st = INV ? plan.n_stages - 1 - k : k;
    const int n = plan.n, L = plan.len[st];
    switch (plan.radix[st]) {
      case 8: TileStage<8, INV, TC>(s, n, L, tw); break;
      case 4: TileStage<4, INV, TC>(s, n, L, tw); break;
      case 2: TileStage<2, INV, TC>(s, n, L, tw); break;
      case 3: TileStage<3, INV, TC>(s, n, L, tw); break;
      case 5: TileStage<5, INV, TC>(s, n, L, tw); break;
      case 7: TileStage<7, INV, TC>(s, n, L, tw); break;
      default:
        if (GENERIC) TileGenericStage<INV, TC>(s, n, L, plan.radix[st], tw);
        break;
    }
    __syncthreads();
  }
}

__device__ __forceinline__ void StageTwiddles(float2* dst, const Plan& plan) {
  for (int i = threadIdx.x; i < plan.n; i += blockDim.x) dst[i] = __ldg(plan.tw + i);
}

struct RowParams {
  // real image side
  const float* src;  // forward: source image (sw x sh)
  float* dst;        // inverse: destination image (sw x sh)
  int sw, sh;
  int ox, oy;        // position of the image inside the W x H frame (Untrim / Trim offsets)
  int W, H;
  float2* spec;      // [H][pitch] half spectrum, 2n = W
  int pitch;
  Plan plan;         // n = W / 2
  const uint32_t* perm;     // perm[k] = slot of frequency k after the forward pass
  const float2* split_tw;   // exp(-2 pi i k / W), k in [0, n]
  int rows_per_cta;
  int seq_stride;
  int mode;          // inverse epilogue: 0 store, 1 dst -= v, 2 dst += v
};

template <bool GENERIC>
__global__ void __launch_bounds__(256, 3) RowForwardKernel(const __grid_constant__ RowParams p) {
  extern __shared__ __align__(16) float2 smem[];
  const int n = p.plan.n;
  const int row0 = blockIdx.x * p.rows_per_cta;
  const int n_seq = min(p.rows_per_cta, p.sh - row0);
  // 128-bit loads when the row, its offset in the frame and the source are 16-byte aligned
  const bool vec = ((p.ox | p.sw | p.W) & 3) == 0 && (reinterpret_cast<uintptr_t>(p.src) & 15) == 0;
  for (int c = 0; c < n_seq; ++c) {
    const float* src = p.src + size_t(row0 + c) * p.sw;
    if (vec) {
      // four independent 128-bit loads in flight per thread before the first one is consumed
      float2* s2 = smem + c * p.seq_stride;
      const int step = 4 * blockDim.x;
      for (int x = 4 * threadIdx.x; x < p.W; x += 4 * step) {
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int xs = x + u * step - p.ox;
          v[u] = (x + u * step < p.W && xs >= 0 && xs < p.sw)
                     ? __ldg(reinterpret_cast<const float4*>(src + xs))
                     : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int xx = x + u * step;
          if (xx < p.W) {
            s2[Pad(xx >> 1)] = make_float2(v[u].x, v[u].y);
            s2[Pad((xx >> 1) + 1)] = make_float2(v[u].z, v[u].w);
          }
        }
      }
    } else {
      float* s = reinterpret_cast<float*>(smem + c * p.seq_stride);
      for (int x = threadIdx.x; x < p.W; x += blockDim.x) {
        const int xs = x - p.ox;
        const float v = (xs >= 0 && xs < p.sw) ? __ldg(src + xs) : 0.0f;
        s[2 * Pad(x >> 1) + (x & 1)] = v;
      }
    }
  }
  __syncthreads();
  RunFft<false, GENERIC>(smem, n_seq, p.seq_stride, p.plan);
  // split: 2 X[k] = (Z[k] + conj Z[n-k]) - i w^k (Z[k] - conj Z[n-k]),  k = 0..n
  for (int c = 0; c < n_seq; ++c) {
    const float2* s = smem + c * p.seq_stride;
    float2* out = p.spec + size_t(p.oy + row0 + c) * p.pitch;
    for (int k0 = threadIdx.x; k0 <= n; k0 += 4 * blockDim.x) {
      uint32_t pa[4], pb[4];
      float2 w[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {  // table lookups of four outputs in flight together
        const int k = k0 + u * blockDim.x;
        if (k <= n) {
          pa[u] = __ldg(p.perm + (k == n ? 0 : k));
          pb[u] = __ldg(p.perm + (k == 0 ? 0 : n - k));
          w[u] = __ldg(p.split_tw + k);
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int k = k0 + u * blockDim.x;
        if (k <= n) {
          const float2 a = s[Pad(pa[u])], b = s[Pad(pb[u])];
          const float2 sum = make_float2(a.x + b.x, a.y - b.y);
          const float2 dif = make_float2(a.x - b.x, a.y + b.y);
          const float2 t = CMul(w[u], dif);
          out[k] = make_float2(sum.x + t.y, sum.y - t.x);
        }
      }
    }
  }
}

template <bool GENERIC>
__global__ void __launch_bounds__(256, 3) RowInverseKernel(const __grid_constant__ RowParams p) {
  extern __shared__ __align__(16) float2 smem[];
  const int n = p.plan.n;
  const int row0 = blockIdx.x * p.rows_per_cta;
  const int n_seq = min(p.rows_per_cta, p.sh - row0);
  // inverse split: 2 Z[k] = (Y[k] + conj Y[n-k]) + i conj(w^k) (Y[k] - conj Y[n-k]),  k = 0..n-1
  for (int c = 0; c < n_seq; ++c) {
    float2* s = smem + c * p.seq_stride;
    const float2* in = p.spec + size_t(p.oy + row0 + c) * p.pitch;
    for (int k0 = threadIdx.x; k0 < n; k0 += 4 * blockDim.x) {
      float2 a[4], b[4], w[4];
      uint32_t slot[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {  // all global loads first
        const int k = k0 + u * blockDim.x;
        if (k < n) {
          a[u] = in[k];
          b[u] = in[n - k];
          w[u] = __ldg(p.split_tw + k);
          slot[u] = __ldg(p.perm + k);
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int k = k0 + u * blockDim.x;
        if (k < n) {
          const float2 sum = make_float2(a[u].x + b[u].x, a[u].y - b[u].y);
          const float2 dif = make_float2(a[u].x - b[u].x, a[u].y + b[u].y);
          const float2 t = CMul(Conj(w[u]), dif);
          s[Pad(slot[u])] = make_float2(sum.x - t.y, sum.y + t.x);
        }
      }
    }
  }
  __syncthreads();
  RunFft<true, GENERIC>(smem, n_seq, p.seq_stride, p.plan);
  const bool vec = ((p.ox | p.sw) & 3) == 0 && (reinterpret_cast<uintptr_t>(p.dst) & 15) == 0;
  for (int c = 0; c < n_seq; ++c) {
    const float* s = reinterpret_cast<const float*>(smem + c * p.seq_stride);
    float* dst = p.dst + size_t(row0 + c) * p.sw;
    if (vec) {
      const float2* s2 = smem + c * p.seq_stride;
      for (int xs = 4 * threadIdx.x; xs < p.sw; xs += 4 * blockDim.x) {
        const int x = xs + p.ox;
        const float2 a = s2[Pad(x >> 1)], b = s2[Pad((x >> 1) + 1)];
        float4* d4 = reinterpret_cast<float4*>(dst + xs);
        float4 v = make_float4(a.x, a.y, b.x, b.y);
        if (p.mode != 0) {
          const float4 o = *d4;
          v = p.mode == 1 ? make_float4(o.x - v.x, o.y - v.y, o.z - v.z, o.w - v.w)
                          : make_float4(o.x + v.x, o.y + v.y, o.z + v.z, o.w + v.w);
        }
        *d4 = v;
      }
      continue;
    }
    for (int xs = threadIdx.x; xs < p.sw; xs += blockDim.x) {
      const int x = xs + p.ox;
      const float v = s[2 * Pad(x >> 1) + (x & 1)];
      if (p.mode == 0)
        dst[xs] = v;
      else if (p.mode == 1)
        dst[xs] = dst[xs] - v;
      else
        dst[xs] = dst[xs] + v;
    }
  }
}

struct ColParams {
  float2* spec;          // work spectrum, transformed in place
  const float2* kspec;   // kernel spectrum to multiply with (convolve mode)
  float2* kspec_out;     // forward-only mode: where the scaled spectrum goes
  int pitch, H;
  int H1, H2;            // H = H1 * H2: column index y = r + H1 * j  (r < H1, j < H2)
  int y0, y1;            // rows of `spec` that hold data (others are zero by construction)
  int out_y0, out_y1;    // rows the inverse row pass will read
  float scale;
  Plan plan1;            // n = H1 (contiguous row blocks)
  Plan plan2;            // n = H2 (rows at stride H1)
  const float2* tw_h;    // exp(-2 pi i t / H), t in [0, H)
  const uint32_t* freq2; // frequency held by slot s after the forward H2 pass
};

constexpr int kTileCols = 16;  // 16 complex columns = one 128-byte line per spectrum row

// Four-step column transform, step 1 (forward) / step 3 (inverse): the H2-point transforms
// over rows r, r + H1, r + 2 H1, ... of one column tile, fused with the step twiddle
// w_H^(r k2). Forward leaves frequency k2 in the row block of its slot, so step 2 works on H1
// CONTIGUOUS rows. One CTA per (column tile, r).
template <bool INV, int TC, bool GENERIC>
__global__ void __launch_bounds__(256) ColumnStridedKernel(const __grid_constant__ ColParams p) {
  extern __shared__ __align__(16) float2 smem[];
  const int r = blockIdx.y;
  const int c0 = blockIdx.x * TC;
  const int H1 = p.H1, H2 = p.H2;
  constexpr int Q = TC / 2;  // 128-bit (two complex) elements per tile row
  float4* base4 = reinterpret_cast<float4*>(p.spec + size_t(r) * p.pitch + c0);
  const size_t step4 = size_t(H1) * p.pitch / 2;
  float4* sm4 = reinterpret_cast<float4*>(smem);
  float2* tw = smem + H2 * TC;
  StageTwiddles(tw, p.plan2);
  const int n4 = H2 * Q, bd = blockDim.x;
  // loads: four independent 128-bit requests per thread in flight
  for (int i0 = threadIdx.x; i0 < n4; i0 += 4 * bd) {
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * bd;
      if (i < n4) {
        const int j = i / Q, c = i - j * Q;
        const int y = r + H1 * j;
        v[u] = (INV || (y >= p.y0 && y < p.y1)) ? base4[j * step4 + c]
                                                : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * bd;
      if (i < n4) {
        if (INV) {  // undo the step twiddle on the way in
          const float2 w = Conj(__ldg(p.tw_h + r * int(__ldg(p.freq2 + i / Q))));
          const float2 a = CMul(make_float2(v[u].x, v[u].y), w);
          const float2 b = CMul(make_float2(v[u].z, v[u].w), w);
          v[u] = make_float4(a.x, a.y, b.x, b.y);
        }
        sm4[i] = v[u];
      }
    }
  }
  __syncthreads();
  RunTileFft<INV, TC, GENERIC>(smem, p.plan2, tw);
  for (int i = threadIdx.x; i < n4; i += bd) {
    const int j = i / Q, c = i - j * Q;
    float4 v = sm4[i];
    if (!INV) {
      const float2 w = __ldg(p.tw_h + r * int(__ldg(p.freq2 + j)));
      const float2 a = CMul(make_float2(v.x, v.y), w), b = CMul(make_float2(v.z, v.w), w);
      base4[j * step4 + c] = make_float4(a.x, a.y, b.x, b.y);
    } else {
      const int y = r + H1 * j;
      if (y >= p.out_y0 && y < p.out_y1) base4[j * step4 + c] = v;
    }
  }
}

// Step 2: H1-point transforms over the contiguous rows of slot s, times the kernel spectrum,
// and straight back (inverse H1-point transforms) — or, for a kernel image, the scaled
// spectrum written out. With H2 == 1 this is the whole column transform in one pass.
// The kernel-spectrum tile is fetched into registers BEFORE the forward transform so its
// latency hides behind the butterflies.
template <bool CONVOLVE, int TC, bool GENERIC>
__global__ void __launch_bounds__(512) ColumnBlockKernel(const __grid_constant__ ColParams p) {
  extern __shared__ __align__(16) float2 smem[];
  const int c0 = blockIdx.x * TC;
  const int H1 = p.H1;
  constexpr int Q = TC / 2;
  const size_t row0 = size_t(blockIdx.y) * H1;
  const size_t pitch4 = size_t(p.pitch) / 2;
  float4* base4 = reinterpret_cast<float4*>(p.spec + row0 * p.pitch + c0);
  float4* sm4 = reinterpret_cast<float4*>(smem);
  const bool whole = p.H2 == 1;
  float2* tw = smem + H1 * TC;
  StageTwiddles(tw, p.plan1);
  const int n4 = H1 * Q, bd = blockDim.x;
  for (int i0 = threadIdx.x; i0 < n4; i0 += 4 * bd) {
    float4 v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * bd;
      if (i < n4) {
        const int y = i / Q, c = i - y * Q;
        v[u] = (!whole || (y >= p.y0 && y < p.y1)) ? base4[size_t(y) * pitch4 + c]
                                                    : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = i0 + u * bd;
      if (i < n4) sm4[i] = v[u];
    }
  }
  const float4* k4 =
      CONVOLVE ? reinterpret_cast<const float4*>(p.kspec + row0 * p.pitch + c0) : nullptr;
  const bool prefetch = CONVOLVE && n4 <= 4 * bd;
  float4 kreg[4];
  if (prefetch) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = threadIdx.x + u * bd;
      if (i < n4) kreg[u] = __ldg(k4 + size_t(i / Q) * pitch4 + (i % Q));
    }
  }
  __syncthreads();
  RunTileFft<false, TC, GENERIC>(smem, p.plan1, tw);
  if (!CONVOLVE) {
    float4* out4 = reinterpret_cast<float4*>(p.kspec_out + row0 * p.pitch + c0);
    for (int i = threadIdx.x; i < n4; i += bd) {
      const int y = i / Q, c = i - y * Q;
      const float4 v = sm4[i];
      out4[size_t(y) * pitch4 + c] =
          make_float4(p.scale * v.x, p.scale * v.y, p.scale * v.z, p.scale * v.w);
    }
    return;
  }
  auto mul2 = [](float4 v, float4 k) {
    const float2 a = CMul(make_float2(v.x, v.y), make_float2(k.x, k.y));
    const float2 b = CMul(make_float2(v.z, v.w), make_float2(k.z, k.w));
    return make_float4(a.x, a.y, b.x, b.y);
  };
  if (prefetch) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int i = threadIdx.x + u * bd;
      if (i < n4) sm4[i] = mul2(sm4[i], kreg[u]);
    }
  } else {
    for (int i = threadIdx.x; i < n4; i += bd)
      sm4[i] = mul2(sm4[i], __ldg(k4 + size_t(i / Q) * pitch4 + (i % Q)));
  }
  __syncthreads();
  RunTileFft<true, TC, GENERIC>(smem, p.plan1, tw);
  for (int i = threadIdx.x; i < n4; i += bd) {
    const int y = i / Q, c = i - y * Q;
    if (!whole || (y >= p.out_y0 && y < p.out_y1)) base4[size_t(y) * pitch4 + c] = sm4[i];
  }
}

// dst (w x h) = PrepareSmallConvolutionKernel(k, n): an n x n kernel wrapped around the origin
__global__ void WrapSmallKernelKernel(float* dst, int w, int h, const float* k, int n) {
  const int half = n / 2;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n * n; i += gridDim.x * blockDim.x) {
    const int ky = i / n, kx = i - ky * n;
    const int y = (ky + h - half) % h, x = (kx + w - half) % w;
    dst[size_t(y) * w + x] = k[i];
  }
}

// dst (w x h) = PrepareConvolutionKernel(Untrim(psf, pw x ph -> w x h)): centred zero-pad, then
// the circular shift that moves (w/2, h/2) to the origin (subminor_loop.cc:199-202).
__global__ void PreparePsfKernelKernel(float* dst, int w, int h, const float* psf, int pw, int ph) {
  const int sx = (w - pw) / 2, sy = (h - ph) / 2;
  const size_t total = size_t(w) * h;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += size_t(gridDim.x) * blockDim.x) {
    const int y = int(i / w), x = int(i - size_t(y) * w);
    const int yy = (y + h / 2) % h - sy, xx = (x + w / 2) % w - sx;
    dst[i] = (yy >= 0 && yy < ph && xx >= 0 && xx < pw) ? __ldg(psf + size_t(yy) * pw + xx) : 0.0f;
  }
}

// ---- register-radix engine (fft_fast.cuh): 2-3-5-7-smooth frames ---------------------------------
__global__ void __launch_bounds__(128, 4)
    FastRowForwardKernel(const __grid_constant__ fast::RowPlan p, const __grid_constant__ fast::RowIo io) {
  extern __shared__ __align__(16) float2 smem[];
  fast::RowForwardPhase<0>(p, io, smem, blockIdx.x, threadIdx.x, blockDim.x);
  __syncthreads();
  fast::RowForwardPhase<1>(p, io, smem, blockIdx.x, threadIdx.x, blockDim.x);
  __syncthreads();
  fast::RowForwardPhase<2>(p, io, smem, blockIdx.x, threadIdx.x, blockDim.x);
}

__global__ void __launch_bounds__(128, 4)
    FastRowInverseKernel(const __grid_constant__ fast::RowPlan p, const __grid_constant__ fast::RowIo io) {
  extern __shared__ __align__(16) float2 smem[];
  fast::RowInversePhase<0>(p, io, smem, blockIdx.x, threadIdx.x, blockDim.x);
  __syncthreads();
  fast::RowInversePhase<1>(p, io, smem, blockIdx.x, threadIdx.x, blockDim.x);
  __syncthreads();
  fast::RowInversePhase<2>(p, io, smem, blockIdx.x, threadIdx.x, blockDim.x);
}

// grid: (column tiles of a band, row groups); tile0 = first tile of the band
template <bool INV>
__global__ void __launch_bounds__(fast::kTile * fast::kStridedRows, 4)
    FastColumnStridedKernel(const __grid_constant__ fast::ColPlan p, const __grid_constant__ fast::ColIo io,
                            int tile0) {
  fast::ColumnStridedThread<INV>(p, io, tile0 + blockIdx.x, blockIdx.y, threadIdx.x);
}

// grid: (column tiles of a band, H2 row blocks)
template <bool CONVOLVE>
__global__ void __launch_bounds__(256, 2)
    FastColumnBlockKernel(const __grid_constant__ fast::ColPlan p, const __grid_constant__ fast::ColIo io,
                          int tile0) {
  extern __shared__ __align__(16) float2 smem[];
  const int tile = tile0 + blockIdx.x;
  fast::ColumnBlockPhase<0, CONVOLVE>(p, io, smem, tile, blockIdx.y, threadIdx.x, blockDim.x);
  __syncthreads();
  fast::ColumnBlockPhase<1, CONVOLVE>(p, io, smem, tile, blockIdx.y, threadIdx.x, blockDim.x);
  if (!CONVOLVE) return;
  __syncthreads();
  fast::ColumnBlockPhase<2, CONVOLVE>(p, io, smem, tile, blockIdx.y, threadIdx.x, blockDim.x);
}

struct HostPlan {
  Plan plan{};
  DeviceBuffer<float2> tw;
  std::vector<int> radices;

  void Build(int n) {
    if (n < 1) throw std::invalid_argument("FFT length must be positive");
    std::vector<int> generic, pow2, odd;
    int rest = n;
    while (rest % 8 == 0) { pow2.push_back(8); rest /= 8; }
    while (rest % 4 == 0) { pow2.push_back(4); rest /= 4; }
    while (rest % 2 == 0) { pow2.push_back(2); rest /= 2; }
    for (int p : {7, 5, 3})
      while (rest % p == 0) { odd.push_back(p); rest /= p; }
    for (int p = 11; rest > 1; p += 2) {
      if (p * p > rest) p = rest;
      while (rest % p == 0) {
        if (p > kMaxGenericRadix)
          throw std::domain_error("FFT length " + std::to_string(n) + " has the prime factor " +
                                  std::to_string(p) + " > " + std::to_string(kMaxGenericRadix) +
                                  ", which this build does not support");
        generic.push_back(p);
        rest /= p;
      }
    }
    // generic primes first (largest strides), powers of two next, small odd radices last:
    // the late, short-stride stages then have odd strides and stay off the same banks.
    radices = generic;
    radices.insert(radices.end(), pow2.begin(), pow2.end());
    radices.insert(radices.end(), odd.begin(), odd.end());
    if (radices.empty()) radices.push_back(1);
    if (int(radices.size()) > kMaxStages) throw std::domain_error("FFT plan has too many stages");
    plan.n = n;
    plan.n_stages = 0;
    int len = n;
    for (int r : radices) {
      if (r == 1) continue;
      plan.radix[plan.n_stages] = r;
      plan.len[plan.n_stages] = len;
      len /= r;
      ++plan.n_stages;
    }
    std::vector<float2> t(n);
    for (int i = 0; i < n; ++i) {
      const double a = -2.0 * M_PI * double(i) / double(n);
      t[i] = make_float2(float(std::cos(a)), float(std::sin(a)));
    }
    tw.Resize(n);
    RDL_CUDA(cudaMemcpy(tw.get(), t.data(), n * sizeof(float2), cudaMemcpyHostToDevice));
    plan.tw = tw.get();
  }

  // slot of frequency k after the forward (DIF) pass
  uint32_t Slot(uint32_t k) const {
    uint32_t pos = 0;
    int len = plan.n;
    for (int st = 0; st < plan.n_stages; ++st) {
      const int r = plan.radix[st], m = len / r;
      pos += (k % r) * m;
      k /= r;
      len = m;
    }
    return pos;
  }
};

}  // namespace

struct FftConvolver::Impl {
  int W, H, n, pitch;
  int H1 = 1, H2 = 1;
  bool row_generic = false, col_generic = false;  // a prime radix > 7 somewhere in the plans
  HostPlan row, col, col1, col2;
  DeviceBuffer<uint32_t> perm, freq2;
  DeviceBuffer<float2> split_tw;
  DeviceBuffer<float2> work, kernel_spectrum;
  DeviceBuffer<float> kernel_image;
  int rows_per_cta = 1, row_threads = 256, row_stride = 0;
  int threads1 = 256, threads2 = 256;
  size_t row_smem = 0, smem1 = 0, smem2 = 0;
  // register-radix engine (2-3-5-7-smooth frames)
  bool use_fast = false;
  bool use_spec = true;  // radix-specialised instances (fft_spec.cu) where one exists; RDL_FFT_SPEC=0: off
  fast::HostFastPlan fp;
  fast::RowPlan frp{};
  fast::ColPlan fcp{};
  DeviceBuffer<float2> f_tw_row, f_tw_row2, f_split, f_tw_h1, f_tw_h, f_tw_wa, f_tw_wb;
  // whole columns in shared memory (one kernel instead of strided + block + strided) where the
  // plan allows and fft_spec.cu has an instance; RDL_FFT_WHOLE=0: off
  bool use_whole = false;
  fast::WholePlan fwp{};
  DeviceBuffer<uint4> f_pairs;
  int band_tiles = 0;  // column tiles per band of the column passes (L2-resident bands)

  fast::ColIo MakeFastColIo() const {
    fast::ColIo c{};
    c.spec = work.get();
    c.pitch = pitch;
    c.y0 = 0;
    c.y1 = H;
    c.out_y0 = 0;
    c.out_y1 = H;
    c.scale = 1.0f;
    return c;
  }

  ColParams MakeColParams() const {
    ColParams c{};
    c.spec = work.get();
    c.pitch = pitch;
    c.H = H;
    c.H1 = H1;
    c.H2 = H2;
    c.y0 = 0;
    c.y1 = H;
    c.out_y0 = 0;
    c.out_y1 = H;
    c.scale = 1.0f;
    c.plan1 = col1.plan;
    c.plan2 = col2.plan;
    c.tw_h = col.plan.tw;
    c.freq2 = freq2.get();
    return c;
  }
};

namespace {
int TileThreads(int len) {
  const int t = (len * kTileCols / 8 + 31) / 32 * 32;
  return std::clamp(t, 64, 256);
}
}  // namespace

// Embedded frames: K[d mod Q] = k[d mod P] for the displacements |dx| < cw, |dy| < ch, else 0.
__global__ void ExpandKernelKernel(float* dst, int qw, int qh, const float* k, int w, int h, int cw,
                                   int ch) {
  const size_t total = size_t(qw) * qh;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < total;
       i += size_t(gridDim.x) * blockDim.x) {
    const int jy = int(i / qw), jx = int(i - size_t(jy) * qw);
    const int dy = jy < ch ? jy : jy - qh, dx = jx < cw ? jx : jx - qw;
    float v = 0.0f;
    if (dy > -ch && dx > -cw) v = __ldg(k + size_t(dy < 0 ? dy + h : dy) * w + (dx < 0 ? dx + w : dx));
    dst[i] = v;
  }
}

namespace {
bool PlanSupported(size_t n) {
  for (size_t p : {2, 3, 5, 7})
    while (n % p == 0) n /= p;
  for (size_t p = 11; n > 1; p += 2) {
    if (p * p > n) p = n;
    while (n % p == 0) {
      if (p > size_t(kMaxGenericRadix)) return false;
      n /= p;
    }
  }
  return true;
}
}  // namespace

bool FftConvolver::HasNativePlan(size_t width, size_t height) {
  if (width < 2 || height < 1 || width % 2 != 0) return false;
  if (!PlanSupported(width / 2) || !PlanSupported(height)) return false;
  if (size_t(PaddedLength(int(width / 2))) * sizeof(float2) > kMaxSmemBytes) return false;
  if (height > 512) {  // four-step columns: the contiguous factor has to fit in shared memory
    size_t best = 1;
    for (size_t f = 1; f * f <= height; ++f)
      if (height % f == 0) best = f;
    if ((height / best) * (kTileCols * sizeof(float2) + sizeof(float2)) > kMaxSmemBytes) return false;
  }
  return true;
}

FftConvolver::FftConvolver(size_t width, size_t height, size_t content_width, size_t content_height)
    : impl_(new Impl), width_(width), height_(height),
      content_w_(content_width ? std::min(content_width, width) : width),
      content_h_(content_height ? std::min(content_height, height) : height) {
  Impl& d = *impl_;
  try {
    if (width < 1 || height < 1) throw std::invalid_argument("empty FFT frame");
    d.W = int(width);
    d.H = int(height);
    if (!HasNativePlan(width, height)) {
      const size_t qw = CalculateGoodFFTSize(std::max<size_t>(2 * content_w_ - 1, 2));
      const size_t qh = CalculateGoodFFTSize(std::max<size_t>(2 * content_h_ - 1, 2));
      if (!HasNativePlan(qw, qh))
        throw std::domain_error("no FFT plan for a " + std::to_string(width) + " x " +
                                std::to_string(height) + " frame");
      inner_ = new FftConvolver(qw, qh);
      return;
    }
    d.n = d.W / 2;
    // whole 128-byte lines per spectrum row: the column passes work on 16-column tiles
    d.pitch = (d.n + 1 + kTileCols - 1) / kTileCols * kTileCols;
    const char* engine = std::getenv("RDL_FFT_ENGINE");
    if (!engine || std::string(engine) != "classic") d.fp.Build(d.W, d.H);
    if (d.fp.ok) {
      auto upload = [](auto& buf, const auto& host) {
        buf.Resize(host.size());
        RDL_CUDA(cudaMemcpy(buf.get(), host.data(), host.size() * sizeof(host[0]), cudaMemcpyHostToDevice));
      };
      upload(d.f_tw_row, d.fp.tw_row1);
      upload(d.f_tw_row2, d.fp.tw_row2);
      upload(d.f_split, d.fp.tw_split);
      upload(d.f_tw_h1, d.fp.tw_h1);
      upload(d.f_tw_h, d.fp.tw_h);
      upload(d.f_pairs, d.fp.pairs);
      d.frp = d.fp.MakeRowPlan(d.f_tw_row.get(), d.f_tw_row2.get(), d.f_split.get(), d.f_pairs.get());
      d.fcp = d.fp.MakeColPlan(d.f_tw_h1.get(), d.f_tw_h.get());
      const auto attr = cudaFuncAttributeMaxDynamicSharedMemorySize;
      const int smem_max = int(kMaxSmemBytes);
      RDL_CUDA(cudaFuncSetAttribute(FastRowForwardKernel, attr, smem_max));
      RDL_CUDA(cudaFuncSetAttribute(FastRowInverseKernel, attr, smem_max));
      RDL_CUDA(cudaFuncSetAttribute(FastColumnBlockKernel<true>, attr, smem_max));
      RDL_CUDA(cudaFuncSetAttribute(FastColumnBlockKernel<false>, attr, smem_max));
      // The column passes can run band by band so that the three passes of a band find it in L2
      // (RDL_FFT_BAND_MB = spectrum megabytes per band). Measured on B200 (profiles/
      // r02_fft_bench_bands.jsonl): 8192^2 1.02 ms unbanded, 1.09 / 1.16 / 1.27 ms with 70 / 40 /
      // 20 MB bands -- the extra launches and tails cost more than the L2 hits save, so the default
      // is one band.
      double band_mb = 0.0;
      if (const char* e = std::getenv("RDL_FFT_BAND_MB")) band_mb = std::atof(e);
      const int tiles = d.pitch / kTileCols;
      const double tile_mb = double(d.H) * kTileCols * sizeof(float2) / 1.0e6;
      d.band_tiles = tiles;
      if (band_mb > 0.0 && d.fcp.H2 > 1 && tile_mb * tiles > 1.5 * band_mb)
        d.band_tiles = std::max(1, int(band_mb / tile_mb));
      d.use_fast = true;
      if (const char* e = std::getenv("RDL_FFT_SPEC")) d.use_spec = std::atoi(e) != 0;
      // off unless asked for: measured SLOWER than strided + block + strided on B200 (8192^2 1.00 vs
      // 0.79 ms, 9216^2 1.31 vs 0.90, 4096^2 0.224 vs 0.182; profiles/r02_fft_bench_whole_column.jsonl):
      // one CTA per SM with load, five shared-memory passes and store in series loses more than
      // the 16 B/px it saves
      bool want_whole = false;
      if (const char* e = std::getenv("RDL_FFT_WHOLE")) want_whole = d.use_spec && std::atoi(e) != 0;
      if (want_whole && d.fp.whole_ok &&
          fast::HasWholeColumnSpec(d.fp.whole.ra, d.fp.whole.rb, d.fp.whole.rc)) {
        upload(d.f_tw_wa, d.fp.tw_whole_a);
        upload(d.f_tw_wb, d.fp.tw_whole_b);
        d.fwp = d.fp.MakeWholePlan(d.f_tw_wa.get(), d.f_tw_wb.get());
        d.use_whole = true;
      }
      d.work.Resize(size_t(d.H) * d.pitch);
      RDL_CUDA(cudaMemset(d.work.get(), 0, size_t(d.H) * d.pitch * sizeof(float2)));
      return;
    }
    d.row.Build(d.n);
    d.col.Build(d.H);  // validates the factors and holds the exp(-2 pi i t / H) table
    std::vector<uint32_t> perm(d.n);
    for (int k = 0; k < d.n; ++k) perm[k] = d.row.Slot(uint32_t(k));
    d.perm.Resize(d.n);
    RDL_CUDA(cudaMemcpy(d.perm.get(), perm.data(), d.n * sizeof(uint32_t), cudaMemcpyHostToDevice));
    std::vector<float2> st(d.n + 1);
    for (int k = 0; k <= d.n; ++k) {
      const double a = -2.0 * M_PI * double(k) / double(d.W);
      st[k] = make_float2(float(std::cos(a)), float(std::sin(a)));
    }
    d.split_tw.Resize(d.n + 1);
    RDL_CUDA(cudaMemcpy(d.split_tw.get(), st.data(), (d.n + 1) * sizeof(float2), cudaMemcpyHostToDevice));

    d.row_stride = PaddedLength(d.n);
    const size_t row_bytes = size_t(d.row_stride) * sizeof(float2);
    if (row_bytes > kMaxSmemBytes) throw std::domain_error("image row does not fit in shared memory");
    size_t row_budget_kb = 75;
    const char* budget_override = std::getenv("RDL_FFT_ROW_SMEM_KB");
    if (budget_override) row_budget_kb = size_t(std::atoi(budget_override));
    d.rows_per_cta = int(std::clamp<size_t>(row_budget_kb * 1024 / row_bytes, 1, 8));
    // long rows: two per CTA (and two CTAs per SM) still beat one (9408: 2.23 -> 2.08 ms)
    if (!budget_override && d.rows_per_cta == 1 && 2 * row_bytes <= 113 * 1024) d.rows_per_cta = 2;
    d.row_smem = row_bytes * d.rows_per_cta;
    d.row_threads = d.n * d.rows_per_cta >= 4096 ? 256 : 128;

    // Column split H = H1 * H2 (four-step). Short columns are done in one pass (H2 = 1);
    // otherwise the most balanced factorisation, the larger factor on the contiguous side.
    const size_t tile_row_bytes = kTileCols * sizeof(float2);
    const int single_pass_max = 512;
    d.H1 = d.H;
    d.H2 = 1;
    if (d.H > single_pass_max) {
      int best = 0;
      for (int f = 1; f * f <= d.H; ++f)
        if (d.H % f == 0) best = f;
      d.H2 = best;
      d.H1 = d.H / best;
    }
    if (size_t(d.H1) * (tile_row_bytes + sizeof(float2)) > kMaxSmemBytes)
      throw std::domain_error("image height " + std::to_string(d.H) +
                              " has no factorisation that fits in shared memory");
    d.col1.Build(d.H1);
    d.col2.Build(d.H2);
    std::vector<uint32_t> freq2(d.H2);
    for (int k = 0; k < d.H2; ++k) freq2[d.col2.Slot(uint32_t(k))] = uint32_t(k);
    d.freq2.Resize(d.H2);
    RDL_CUDA(cudaMemcpy(d.freq2.get(), freq2.data(), d.H2 * sizeof(uint32_t), cudaMemcpyHostToDevice));
    // tile + the transform's twiddle table
    d.smem1 = size_t(d.H1) * (tile_row_bytes + sizeof(float2));
    d.smem2 = size_t(d.H2) * (tile_row_bytes + sizeof(float2));
    d.threads1 = std::min(2 * TileThreads(d.H1), 512);
    d.threads2 = TileThreads(d.H2);

    auto has_generic = [](const HostPlan& hp) {
      return std::any_of(hp.radices.begin(), hp.radices.end(), [](int r) { return r > 8; });
    };
    d.row_generic = has_generic(d.row);
    d.col_generic = has_generic(d.col1) || has_generic(d.col2);
    const int smem_max = int(kMaxSmemBytes);
    const auto attr = cudaFuncAttributeMaxDynamicSharedMemorySize;
    RDL_CUDA(cudaFuncSetAttribute(RowForwardKernel<false>, attr, smem_max));
    RDL_CUDA(cudaFuncSetAttribute(RowForwardKernel<true>, attr, smem_max));
    RDL_CUDA(cudaFuncSetAttribute(RowInverseKernel<false>, attr, smem_max));
    RDL_CUDA(cudaFuncSetAttribute(RowInverseKernel<true>, attr, smem_max));
    RDL_CUDA(cudaFuncSetAttribute(ColumnBlockKernel<true, kTileCols, false>, attr, smem_max));
    RDL_CUDA(cudaFuncSetAttribute(ColumnBlockKernel<true, kTileCols, true>, attr, smem_max));
    RDL_CUDA(cudaFuncSetAttribute(ColumnBlockKernel<false, kTileCols, false>, attr, smem_max));
    RDL_CUDA(cudaFuncSetAttribute(ColumnBlockKernel<false, kTileCols, true>, attr, smem_max));
    RDL_CUDA(cudaFuncSetAttribute(ColumnStridedKernel<true, kTileCols, false>, attr, smem_max));
    RDL_CUDA(cudaFuncSetAttribute(ColumnStridedKernel<true, kTileCols, true>, attr, smem_max));
    RDL_CUDA(cudaFuncSetAttribute(ColumnStridedKernel<false, kTileCols, false>, attr, smem_max));
    RDL_CUDA(cudaFuncSetAttribute(ColumnStridedKernel<false, kTileCols, true>, attr, smem_max));
    d.work.Resize(size_t(d.H) * d.pitch);
    // columns n+1 .. pitch-1 ride along in the column tiles: keep them finite
    RDL_CUDA(cudaMemset(d.work.get(), 0, size_t(d.H) * d.pitch * sizeof(float2)));
  } catch (...) {
    delete impl_;
    impl_ = nullptr;
    delete inner_;
    inner_ = nullptr;
    throw;
  }
}

FftConvolver::~FftConvolver() {
  delete impl_;
  delete inner_;
}

size_t FftConvolver::SpectrumSize() const {
  return inner_ ? inner_->SpectrumSize() : size_t(impl_->H) * impl_->pitch;
}

// embedded frames: impl_->kernel_image (W x H, origin at 0,0) -> spectrum of the inner frame
void FftConvolver::ExpandKernel(float2* spectrum_out, cudaStream_t stream) {
  Impl& d = *impl_;
  Impl& q = *inner_->impl_;
  q.kernel_image.Reserve(size_t(q.W) * q.H);
  ExpandKernelKernel<<<1184, 256, 0, stream>>>(q.kernel_image.get(), q.W, q.H, d.kernel_image.get(),
                                               d.W, d.H, int(content_w_), int(content_h_));
  ++launches_;
  inner_->MakeKernelSpectrum(q.kernel_image.get(), spectrum_out, stream);
}

void FftConvolver::ForwardRows(const float* src, size_t sw, size_t sh, size_t ox, size_t oy,
                               cudaStream_t stream) {
  Impl& d = *impl_;
  if (ox + sw > size_t(d.W) || oy + sh > size_t(d.H))
    throw std::invalid_argument("source larger than the FFT frame");
  if (d.use_fast) {
    fast::RowIo io{};
    io.src = src;
    io.sw = int(sw);
    io.sh = int(sh);
    io.ox = int(ox);
    io.oy = int(oy);
    io.spec = d.work.get();
    io.pitch = d.pitch;
    const int grid = (int(sh) + d.fp.rows_per_cta - 1) / d.fp.rows_per_cta;
    if (!d.use_spec ||
        !fast::LaunchRowForwardSpec(d.frp, io, grid, d.fp.row_threads, d.fp.row_smem, stream))
      FastRowForwardKernel<<<grid, d.fp.row_threads, d.fp.row_smem, stream>>>(d.frp, io);
    ++launches_;
    return;
  }
  RowParams p{};
  p.src = src;
  p.sw = int(sw);
  p.sh = int(sh);
  p.ox = int(ox);
  p.oy = int(oy);
  p.W = d.W;
  p.H = d.H;
  p.spec = d.work.get();
  p.pitch = d.pitch;
  p.plan = d.row.plan;
  p.perm = d.perm.get();
  p.split_tw = d.split_tw.get();
  p.rows_per_cta = d.rows_per_cta;
  p.seq_stride = d.row_stride;
  const int grid = (int(sh) + d.rows_per_cta - 1) / d.rows_per_cta;
  if (d.row_generic)
    RowForwardKernel<true><<<grid, d.row_threads, d.row_smem, stream>>>(p);
  else
    RowForwardKernel<false><<<grid, d.row_threads, d.row_smem, stream>>>(p);
  ++launches_;
}

void FftConvolver::MakeKernelSpectrum(const float* kernel_image, float2* spectrum_out, cudaStream_t stream) {
  Impl& d = *impl_;
  if (inner_) {
    // the caller's W x H kernel image (origin at 0,0) written out over the displacements
    if (kernel_image != d.kernel_image.get()) {
      d.kernel_image.Reserve(size_t(d.W) * d.H);
      RDL_CUDA(cudaMemcpyAsync(d.kernel_image.get(), kernel_image, size_t(d.W) * d.H * sizeof(float),
                               cudaMemcpyDeviceToDevice, stream));
    }
    ExpandKernel(spectrum_out, stream);
    return;
  }
  ForwardRows(kernel_image, d.W, d.H, 0, 0, stream);
  if (d.use_fast) {
    fast::ColIo io = d.MakeFastColIo();
    io.kspec_out = spectrum_out;
    io.scale = float(1.0 / (4.0 * double(d.W) * double(d.H)));
    if (d.use_whole) {
      fast::LaunchWholeColumnSpec(false, d.fwp, io, d.pitch / d.fwp.tw, d.fp.whole_threads,
                                  d.fp.whole_smem, stream);
      ++launches_;
      return;
    }
    const int tiles = d.pitch / kTileCols;
    const int groups = (d.fcp.H1 + fast::kStridedRows - 1) / fast::kStridedRows;
    for (int t0 = 0; t0 < tiles; t0 += d.band_tiles) {
      const int nt = std::min(d.band_tiles, tiles - t0);
      if (d.fcp.H2 > 1) {
        if (!d.use_spec ||
            !fast::LaunchColumnStridedSpec(false, d.fcp, io, t0, dim3(nt, groups),
                                           fast::kTile * fast::kStridedRows, stream))
          FastColumnStridedKernel<false><<<dim3(nt, groups), fast::kTile * fast::kStridedRows, 0, stream>>>(
              d.fcp, io, t0);
        ++launches_;
      }
      if (!d.use_spec || !fast::LaunchColumnBlockSpec(false, d.fcp, io, t0, dim3(nt, d.fcp.H2),
                                                      d.fp.block_threads, d.fp.block_smem, stream))
        FastColumnBlockKernel<false><<<dim3(nt, d.fcp.H2), d.fp.block_threads, d.fp.block_smem, stream>>>(
            d.fcp, io, t0);
      ++launches_;
    }
    return;
  }
  ColParams c = d.MakeColParams();
  c.kspec_out = spectrum_out;
  c.scale = float(1.0 / (4.0 * double(d.W) * double(d.H)));
  const int tiles = d.pitch / kTileCols;
  if (d.H2 > 1) {
    if (d.col_generic)
      ColumnStridedKernel<false, kTileCols, true><<<dim3(tiles, d.H1), d.threads2, d.smem2, stream>>>(c);
    else
      ColumnStridedKernel<false, kTileCols, false><<<dim3(tiles, d.H1), d.threads2, d.smem2, stream>>>(c);
    ++launches_;
  }
  if (d.col_generic)
    ColumnBlockKernel<false, kTileCols, true><<<dim3(tiles, d.H2), d.threads1, d.smem1, stream>>>(c);
  else
    ColumnBlockKernel<false, kTileCols, false><<<dim3(tiles, d.H2), d.threads1, d.smem1, stream>>>(c);
  ++launches_;
}

void FftConvolver::MakeSmallKernelSpectrum(const float* small_kernel, size_t n, float2* spectrum_out,
                                           cudaStream_t stream) {
  Impl& d = *impl_;
  if (n > size_t(d.W) || n > size_t(d.H)) throw std::invalid_argument("convolution kernel larger than the image");
  d.kernel_image.Reserve(size_t(d.W) * d.H);
  RDL_CUDA(cudaMemsetAsync(d.kernel_image.get(), 0, size_t(d.W) * d.H * sizeof(float), stream));
  const int total = int(n * n);
  WrapSmallKernelKernel<<<std::min((total + 255) / 256, 1184), 256, 0, stream>>>(
      d.kernel_image.get(), d.W, d.H, small_kernel, int(n));
  ++launches_;
  if (inner_) return ExpandKernel(spectrum_out, stream);
  MakeKernelSpectrum(d.kernel_image.get(), spectrum_out, stream);
}

void FftConvolver::MakePsfKernelSpectrum(const float* psf, size_t pw, size_t ph, float2* spectrum_out,
                                         cudaStream_t stream) {
  Impl& d = *impl_;
  if (pw > size_t(d.W) || ph > size_t(d.H)) throw std::invalid_argument("PSF larger than the FFT frame");
  d.kernel_image.Reserve(size_t(d.W) * d.H);
  PreparePsfKernelKernel<<<1184, 256, 0, stream>>>(d.kernel_image.get(), d.W, d.H, psf, int(pw), int(ph));
  ++launches_;
  if (inner_) return ExpandKernel(spectrum_out, stream);
  MakeKernelSpectrum(d.kernel_image.get(), spectrum_out, stream);
}

void FftConvolver::ConvolveWith(const float* src, size_t sw, size_t sh, const float2* kernel_spectrum,
                                float* dst, size_t dw, size_t dh, int mode, cudaStream_t stream) {
  if (sw > width_ || sh > height_) throw std::invalid_argument("source larger than the FFT frame");
  if (dw > width_ || dh > height_) throw std::invalid_argument("destination larger than the FFT frame");
  // Image::Untrim / Trim: centred, start = (outer - inner) / 2
  ConvolveAt(src, sw, sh, (width_ - sw) / 2, (height_ - sh) / 2, kernel_spectrum, dst, dw, dh,
             (width_ - dw) / 2, (height_ - dh) / 2, mode, stream);
}

void FftConvolver::ConvolveAt(const float* src, size_t sw, size_t sh, size_t sox, size_t soy,
                              const float2* kernel_spectrum, float* dst, size_t dw, size_t dh,
                              size_t dox, size_t doy, int mode, cudaStream_t stream) {
  if (inner_) {
    // only the relative position matters for a circular convolution: shift both to the corner
    if (sw > content_w_ || dw > content_w_ || sh > content_h_ || dh > content_h_)
      throw std::invalid_argument("image larger than the content size of an embedded FFT frame");
    const size_t bx = std::min(sox, dox), by = std::min(soy, doy);
    if (sox - bx + sw > content_w_ || dox - bx + dw > content_w_ || soy - by + sh > content_h_ ||
        doy - by + dh > content_h_)
      throw std::invalid_argument("source and destination are further apart than the content size");
    inner_->ConvolveAt(src, sw, sh, sox - bx, soy - by, kernel_spectrum, dst, dw, dh, dox - bx,
                       doy - by, mode, stream);
    return;
  }
  Impl& d = *impl_;
  if (dox + dw > size_t(d.W) || doy + dh > size_t(d.H))
    throw std::invalid_argument("destination larger than the FFT frame");
  ForwardRows(src, sw, sh, sox, soy, stream);
  if (d.use_fast) {
    fast::ColIo io = d.MakeFastColIo();
    io.kspec = kernel_spectrum;
    io.y0 = int(soy);
    io.y1 = io.y0 + int(sh);
    io.out_y0 = int(doy);
    io.out_y1 = io.out_y0 + int(dh);
    const int tiles = d.pitch / kTileCols;
    const int groups = (d.fcp.H1 + fast::kStridedRows - 1) / fast::kStridedRows;
    const int strided_threads = fast::kTile * fast::kStridedRows;
    for (int t0 = 0; t0 < tiles && !d.use_whole; t0 += d.band_tiles) {
      const int nt = std::min(d.band_tiles, tiles - t0);
      if (d.fcp.H2 > 1) {
        if (!d.use_spec || !fast::LaunchColumnStridedSpec(false, d.fcp, io, t0, dim3(nt, groups),
                                                          strided_threads, stream))
          FastColumnStridedKernel<false><<<dim3(nt, groups), strided_threads, 0, stream>>>(d.fcp, io, t0);
        ++launches_;
      }
      if (!d.use_spec || !fast::LaunchColumnBlockSpec(true, d.fcp, io, t0, dim3(nt, d.fcp.H2),
                                                      d.fp.block_threads, d.fp.block_smem, stream))
        FastColumnBlockKernel<true><<<dim3(nt, d.fcp.H2), d.fp.block_threads, d.fp.block_smem, stream>>>(
            d.fcp, io, t0);
      ++launches_;
      if (d.fcp.H2 > 1) {
        if (!d.use_spec || !fast::LaunchColumnStridedSpec(true, d.fcp, io, t0, dim3(nt, groups),
                                                          strided_threads, stream))
          FastColumnStridedKernel<true><<<dim3(nt, groups), strided_threads, 0, stream>>>(d.fcp, io, t0);
        ++launches_;
      }
    }
    if (d.use_whole) {
      fast::LaunchWholeColumnSpec(true, d.fwp, io, d.pitch / d.fwp.tw, d.fp.whole_threads,
                                  d.fp.whole_smem, stream);
      ++launches_;
    }
    fast::RowIo rio{};
    rio.dst = dst;
    rio.sw = int(dw);
    rio.sh = int(dh);
    rio.ox = int(dox);
    rio.oy = int(doy);
    rio.spec = d.work.get();
    rio.pitch = d.pitch;
    rio.mode = mode;
    const int grid = (int(dh) + d.fp.rows_per_cta - 1) / d.fp.rows_per_cta;
    if (!d.use_spec ||
        !fast::LaunchRowInverseSpec(d.frp, rio, grid, d.fp.row_threads, d.fp.row_smem, stream))
      FastRowInverseKernel<<<grid, d.fp.row_threads, d.fp.row_smem, stream>>>(d.frp, rio);
    ++launches_;
    return;
  }
  ColParams c = d.MakeColParams();
  c.kspec = kernel_spectrum;
  c.y0 = int(soy);
  c.y1 = c.y0 + int(sh);
  c.out_y0 = int(doy);
  c.out_y1 = c.out_y0 + int(dh);
  const int tiles = d.pitch / kTileCols;
  if (d.H2 > 1) {
    if (d.col_generic)
      ColumnStridedKernel<false, kTileCols, true><<<dim3(tiles, d.H1), d.threads2, d.smem2, stream>>>(c);
    else
      ColumnStridedKernel<false, kTileCols, false><<<dim3(tiles, d.H1), d.threads2, d.smem2, stream>>>(c);
    ++launches_;
  }
  if (d.col_generic)
    ColumnBlockKernel<true, kTileCols, true><<<dim3(tiles, d.H2), d.threads1, d.smem1, stream>>>(c);
  else
    ColumnBlockKernel<true, kTileCols, false><<<dim3(tiles, d.H2), d.threads1, d.smem1, stream>>>(c);
  ++launches_;
  if (d.H2 > 1) {
    if (d.col_generic)
      ColumnStridedKernel<true, kTileCols, true><<<dim3(tiles, d.H1), d.threads2, d.smem2, stream>>>(c);
    else
      ColumnStridedKernel<true, kTileCols, false><<<dim3(tiles, d.H1), d.threads2, d.smem2, stream>>>(c);
    ++launches_;
  }
  RowParams p{};
  p.dst = dst;
  p.sw = int(dw);
  p.sh = int(dh);
  p.ox = int(dox);
  p.oy = c.out_y0;
  p.W = d.W;
  p.H = d.H;
  p.spec = d.work.get();
  p.pitch = d.pitch;
  p.plan = d.row.plan;
  p.perm = d.perm.get();
  p.split_tw = d.split_tw.get();
  p.rows_per_cta = d.rows_per_cta;
  p.seq_stride = d.row_stride;
  p.mode = mode;
  const int grid = (int(dh) + d.rows_per_cta - 1) / d.rows_per_cta;
  if (d.row_generic)
    RowInverseKernel<true><<<grid, d.row_threads, d.row_smem, stream>>>(p);
  else
    RowInverseKernel<false><<<grid, d.row_threads, d.row_smem, stream>>>(p);
  ++launches_;
}

void FftConvolver::SetKernel(const float* kernel, cudaStream_t stream) {
  Impl& d = *impl_;
  d.kernel_spectrum.Reserve(SpectrumSize());
  MakeKernelSpectrum(kernel, d.kernel_spectrum.get(), stream);
}

void FftConvolver::Convolve(const float* in, float* out, cudaStream_t stream) {
  Impl& d = *impl_;
  if (d.kernel_spectrum.empty()) throw std::invalid_argument("FftConvolver::Convolve without a kernel");
  ConvolveWith(in, width_, height_, d.kernel_spectrum.get(), out, width_, height_, 0, stream);
}

FftConvolver& GetConvolver(rdl_handle& h, size_t width, size_t height, size_t content_width,
                           size_t content_height) {
  const std::array<size_t, 4> key{width, height, content_width ? content_width : width,
                                  content_height ? content_height : height};
  auto it = h.convolvers.find(key);
  if (it == h.convolvers.end())
    it = h.convolvers
             .emplace(key, std::make_shared<FftConvolver>(width, height, key[2], key[3]))
             .first;
  return *it->second;
}

}  // namespace rdl
