// Register-radix FFT passes for the convolution engine (fft.cu). Every pass keeps one whole
// radix-R butterfly (R up to 32, composite radices built at compile time from 2/3/4/5/7/8-point
// kernels with constant twiddles) in registers, so a 4096..4704-point row transform is THREE
// passes with two shared-memory exchanges instead of five or six radix <= 8 passes, the first pass
// reads global memory directly, and the last forward / first inverse pass carries the real-FFT
// split, computed on pairs of butterflies (k, n - k) that one thread holds in registers: no
// permutation tables, no digit-reversed gathers.
//
// The pass bodies are __host__ __device__ functions of (thread index, thread count) so that
// tests/fft_emulation.cu can run them on the CPU, phase by phase, against a float64 DFT.
#ifndef RDL_FFT_FAST_CUH_
#define RDL_FFT_FAST_CUH_

#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>
#include <type_traits>
#include <utility>

#if defined(__CUDA_ARCH__)
#define RDL_LDG(p) __ldg(p)
#else
#define RDL_LDG(p) (*(p))
#endif
#define RDL_HD __host__ __device__ __forceinline__

namespace rdl {
namespace fast {

// ---- compile-time loops and trigonometry -----------------------------------------------------
template <typename F, int... I>
RDL_HD void StaticForImpl(F&& f, std::integer_sequence<int, I...>) {
  (f(std::integral_constant<int, I>{}), ...);
}
template <int N, typename F>
RDL_HD void StaticFor(F&& f) {
  StaticForImpl(f, std::make_integer_sequence<int, N>{});
}

constexpr double kTau = 6.283185307179586476925286766559;
constexpr double SinSeries(double x) {  // |x| <= pi/4
  const double x2 = x * x;
  double term = x, sum = x;
  for (int k = 1; k < 14; ++k) {
    term *= -x2 / double((2 * k) * (2 * k + 1));
    sum += term;
  }
  return sum;
}
constexpr double CosSeries(double x) {  // |x| <= pi/4
  const double x2 = x * x;
  double term = 1.0, sum = 1.0;
  for (int k = 1; k < 14; ++k) {
    term *= -x2 / double((2 * k - 1) * (2 * k));
    sum += term;
  }
  return sum;
}
constexpr double SinTurn(double f);
constexpr double CosTurn(double f) {  // cos(2 pi f), f in [0, 1)
  if (f > 0.5) f = 1.0 - f;
  if (f > 0.25) return -CosTurn(0.5 - f);
  return f > 0.125 ? SinSeries(kTau * (0.25 - f)) : CosSeries(kTau * f);
}
constexpr double SinTurn(double f) {  // sin(2 pi f), f in [0, 1)
  if (f > 0.5) return -SinTurn(1.0 - f);
  if (f > 0.25) f = 0.5 - f;
  return f > 0.125 ? CosSeries(kTau * (0.25 - f)) : SinSeries(kTau * f);
}

RDL_HD float2 CAdd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
RDL_HD float2 CSub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }
RDL_HD float2 CScale(float s, float2 a) { return make_float2(s * a.x, s * a.y); }
RDL_HD float2 CMulF(float2 a, float2 b) {
  return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x);
}
RDL_HD float2 CMulConj(float2 a, float2 w) {  // a * conj(w)
  return make_float2(a.x * w.x + a.y * w.y, a.y * w.x - a.x * w.y);
}

// a * exp(-/+ 2 pi i NUM / DEN) with the constant folded at compile time
template <int NUM, int DEN, bool INV>
RDL_HD float2 MulW(float2 a) {
  constexpr int num = ((NUM % DEN) + DEN) % DEN;
  if constexpr (num == 0) {
    return a;
  } else if constexpr (2 * num == DEN) {
    return make_float2(-a.x, -a.y);
  } else if constexpr (4 * num == DEN) {  // forward: * (-i)
    return INV ? make_float2(-a.y, a.x) : make_float2(a.y, -a.x);
  } else if constexpr (4 * num == 3 * DEN) {  // forward: * (+i)
    return INV ? make_float2(a.y, -a.x) : make_float2(-a.y, a.x);
  } else {
    constexpr float c = float(CosTurn(double(num) / double(DEN)));
    constexpr float s = float(SinTurn(double(num) / double(DEN)));
    // forward w = c - i s, inverse w = c + i s
    return INV ? make_float2(a.x * c - a.y * s, a.y * c + a.x * s)
               : make_float2(a.x * c + a.y * s, a.y * c - a.x * s);
  }
}

// ---- register DFTs, natural order in and out --------------------------------------------------
template <int R, bool INV>
struct Dft;

template <bool INV>
struct Dft<1, INV> {
  RDL_HD static void Run(float2 (&)[1]) {}
};
template <bool INV>
struct Dft<2, INV> {
  RDL_HD static void Run(float2 (&v)[2]) {
    const float2 t = v[0];
    v[0] = CAdd(t, v[1]);
    v[1] = CSub(t, v[1]);
  }
};
template <bool INV>
struct Dft<4, INV> {
  RDL_HD static void Run(float2 (&v)[4]) {
    const float2 t0 = CAdd(v[0], v[2]), t1 = CSub(v[0], v[2]), t2 = CAdd(v[1], v[3]);
    const float2 t3 = MulW<1, 4, INV>(CSub(v[1], v[3]));
    v[0] = CAdd(t0, t2);
    v[1] = CAdd(t1, t3);
    v[2] = CSub(t0, t2);
    v[3] = CSub(t1, t3);
  }
};

// odd prime P by the symmetric (cos / sin) form: (P-1)^2 / 2 real multiplications
template <int P, bool INV>
struct DftOdd {
  RDL_HD static void Run(float2 (&v)[P]) {
    constexpr int H = (P - 1) / 2;
    float2 a[H], b[H];
    StaticFor<H>([&](auto m) {
      a[m] = CAdd(v[m + 1], v[P - 1 - m]);
      b[m] = CSub(v[m + 1], v[P - 1 - m]);
    });
    const float2 x0 = v[0];
    float2 sum = x0;
    StaticFor<H>([&](auto m) { sum = CAdd(sum, a[m]); });
    StaticFor<H>([&](auto qq) {
      constexpr int q = decltype(qq)::value + 1;
      float2 c = x0, s = make_float2(0.0f, 0.0f);
      StaticFor<H>([&](auto mm) {
        constexpr int m = decltype(mm)::value + 1;
        constexpr int t = (q * m) % P;
        constexpr float ct = float(CosTurn(double(t) / double(P)));
        constexpr float st = float(SinTurn(double(t) / double(P)));
        c = CAdd(c, CScale(ct, a[m - 1]));
        s = CAdd(s, CScale(st, b[m - 1]));
      });
      // forward: X_q = c - i s, X_{P-q} = c + i s; inverse: swapped
      const float2 lo = make_float2(c.x + s.y, c.y - s.x), hi = make_float2(c.x - s.y, c.y + s.x);
      v[q] = INV ? hi : lo;
      v[P - q] = INV ? lo : hi;
    });
    v[0] = sum;
  }
};
template <bool INV>
struct Dft<3, INV> : DftOdd<3, INV> {};
template <bool INV>
struct Dft<5, INV> : DftOdd<5, INV> {};
template <bool INV>
struct Dft<7, INV> : DftOdd<7, INV> {};

// R = A * B, Cooley-Tukey in registers: B-point transforms of the A interleaved subsequences,
// constant twiddles, A-point transforms across them.
template <int A, int B, bool INV>
struct DftCT {
  static constexpr int R = A * B;
  RDL_HD static void Run(float2 (&v)[R]) {
    float2 y[R];
    StaticFor<A>([&](auto aa) {
      constexpr int a = decltype(aa)::value;
      float2 u[B];
      StaticFor<B>([&](auto b) { u[b] = v[a + A * b]; });
      Dft<B, INV>::Run(u);
      StaticFor<B>([&](auto kk) {
        constexpr int kb = decltype(kk)::value;
        y[a * B + kb] = MulW<a * kb, R, INV>(u[kb]);
      });
    });
    StaticFor<B>([&](auto kk) {
      constexpr int kb = decltype(kk)::value;
      float2 t[A];
      StaticFor<A>([&](auto a) { t[a] = y[a * B + kb]; });
      Dft<A, INV>::Run(t);
      StaticFor<A>([&](auto ka) { v[kb + B * ka] = t[ka]; });
    });
  }
};
#define RDL_DFT_CT(R, A, B) \
  template <bool INV>       \
  struct Dft<R, INV> : DftCT<A, B, INV> {};
RDL_DFT_CT(6, 2, 3)
RDL_DFT_CT(8, 2, 4)
RDL_DFT_CT(9, 3, 3)
RDL_DFT_CT(10, 2, 5)
RDL_DFT_CT(12, 3, 4)
RDL_DFT_CT(14, 2, 7)
RDL_DFT_CT(15, 3, 5)
RDL_DFT_CT(16, 4, 4)
RDL_DFT_CT(18, 2, 9)
RDL_DFT_CT(20, 4, 5)
RDL_DFT_CT(21, 3, 7)
RDL_DFT_CT(24, 3, 8)
RDL_DFT_CT(25, 5, 5)
RDL_DFT_CT(27, 3, 9)
RDL_DFT_CT(28, 4, 7)
RDL_DFT_CT(30, 5, 6)
RDL_DFT_CT(32, 4, 8)
#undef RDL_DFT_CT

// Radices a pass may use, and the dispatcher from a runtime radix to the compiled body.
constexpr int kRadices[] = {1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 12, 14, 15, 16, 18, 20, 21, 24, 25, 27, 28, 30, 32};
constexpr int kMaxLastRadix = 18;  // the split pass holds TWO butterflies in registers

#define RDL_RADIX_CASE(R, CALL) \
  case R: {                     \
    constexpr int kR = R;       \
    CALL;                       \
  } break;
#define RDL_DISPATCH_RADIX(r, CALL)                                                              \
  switch (r) {                                                                                   \
    RDL_RADIX_CASE(1, CALL) RDL_RADIX_CASE(2, CALL) RDL_RADIX_CASE(3, CALL)                      \
    RDL_RADIX_CASE(4, CALL) RDL_RADIX_CASE(5, CALL) RDL_RADIX_CASE(6, CALL)                      \
    RDL_RADIX_CASE(7, CALL) RDL_RADIX_CASE(8, CALL) RDL_RADIX_CASE(9, CALL)                      \
    RDL_RADIX_CASE(10, CALL) RDL_RADIX_CASE(12, CALL) RDL_RADIX_CASE(14, CALL)                   \
    RDL_RADIX_CASE(15, CALL) RDL_RADIX_CASE(16, CALL) RDL_RADIX_CASE(18, CALL)                   \
    RDL_RADIX_CASE(20, CALL) RDL_RADIX_CASE(21, CALL) RDL_RADIX_CASE(24, CALL)                   \
    RDL_RADIX_CASE(25, CALL) RDL_RADIX_CASE(27, CALL) RDL_RADIX_CASE(28, CALL)                   \
    RDL_RADIX_CASE(30, CALL) RDL_RADIX_CASE(32, CALL)                                            \
    default: break;                                                                              \
  }
#define RDL_DISPATCH_LAST_RADIX(r, CALL)                                                         \
  switch (r) {                                                                                   \
    RDL_RADIX_CASE(1, CALL) RDL_RADIX_CASE(2, CALL) RDL_RADIX_CASE(3, CALL)                      \
    RDL_RADIX_CASE(4, CALL) RDL_RADIX_CASE(5, CALL) RDL_RADIX_CASE(6, CALL)                      \
    RDL_RADIX_CASE(7, CALL) RDL_RADIX_CASE(8, CALL) RDL_RADIX_CASE(9, CALL)                      \
    RDL_RADIX_CASE(10, CALL) RDL_RADIX_CASE(12, CALL) RDL_RADIX_CASE(14, CALL)                   \
    RDL_RADIX_CASE(15, CALL) RDL_RADIX_CASE(16, CALL) RDL_RADIX_CASE(18, CALL)                   \
    default: break;                                                                              \
  }

// ---- rows: real <-> half spectrum through an n = W / 2 point complex transform ------------------
// n = ra * rb * rc. In-place decimation in frequency: after pass 1 the outputs of butterfly j sit
// in block ka (length l2 = rb * rc), after pass 2 in sub-block kb (length rc); frequency
// k = ka + ra * kb + ra * rb * kc ends at position (ka * rb + kb) * rc + kc. One padding slot per
// l2 block makes the block stride odd, so the last pass (lanes along ka) is conflict-free.
struct RowPlan {
  int n, ra, rb, rc, l2;
  int pitch_l2;  // l2 + pad
  int n_pairs;
  int rows_per_cta, row_stride;  // shared-memory elements per row
  // twiddles by (output index q, butterfly j) so that neighbouring threads (j) read neighbouring
  // entries: tw1[(q - 1) * l2 + j] = exp(-2 pi i j q / n), tw2[(q - 1) * rc + j] = exp(-2 pi i j q / l2)
  const float2* tw1;
  const float2* tw2;
  const float2* split_tw;  // exp(-2 pi i k / (2 n)), k in [0, n]
  const uint4* pairs;      // last pass: (offset A, offset B, k_low A, k_low B); entry 0 = block 0
};

struct RowIo {
  const float* src;  // forward: source image (sw x sh)
  float* dst;        // inverse: destination image (sw x sh)
  int sw, sh;
  int ox, oy;  // position of the image inside the frame
  float2* spec;  // [H][pitch] half spectrum
  int pitch;
  int mode;  // inverse epilogue: 0 store, 1 dst -= v, 2 dst += v
};

// 2 X[k] = (Z[k] + conj Z[n-k]) - i w^k (Z[k] - conj Z[n-k])
RDL_HD float2 SplitForward(float2 z, float2 zp, float2 w) {
  const float2 sum = make_float2(z.x + zp.x, z.y - zp.y);
  const float2 dif = make_float2(z.x - zp.x, z.y + zp.y);
  const float2 t = CMulF(w, dif);
  return make_float2(sum.x + t.y, sum.y - t.x);
}
// 2 Z[k] = (Y[k] + conj Y[n-k]) + i conj(w^k) (Y[k] - conj Y[n-k])
RDL_HD float2 SplitInverse(float2 y, float2 yp, float2 w) {
  const float2 sum = make_float2(y.x + yp.x, y.y - yp.y);
  const float2 dif = make_float2(y.x - yp.x, y.y + yp.y);
  const float2 t = CMulConj(dif, w);
  return make_float2(sum.x - t.y, sum.y + t.x);
}

// complex element i of a frame row = pixels (2 i, 2 i + 1); outside the image: 0 (Untrim)
RDL_HD float2 LoadPixelPair(const float* row, int i, int ox, int sw, bool fast) {
  const int xs = 2 * i - ox;
  if (fast) {
    return (xs >= 0 && xs < sw) ? RDL_LDG(reinterpret_cast<const float2*>(row + xs))
                                : make_float2(0.0f, 0.0f);
  }
  float2 v;
  v.x = (xs >= 0 && xs < sw) ? RDL_LDG(row + xs) : 0.0f;
  v.y = (xs + 1 >= 0 && xs + 1 < sw) ? RDL_LDG(row + xs + 1) : 0.0f;
  return v;
}

template <int RA>
RDL_HD void RowForwardPass1(const RowPlan& p, const RowIo& io, float2* smem, int row0, int n_rows,
                            int tid, int n_threads) {
  const int tasks = p.l2;
  // 64-bit loads when the pixel pairs are 8-byte aligned in the source
  const bool fast = ((io.ox | io.sw) & 1) == 0 && (reinterpret_cast<uintptr_t>(io.src) & 7) == 0;
  for (int idx = tid; idx < n_rows * tasks; idx += n_threads) {
    const int r = idx / tasks, j = idx - r * tasks;
    const float* row = io.src + size_t(row0 + r) * io.sw;
    float2 v[RA];
    StaticFor<RA>([&](auto q) { v[q] = LoadPixelPair(row, j + p.l2 * q, io.ox, io.sw, fast); });
    Dft<RA, false>::Run(v);
    float2* s = smem + r * p.row_stride + j;
    s[0] = v[0];
    StaticFor<RA - 1>([&](auto qq) {
      constexpr int q = decltype(qq)::value + 1;
      s[q * p.pitch_l2] = CMulF(v[q], RDL_LDG(p.tw1 + (q - 1) * p.l2 + j));
    });
  }
}

template <int RB>
RDL_HD void RowForwardPass2(const RowPlan& p, float2* smem, int n_rows, int tid, int n_threads) {
  const int tasks = p.ra * p.rc;
  for (int idx = tid; idx < n_rows * tasks; idx += n_threads) {
    const int r = idx / tasks, t = idx - r * tasks;
    const int b = t / p.rc, j = t - b * p.rc;
    float2* s = smem + r * p.row_stride + b * p.pitch_l2 + j;
    float2 v[RB];
    StaticFor<RB>([&](auto q) { v[q] = s[q * p.rc]; });
    Dft<RB, false>::Run(v);
    s[0] = v[0];
    StaticFor<RB - 1>([&](auto qq) {
      constexpr int q = decltype(qq)::value + 1;
      s[q * p.rc] = CMulF(v[q], RDL_LDG(p.tw2 + (q - 1) * p.rc + j));
    });
  }
}

template <int RC>
RDL_HD void RowForwardPass3(const RowPlan& p, const RowIo& io, const float2* smem, int row0,
                            int n_rows, int tid, int n_threads) {
  const int tasks = p.n_pairs;
  const int stride = p.ra * p.rb;
  for (int idx = tid; idx < n_rows * tasks; idx += n_threads) {
    const int r = idx / tasks, t = idx - r * tasks;
    const uint4 pr = RDL_LDG(p.pairs + t);
    const float2* s = smem + r * p.row_stride;
    float2* out = io.spec + size_t(io.oy + row0 + r) * io.pitch;
    float2 a[RC];
    StaticFor<RC>([&](auto q) { a[q] = s[pr.x + q]; });
    Dft<RC, false>::Run(a);
    if (t == 0) {
      // block (ka, kb) = (0, 0): k = stride * q pairs with stride * (RC - q); k = 0 also gives X[n]
      StaticFor<RC>([&](auto qq) {
        constexpr int q = decltype(qq)::value;
        const int k = stride * q;
        out[k] = SplitForward(a[q], a[(RC - q) % RC], RDL_LDG(p.split_tw + k));
      });
      out[p.n] = SplitForward(a[0], a[0], make_float2(-1.0f, 0.0f));
      continue;
    }
    const bool self = pr.x == pr.y;
    float2 b[RC];
    if (!self) {
      StaticFor<RC>([&](auto q) { b[q] = s[pr.y + q]; });
      Dft<RC, false>::Run(b);
    } else {
      StaticFor<RC>([&](auto q) { b[q] = a[q]; });
    }
    StaticFor<RC>([&](auto qq) {
      constexpr int q = decltype(qq)::value;
      const int k = int(pr.z) + stride * q;           // n - k = pr.w + stride * (RC - 1 - q)
      const float2 w = RDL_LDG(p.split_tw + k);
      out[k] = SplitForward(a[q], b[RC - 1 - q], w);
      // w^(n-k) = -conj(w^k)
      if (!self) out[p.n - k] = SplitForward(b[RC - 1 - q], a[q], make_float2(-w.x, w.y));
    });
  }
}

template <int RC>
RDL_HD void RowInversePass1(const RowPlan& p, const RowIo& io, float2* smem, int row0, int n_rows,
                            int tid, int n_threads) {
  const int tasks = p.n_pairs;
  const int stride = p.ra * p.rb;
  for (int idx = tid; idx < n_rows * tasks; idx += n_threads) {
    const int r = idx / tasks, t = idx - r * tasks;
    const uint4 pr = RDL_LDG(p.pairs + t);
    float2* s = smem + r * p.row_stride;
    const float2* in = io.spec + size_t(io.oy + row0 + r) * io.pitch;
    float2 a[RC];
    if (t == 0) {
      float2 y[RC + 1];
      StaticFor<RC>([&](auto q) { y[q] = in[stride * q]; });
      y[RC] = in[p.n];
      StaticFor<RC>([&](auto qq) {
        constexpr int q = decltype(qq)::value;
        a[q] = SplitInverse(y[q], y[RC - q], RDL_LDG(p.split_tw + stride * q));
      });
      Dft<RC, true>::Run(a);
      StaticFor<RC>([&](auto q) { s[pr.x + q] = a[q]; });
      continue;
    }
    const bool self = pr.x == pr.y;
    float2 b[RC];
    StaticFor<RC>([&](auto qq) {
      constexpr int q = decltype(qq)::value;
      const int k = int(pr.z) + stride * q;
      const float2 ya = in[k], yb = in[p.n - k];
      const float2 w = RDL_LDG(p.split_tw + k);
      a[q] = SplitInverse(ya, yb, w);
      b[RC - 1 - q] = SplitInverse(yb, ya, make_float2(-w.x, w.y));
    });
    Dft<RC, true>::Run(a);
    StaticFor<RC>([&](auto q) { s[pr.x + q] = a[q]; });
    if (!self) {
      Dft<RC, true>::Run(b);
      StaticFor<RC>([&](auto q) { s[pr.y + q] = b[q]; });
    }
  }
}

template <int RB>
RDL_HD void RowInversePass2(const RowPlan& p, float2* smem, int n_rows, int tid, int n_threads) {
  const int tasks = p.ra * p.rc;
  for (int idx = tid; idx < n_rows * tasks; idx += n_threads) {
    const int r = idx / tasks, t = idx - r * tasks;
    const int b = t / p.rc, j = t - b * p.rc;
    float2* s = smem + r * p.row_stride + b * p.pitch_l2 + j;
    float2 v[RB];
    v[0] = s[0];
    StaticFor<RB - 1>([&](auto qq) {
      constexpr int q = decltype(qq)::value + 1;
      v[q] = CMulConj(s[q * p.rc], RDL_LDG(p.tw2 + (q - 1) * p.rc + j));
    });
    Dft<RB, true>::Run(v);
    StaticFor<RB>([&](auto q) { s[q * p.rc] = v[q]; });
  }
}

template <int RA>
RDL_HD void RowInversePass3(const RowPlan& p, const RowIo& io, const float2* smem, int row0,
                            int n_rows, int tid, int n_threads) {
  const int tasks = p.l2;
  const bool fast = ((io.ox | io.sw) & 1) == 0 && (reinterpret_cast<uintptr_t>(io.dst) & 7) == 0;
  for (int idx = tid; idx < n_rows * tasks; idx += n_threads) {
    const int r = idx / tasks, j = idx - r * tasks;
    const float2* s = smem + r * p.row_stride + j;
    float2 v[RA];
    v[0] = s[0];
    StaticFor<RA - 1>([&](auto qq) {
      constexpr int q = decltype(qq)::value + 1;
      v[q] = CMulConj(s[q * p.pitch_l2], RDL_LDG(p.tw1 + (q - 1) * p.l2 + j));
    });
    Dft<RA, true>::Run(v);
    float* row = io.dst + size_t(row0 + r) * io.sw;
    StaticFor<RA>([&](auto q) {
      const int xs = 2 * (j + p.l2 * q) - io.ox;
      float2 o = v[q];
      if (fast) {
        if (xs >= 0 && xs < io.sw) {
          float2* d = reinterpret_cast<float2*>(row + xs);
          if (io.mode != 0) {
            const float2 old = *d;
            o = io.mode == 1 ? make_float2(old.x - o.x, old.y - o.y)
                             : make_float2(old.x + o.x, old.y + o.y);
          }
          *d = o;
        }
      } else {
        if (xs >= 0 && xs < io.sw)
          row[xs] = io.mode == 0 ? o.x : (io.mode == 1 ? row[xs] - o.x : row[xs] + o.x);
        if (xs + 1 >= 0 && xs + 1 < io.sw)
          row[xs + 1] = io.mode == 0 ? o.y : (io.mode == 1 ? row[xs + 1] - o.y : row[xs + 1] + o.y);
      }
    });
  }
}

// ---- columns -------------------------------------------------------------------------------------
// H = H1 * H2 (four-step). Strided pass: the H2-point transforms over rows r, r + H1, ... held in
// registers straight from global memory, times the step twiddle w_H^(r k2); frequency k2 goes to
// row r + H1 * k2 (natural order, in place). Block pass: H1 = ra * rb contiguous rows of a
// 16-column tile: pass A (radix ra over rows at stride rb, from global memory), pass B (radix rb,
// kernel spectrum product, inverse radix rb: the same registers), pass C (inverse radix ra, to
// global memory): two shared-memory exchanges for forward + product + inverse.
constexpr int kTile = 16;  // complex columns per tile: one 128-byte line per spectrum row

struct ColPlan {
  int H, H1, H2;
  int ra, rb;  // H1 = ra * rb
  const float2* tw1;   // exp(-2 pi i t / H1)
  const float2* tw_h;  // exp(-2 pi i t / H)
};

struct ColIo {
  float2* spec;          // work spectrum, in place
  const float2* kspec;   // kernel spectrum (convolve)
  float2* kspec_out;     // forward-only: scaled spectrum
  int pitch;
  int y0, y1;            // rows of spec that hold data on the way in (others count as zero)
  int out_y0, out_y1;    // rows the inverse row pass will read
  float scale;
};

// grid: (column tiles, ceil(H1 / rows_per_cta)); thread = (r_local, column)
template <int R2, bool INV>
RDL_HD void ColumnStridedBody(const ColPlan& p, const ColIo& io, int c0, int r, int c) {
  float2* base = io.spec + size_t(r) * io.pitch + c0 + c;
  const size_t step = size_t(p.H1) * io.pitch;
  float2 v[R2];
  if (!INV) {
    StaticFor<R2>([&](auto j) {
      const int y = r + p.H1 * j;
      v[j] = (y >= io.y0 && y < io.y1) ? base[j * step] : make_float2(0.0f, 0.0f);
    });
    Dft<R2, false>::Run(v);
    base[0] = v[0];
    StaticFor<R2 - 1>([&](auto kk) {
      constexpr int k = decltype(kk)::value + 1;
      base[k * step] = CMulF(v[k], RDL_LDG(p.tw_h + r * k));
    });
  } else {
    v[0] = base[0];
    StaticFor<R2 - 1>([&](auto kk) {
      constexpr int k = decltype(kk)::value + 1;
      v[k] = CMulConj(base[k * step], RDL_LDG(p.tw_h + r * k));
    });
    Dft<R2, true>::Run(v);
    StaticFor<R2>([&](auto j) {
      const int y = r + p.H1 * j;
      if (y >= io.out_y0 && y < io.out_y1) base[j * step] = v[j];
    });
  }
}

// tile of rows [row0, row0 + H1) x kTile columns at c0; smem holds H1 * kTile elements
template <int RA>
RDL_HD void ColumnBlockPassA(const ColPlan& p, const ColIo& io, float2* smem, size_t row0, int c0,
                             int tid, int n_threads) {
  const int tasks = p.rb * kTile;
  const bool whole = p.H2 == 1;
  for (int idx = tid; idx < tasks; idx += n_threads) {
    const int j = idx / kTile, c = idx - j * kTile;
    const float2* g = io.spec + (row0 + j) * io.pitch + c0 + c;
    float2 v[RA];
    StaticFor<RA>([&](auto q) {
      const int y = j + p.rb * q;
      v[q] = (!whole || (y >= io.y0 && y < io.y1)) ? g[size_t(p.rb) * q * io.pitch]
                                                  : make_float2(0.0f, 0.0f);
    });
    Dft<RA, false>::Run(v);
    float2* s = smem + j * kTile + c;
    s[0] = v[0];
    StaticFor<RA - 1>([&](auto qq) {
      constexpr int q = decltype(qq)::value + 1;
      s[q * p.rb * kTile] = CMulF(v[q], RDL_LDG(p.tw1 + j * q));
    });
  }
}

template <int RB, bool CONVOLVE>
RDL_HD void ColumnBlockPassB(const ColPlan& p, const ColIo& io, float2* smem, size_t row0, int c0,
                             int tid, int n_threads) {
  const int tasks = p.ra * kTile;
  for (int idx = tid; idx < tasks; idx += n_threads) {
    const int ka = idx / kTile, c = idx - ka * kTile;
    float2* s = smem + ka * p.rb * kTile + c;
    const size_t g0 = (row0 + size_t(ka) * p.rb) * io.pitch + c0 + c;
    float2 v[RB], k[RB];
    if (CONVOLVE) StaticFor<RB>([&](auto q) { k[q] = RDL_LDG(io.kspec + g0 + size_t(q) * io.pitch); });
    StaticFor<RB>([&](auto q) { v[q] = s[q * kTile]; });
    Dft<RB, false>::Run(v);
    if (!CONVOLVE) {
      StaticFor<RB>([&](auto q) {
        io.kspec_out[g0 + size_t(q) * io.pitch] = make_float2(io.scale * v[q].x, io.scale * v[q].y);
      });
      continue;
    }
    StaticFor<RB>([&](auto q) { v[q] = CMulF(v[q], k[q]); });
    Dft<RB, true>::Run(v);
    // the conjugate twiddle of the inverse radix-ra pass (element j of block ka)
    StaticFor<RB>([&](auto jj) {
      constexpr int j = decltype(jj)::value;
      s[j * kTile] = (j == 0 || ka == 0) ? v[j] : CMulConj(v[j], RDL_LDG(p.tw1 + j * ka));
    });
  }
}

template <int RA>
RDL_HD void ColumnBlockPassC(const ColPlan& p, const ColIo& io, const float2* smem, size_t row0,
                             int c0, int tid, int n_threads) {
  const int tasks = p.rb * kTile;
  const bool whole = p.H2 == 1;
  for (int idx = tid; idx < tasks; idx += n_threads) {
    const int j = idx / kTile, c = idx - j * kTile;
    const float2* s = smem + j * kTile + c;
    float2 v[RA];
    StaticFor<RA>([&](auto q) { v[q] = s[q * p.rb * kTile]; });
    Dft<RA, true>::Run(v);
    float2* g = io.spec + (row0 + j) * io.pitch + c0 + c;
    StaticFor<RA>([&](auto q) {
      const int y = j + p.rb * q;
      if (!whole || (y >= io.out_y0 && y < io.out_y1)) g[size_t(p.rb) * q * io.pitch] = v[q];
    });
  }
}

// ---- whole columns in shared memory -----------------------------------------------------------
// The four-step scheme above moves the spectrum through HBM three times (strided, block,
// strided). Where a column tile of the WHOLE height fits in shared memory -- tw = 4 columns up to
// H ~ 6000, 2 columns up to H ~ 12000 -- one kernel does everything: H = ra * rb * rc forward
// (decimation in frequency, three passes), product with the kernel spectrum, three inverse
// passes. A 2-column tile touches 16 bytes per spectrum row, so neighbouring tiles share
// sectors and DRAM bursts: they run on different SMs at about the same time and meet in L2.
// Frequency k = ka + ra * kb + ra * rb * kc sits at position pos = (ka * rb + kb) * rc + kc, in
// shared memory as well as in the kernel spectrum (which the forward-only flavour writes in that
// order). One padding row per rc positions keeps the last pass (one thread per rc-run) off a
// single bank.
struct WholePlan {
  int H, ra, rb, rc, l2;  // l2 = rb * rc
  int tw;                 // tile width in columns (2 or 4)
  // twa[(q - 1) * l2 + j] = exp(-2 pi i j q / H), twb[(q - 1) * rc + j] = exp(-2 pi i j q / l2)
  const float2* twa;
  const float2* twb;
};

RDL_HD int WholeSlot(const WholePlan& p, int pos) { return pos + pos / p.rc; }
RDL_HD size_t WholeSmemElements(const WholePlan& p) { return size_t(p.H + p.H / p.rc) * p.tw; }

template <int RA>
RDL_HD void WholePassA(const WholePlan& p, const ColIo& io, float2* smem, int c0, int tid, int n_threads) {
  const int tasks = p.l2 * p.tw;
  for (int idx = tid; idx < tasks; idx += n_threads) {
    const int j = idx / p.tw, c = idx - j * p.tw;
    const float2* g = io.spec + size_t(j) * io.pitch + c0 + c;
    float2 v[RA];
    StaticFor<RA>([&](auto q) {
      const int y = j + p.l2 * q;
      v[q] = (y >= io.y0 && y < io.y1) ? g[size_t(p.l2) * q * io.pitch] : make_float2(0.0f, 0.0f);
    });
    Dft<RA, false>::Run(v);
    smem[WholeSlot(p, j) * p.tw + c] = v[0];
    StaticFor<RA - 1>([&](auto qq) {
      constexpr int q = decltype(qq)::value + 1;
      smem[WholeSlot(p, j + p.l2 * q) * p.tw + c] = CMulF(v[q], RDL_LDG(p.twa + (q - 1) * p.l2 + j));
    });
  }
}

template <int RB, bool INV>
RDL_HD void WholePassB(const WholePlan& p, float2* smem, int tid, int n_threads) {
  const int per_block = p.rc * p.tw;
  const int tasks = p.ra * per_block;
  for (int idx = tid; idx < tasks; idx += n_threads) {
    const int ka = idx / per_block, t = idx - ka * per_block;
    const int j = t / p.tw, c = t - j * p.tw;
    const int base = ka * p.l2 + j;
    float2 v[RB];
    if (!INV) {
      StaticFor<RB>([&](auto q) { v[q] = smem[WholeSlot(p, base + p.rc * q) * p.tw + c]; });
      Dft<RB, false>::Run(v);
      smem[WholeSlot(p, base) * p.tw + c] = v[0];
      StaticFor<RB - 1>([&](auto qq) {
        constexpr int q = decltype(qq)::value + 1;
        smem[WholeSlot(p, base + p.rc * q) * p.tw + c] = CMulF(v[q], RDL_LDG(p.twb + (q - 1) * p.rc + j));
      });
    } else {
      v[0] = smem[WholeSlot(p, base) * p.tw + c];
      StaticFor<RB - 1>([&](auto qq) {
        constexpr int q = decltype(qq)::value + 1;
        v[q] = CMulConj(smem[WholeSlot(p, base + p.rc * q) * p.tw + c], RDL_LDG(p.twb + (q - 1) * p.rc + j));
      });
      Dft<RB, true>::Run(v);
      StaticFor<RB>([&](auto q) { smem[WholeSlot(p, base + p.rc * q) * p.tw + c] = v[q]; });
    }
  }
}

// last forward pass, product (or the scaled kernel spectrum out), first inverse pass
template <int RC, bool CONVOLVE>
RDL_HD void WholePassC(const WholePlan& p, const ColIo& io, float2* smem, int c0, int tid, int n_threads) {
  const int tasks = p.ra * p.rb * p.tw;
  for (int idx = tid; idx < tasks; idx += n_threads) {
    const int b = idx / p.tw, c = idx - b * p.tw;  // b = ka * rb + kb
    const int pos0 = b * p.rc;
    float2* s = smem + WholeSlot(p, pos0) * p.tw + c;  // the rc positions of a run share one pad offset
    const size_t g0 = size_t(pos0) * io.pitch + c0 + c;
    float2 v[RC], k[RC];
    if (CONVOLVE) StaticFor<RC>([&](auto q) { k[q] = RDL_LDG(io.kspec + g0 + size_t(q) * io.pitch); });
    StaticFor<RC>([&](auto q) { v[q] = s[q * p.tw]; });
    Dft<RC, false>::Run(v);
    if (!CONVOLVE) {
      StaticFor<RC>([&](auto q) {
        io.kspec_out[g0 + size_t(q) * io.pitch] = make_float2(io.scale * v[q].x, io.scale * v[q].y);
      });
      continue;
    }
    StaticFor<RC>([&](auto q) { v[q] = CMulF(v[q], k[q]); });
    Dft<RC, true>::Run(v);
    StaticFor<RC>([&](auto q) { s[q * p.tw] = v[q]; });
  }
}

template <int RA>
RDL_HD void WholePassD(const WholePlan& p, const ColIo& io, const float2* smem, int c0, int tid,
                       int n_threads) {
  const int tasks = p.l2 * p.tw;
  for (int idx = tid; idx < tasks; idx += n_threads) {
    const int j = idx / p.tw, c = idx - j * p.tw;
    float2 v[RA];
    v[0] = smem[WholeSlot(p, j) * p.tw + c];
    StaticFor<RA - 1>([&](auto qq) {
      constexpr int q = decltype(qq)::value + 1;
      v[q] = CMulConj(smem[WholeSlot(p, j + p.l2 * q) * p.tw + c], RDL_LDG(p.twa + (q - 1) * p.l2 + j));
    });
    Dft<RA, true>::Run(v);
    float2* g = io.spec + size_t(j) * io.pitch + c0 + c;
    StaticFor<RA>([&](auto q) {
      const int y = j + p.l2 * q;
      if (y >= io.out_y0 && y < io.out_y1) g[size_t(p.l2) * q * io.pitch] = v[q];
    });
  }
}

// all phases of one tile, thread by thread (the kernels in fft_spec.cu put a __syncthreads()
// between the phases; the CPU emulation runs phase after phase over all threads)
template <int PHASE, int RA, int RB, int RC, bool CONVOLVE>
RDL_HD void WholeColumnPhase(const WholePlan& p, const ColIo& io, float2* smem, int tile, int tid,
                             int n_threads) {
  const int c0 = tile * p.tw;
  if (PHASE == 0) {
    WholePassA<RA>(p, io, smem, c0, tid, n_threads);
  } else if (PHASE == 1) {
    WholePassB<RB, false>(p, smem, tid, n_threads);
  } else if (PHASE == 2) {
    WholePassC<RC, CONVOLVE>(p, io, smem, c0, tid, n_threads);
  } else if (PHASE == 3) {
    WholePassB<RB, true>(p, smem, tid, n_threads);
  } else {
    WholePassD<RA>(p, io, smem, c0, tid, n_threads);
  }
}

// ---- kernel bodies by phase (a __syncthreads() separates the phases of one CTA) -----------------
RDL_HD int MinInt(int a, int b) { return a < b ? a : b; }

template <int PHASE>
RDL_HD void RowForwardPhase(const RowPlan& p, const RowIo& io, float2* smem, int block, int tid,
                            int n_threads) {
  const int row0 = block * p.rows_per_cta;
  const int n_rows = MinInt(p.rows_per_cta, io.sh - row0);
  if (PHASE == 0) {
    RDL_DISPATCH_RADIX(p.ra, (RowForwardPass1<kR>(p, io, smem, row0, n_rows, tid, n_threads)))
  } else if (PHASE == 1) {
    if (p.rb > 1) RDL_DISPATCH_RADIX(p.rb, (RowForwardPass2<kR>(p, smem, n_rows, tid, n_threads)))
  } else {
    RDL_DISPATCH_LAST_RADIX(p.rc, (RowForwardPass3<kR>(p, io, smem, row0, n_rows, tid, n_threads)))
  }
}

template <int PHASE>
RDL_HD void RowInversePhase(const RowPlan& p, const RowIo& io, float2* smem, int block, int tid,
                            int n_threads) {
  const int row0 = block * p.rows_per_cta;
  const int n_rows = MinInt(p.rows_per_cta, io.sh - row0);
  if (PHASE == 0) {
    RDL_DISPATCH_LAST_RADIX(p.rc, (RowInversePass1<kR>(p, io, smem, row0, n_rows, tid, n_threads)))
  } else if (PHASE == 1) {
    if (p.rb > 1) RDL_DISPATCH_RADIX(p.rb, (RowInversePass2<kR>(p, smem, n_rows, tid, n_threads)))
  } else {
    RDL_DISPATCH_RADIX(p.ra, (RowInversePass3<kR>(p, io, smem, row0, n_rows, tid, n_threads)))
  }
}

constexpr int kStridedRows = 8;  // r values per CTA of the strided pass (x kTile columns)

template <bool INV>
RDL_HD void ColumnStridedThread(const ColPlan& p, const ColIo& io, int tile, int row_group, int tid) {
  const int c = tid % kTile, r = row_group * kStridedRows + tid / kTile;
  if (r >= p.H1) return;
  RDL_DISPATCH_RADIX(p.H2, (ColumnStridedBody<kR, INV>(p, io, tile * kTile, r, c)))
}

template <int PHASE, bool CONVOLVE>
RDL_HD void ColumnBlockPhase(const ColPlan& p, const ColIo& io, float2* smem, int tile, int block,
                             int tid, int n_threads) {
  const size_t row0 = size_t(block) * p.H1;
  const int c0 = tile * kTile;
  if (PHASE == 0) {
    RDL_DISPATCH_RADIX(p.ra, (ColumnBlockPassA<kR>(p, io, smem, row0, c0, tid, n_threads)))
  } else if (PHASE == 1) {
    RDL_DISPATCH_RADIX(p.rb, (ColumnBlockPassB<kR, CONVOLVE>(p, io, smem, row0, c0, tid, n_threads)))
  } else if (CONVOLVE) {
    RDL_DISPATCH_RADIX(p.ra, (ColumnBlockPassC<kR>(p, io, smem, row0, c0, tid, n_threads)))
  }
}

}  // namespace fast
}  // namespace rdl
#endif
