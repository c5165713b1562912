// CPU emulation of the register-radix FFT kernels (csrc/fft_fast.cuh): the pass bodies are
// __host__ __device__, so this program runs them thread by thread, phase by phase, exactly as
// the kernels in csrc/fft.cu launch them, and compares the circular convolution with a float64
// DFT. Built and run by tests/test_fft_emulation.py (no GPU needed).
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <random>
#include <string>
#include <vector>

#include "../ska_sdp_func_radler_b200/csrc/fft_fast_plan.h"

using namespace rdl::fast;

namespace {

bool g_whole = false;  // column transforms through the whole-column kernel bodies where planned

struct Engine {
  HostFastPlan hp;
  RowPlan rp;
  ColPlan cp;
  WholePlan wp;
  bool whole = false;
  std::vector<float2> work;

  bool Build(int W, int H) {
    hp.Build(W, H);
    if (!hp.ok) return false;
    rp = hp.MakeRowPlan(hp.tw_row1.data(), hp.tw_row2.data(), hp.tw_split.data(), hp.pairs.data());
    cp = hp.MakeColPlan(hp.tw_h1.data(), hp.tw_h.data());
    whole = g_whole && hp.whole_ok;
    if (g_whole && !hp.whole_ok) return false;
    if (whole) wp = hp.MakeWholePlan(hp.tw_whole_a.data(), hp.tw_whole_b.data());
    work.assign(size_t(H) * hp.pitch, make_float2(0.0f, 0.0f));
    return true;
  }

  void RowForward(const float* src, int sw, int sh, int ox, int oy) {
    RowIo io{};
    io.src = src;
    io.sw = sw;
    io.sh = sh;
    io.ox = ox;
    io.oy = oy;
    io.spec = work.data();
    io.pitch = hp.pitch;
    std::vector<float2> smem(hp.row_smem / sizeof(float2));
    const int grid = (sh + hp.rows_per_cta - 1) / hp.rows_per_cta;
    for (int b = 0; b < grid; ++b) {
      for (int t = 0; t < hp.row_threads; ++t) RowForwardPhase<0>(rp, io, smem.data(), b, t, hp.row_threads);
      for (int t = 0; t < hp.row_threads; ++t) RowForwardPhase<1>(rp, io, smem.data(), b, t, hp.row_threads);
      for (int t = 0; t < hp.row_threads; ++t) RowForwardPhase<2>(rp, io, smem.data(), b, t, hp.row_threads);
    }
  }

  void RowInverse(float* dst, int dw, int dh, int ox, int oy, int mode) {
    RowIo io{};
    io.dst = dst;
    io.sw = dw;
    io.sh = dh;
    io.ox = ox;
    io.oy = oy;
    io.spec = work.data();
    io.pitch = hp.pitch;
    io.mode = mode;
    std::vector<float2> smem(hp.row_smem / sizeof(float2));
    const int grid = (dh + hp.rows_per_cta - 1) / hp.rows_per_cta;
    for (int b = 0; b < grid; ++b) {
      for (int t = 0; t < hp.row_threads; ++t) RowInversePhase<0>(rp, io, smem.data(), b, t, hp.row_threads);
      for (int t = 0; t < hp.row_threads; ++t) RowInversePhase<1>(rp, io, smem.data(), b, t, hp.row_threads);
      for (int t = 0; t < hp.row_threads; ++t) RowInversePhase<2>(rp, io, smem.data(), b, t, hp.row_threads);
    }
  }

  template <bool INV>
  void Strided(const ColIo& io) {
    const int tiles = hp.pitch / kTile, groups = (cp.H1 + kStridedRows - 1) / kStridedRows;
    for (int tile = 0; tile < tiles; ++tile)
      for (int g = 0; g < groups; ++g)
        for (int t = 0; t < kTile * kStridedRows; ++t) ColumnStridedThread<INV>(cp, io, tile, g, t);
  }

  template <bool CONVOLVE>
  void Block(const ColIo& io) {
    const int tiles = hp.pitch / kTile;
    std::vector<float2> smem(hp.block_smem / sizeof(float2));
    for (int tile = 0; tile < tiles; ++tile)
      for (int b = 0; b < cp.H2; ++b) {
        for (int t = 0; t < hp.block_threads; ++t)
          ColumnBlockPhase<0, CONVOLVE>(cp, io, smem.data(), tile, b, t, hp.block_threads);
        for (int t = 0; t < hp.block_threads; ++t)
          ColumnBlockPhase<1, CONVOLVE>(cp, io, smem.data(), tile, b, t, hp.block_threads);
        for (int t = 0; t < hp.block_threads; ++t)
          ColumnBlockPhase<2, CONVOLVE>(cp, io, smem.data(), tile, b, t, hp.block_threads);
      }
  }

  // one tile after the other, phase by phase, every thread of the 512
  template <bool CONVOLVE>
  void Whole(const ColIo& io) {
    const int tiles = hp.pitch / wp.tw, threads = hp.whole_threads;
    std::vector<float2> smem(WholeSmemElements(wp));
    for (int tile = 0; tile < tiles; ++tile) {
      const int c0 = tile * wp.tw;
      for (int t = 0; t < threads; ++t)
        RDL_DISPATCH_RADIX(wp.ra, (WholePassA<kR>(wp, io, smem.data(), c0, t, threads)))
      for (int t = 0; t < threads; ++t)
        RDL_DISPATCH_RADIX(wp.rb, (WholePassB<kR, false>(wp, smem.data(), t, threads)))
      for (int t = 0; t < threads; ++t)
        RDL_DISPATCH_LAST_RADIX(wp.rc, (WholePassC<kR, CONVOLVE>(wp, io, smem.data(), c0, t, threads)))
      if (!CONVOLVE) continue;
      for (int t = 0; t < threads; ++t)
        RDL_DISPATCH_RADIX(wp.rb, (WholePassB<kR, true>(wp, smem.data(), t, threads)))
      for (int t = 0; t < threads; ++t)
        RDL_DISPATCH_RADIX(wp.ra, (WholePassD<kR>(wp, io, smem.data(), c0, t, threads)))
    }
  }

  ColIo MakeIo() {
    ColIo io{};
    io.spec = work.data();
    io.pitch = hp.pitch;
    io.y0 = 0;
    io.y1 = hp.H;
    io.out_y0 = 0;
    io.out_y1 = hp.H;
    io.scale = 1.0f;
    return io;
  }

  void KernelSpectrum(const float* kernel, std::vector<float2>& out) {
    out.assign(size_t(hp.H) * hp.pitch, make_float2(0.0f, 0.0f));
    RowForward(kernel, hp.W, hp.H, 0, 0);
    ColIo io = MakeIo();
    io.kspec_out = out.data();
    io.scale = float(1.0 / (4.0 * double(hp.W) * double(hp.H)));
    if (whole) {
      Whole<false>(io);
      return;
    }
    if (cp.H2 > 1) Strided<false>(io);
    Block<false>(io);
  }

  void Convolve(const float* src, int sw, int sh, int sox, int soy, const std::vector<float2>& k,
                float* dst, int dw, int dh, int dox, int doy, int mode) {
    RowForward(src, sw, sh, sox, soy);
    ColIo io = MakeIo();
    io.kspec = k.data();
    io.y0 = soy;
    io.y1 = soy + sh;
    io.out_y0 = doy;
    io.out_y1 = doy + dh;
    if (whole) {
      Whole<true>(io);
    } else {
      if (cp.H2 > 1) Strided<false>(io);
      Block<true>(io);
      if (cp.H2 > 1) Strided<true>(io);
    }
    RowInverse(dst, dw, dh, dox, doy, mode);
  }
};

// float64 circular convolution of a W x H frame by separable DFTs (O(N^2) per transform)
void Dft1d(std::vector<double>& re, std::vector<double>& im, int n, int stride, int offset, bool inv,
           const std::vector<double>& c, const std::vector<double>& s, std::vector<double>& tr,
           std::vector<double>& ti) {
  for (int k = 0; k < n; ++k) {
    double ar = 0, ai = 0;
    long idx = 0;
    for (int j = 0; j < n; ++j) {
      const double wr = c[idx], wi = inv ? s[idx] : -s[idx];
      const double xr = re[offset + size_t(j) * stride], xi = im[offset + size_t(j) * stride];
      ar += xr * wr - xi * wi;
      ai += xr * wi + xi * wr;
      idx += k;
      if (idx >= n) idx -= n;
    }
    tr[k] = ar;
    ti[k] = ai;
  }
  for (int k = 0; k < n; ++k) {
    re[offset + size_t(k) * stride] = tr[k];
    im[offset + size_t(k) * stride] = ti[k];
  }
}

void Dft2d(std::vector<double>& re, std::vector<double>& im, int W, int H, bool inv) {
  std::vector<double> c(W), s(W), tr(std::max(W, H)), ti(std::max(W, H));
  for (int i = 0; i < W; ++i) {
    c[i] = std::cos(2 * M_PI * i / W);
    s[i] = std::sin(2 * M_PI * i / W);
  }
  for (int y = 0; y < H; ++y) Dft1d(re, im, W, 1, y * W, inv, c, s, tr, ti);
  c.resize(H);
  s.resize(H);
  for (int i = 0; i < H; ++i) {
    c[i] = std::cos(2 * M_PI * i / H);
    s[i] = std::sin(2 * M_PI * i / H);
  }
  for (int x = 0; x < W; ++x) Dft1d(re, im, H, W, x, inv, c, s, tr, ti);
}

std::vector<double> ReferenceConvolve(const std::vector<float>& frame, const std::vector<float>& kernel,
                                      int W, int H) {
  std::vector<double> ar(frame.begin(), frame.end()), ai(size_t(W) * H, 0.0);
  std::vector<double> kr(kernel.begin(), kernel.end()), ki(size_t(W) * H, 0.0);
  Dft2d(ar, ai, W, H, false);
  Dft2d(kr, ki, W, H, false);
  for (size_t i = 0; i < ar.size(); ++i) {
    const double r = ar[i] * kr[i] - ai[i] * ki[i], m = ar[i] * ki[i] + ai[i] * kr[i];
    ar[i] = r;
    ai[i] = m;
  }
  Dft2d(ar, ai, W, H, true);
  for (double& v : ar) v /= double(W) * H;
  return ar;
}

int Check(int W, int H, int sw, int sh, int dw, int dh, int mode, unsigned seed) {
  Engine e;
  if (!e.Build(W, H)) {
    std::printf("%5d x %5d: no fast plan\n", W, H);
    return 0;
  }
  std::mt19937 rng(seed);
  std::normal_distribution<float> dist;
  std::vector<float> src(size_t(sw) * sh), kernel(size_t(W) * H), dst(size_t(dw) * dh);
  for (float& v : src) v = dist(rng);
  for (float& v : kernel) v = dist(rng);
  for (float& v : dst) v = dist(rng);
  const std::vector<float> dst0 = dst;
  const int sox = (W - sw) / 2, soy = (H - sh) / 2, dox = (W - dw) / 2, doy = (H - dh) / 2;
  std::vector<float2> spectrum;
  e.KernelSpectrum(kernel.data(), spectrum);
  e.Convolve(src.data(), sw, sh, sox, soy, spectrum, dst.data(), dw, dh, dox, doy, mode);
  std::vector<float> frame(size_t(W) * H, 0.0f);
  for (int y = 0; y < sh; ++y)
    for (int x = 0; x < sw; ++x) frame[size_t(y + soy) * W + x + sox] = src[size_t(y) * sw + x];
  const std::vector<double> want = ReferenceConvolve(frame, kernel, W, H);
  double err = 0, peak = 0;
  for (int y = 0; y < dh; ++y)
    for (int x = 0; x < dw; ++x) {
      const double w = want[size_t(y + doy) * W + x + dox];
      const double old = dst0[size_t(y) * dw + x];
      const double expect = mode == 0 ? w : (mode == 1 ? old - w : old + w);
      err = std::max(err, std::fabs(expect - double(dst[size_t(y) * dw + x])));
      peak = std::max(peak, std::fabs(w));
    }
  const double rel = err / peak;
  std::printf("%5d x %5d (src %dx%d dst %dx%d mode %d) rows %d*%d*%d cols %d*(%d*%d) threads %d/%d "
              "rows/cta %d: rel err %.3g %s\n",
              W, H, sw, sh, dw, dh, mode, e.hp.row.ra, e.hp.row.rb, e.hp.row.rc, e.hp.col.h2,
              e.hp.col.ra, e.hp.col.rb, e.hp.row_threads, e.hp.block_threads, e.hp.rows_per_cta, rel,
              rel < 2e-6 ? "ok" : "FAIL");
  return rel < 2e-6 ? 0 : 1;
}

}  // namespace

// writes the inputs and the emulated result for tests/test_fft_emulation.py (numpy reference)
int Dump(int W, int H, int sw, int sh, int dw, int dh, int mode, unsigned seed, const char* path) {
  Engine e;
  if (!e.Build(W, H)) return 2;
  std::mt19937 rng(seed);
  std::normal_distribution<float> dist;
  std::vector<float> src(size_t(sw) * sh), kernel(size_t(W) * H), dst(size_t(dw) * dh);
  for (float& v : src) v = dist(rng);
  for (float& v : kernel) v = dist(rng);
  for (float& v : dst) v = dist(rng);
  const std::vector<float> dst0 = dst;
  std::vector<float2> spectrum;
  e.KernelSpectrum(kernel.data(), spectrum);
  e.Convolve(src.data(), sw, sh, (W - sw) / 2, (H - sh) / 2, spectrum, dst.data(), dw, dh,
             (W - dw) / 2, (H - dh) / 2, mode);
  FILE* f = std::fopen(path, "wb");
  if (!f) return 3;
  std::fwrite(src.data(), sizeof(float), src.size(), f);
  std::fwrite(kernel.data(), sizeof(float), kernel.size(), f);
  std::fwrite(dst0.data(), sizeof(float), dst0.size(), f);
  std::fwrite(dst.data(), sizeof(float), dst.size(), f);
  std::fclose(f);
  std::printf("rows %d*%d*%d cols %d*(%d*%d)\n", e.hp.row.ra, e.hp.row.rb, e.hp.row.rc, e.hp.col.h2,
              e.hp.col.ra, e.hp.col.rb);
  return 0;
}

int main(int argc, char** argv) {
  int failures = 0;
  if (argc >= 2 && std::string(argv[1]) == "whole") {  // "whole dump ..." / "whole W H ..."
    g_whole = true;
    --argc;
    ++argv;
  }
  if (argc == 11 && std::string(argv[1]) == "dump")
    return Dump(std::atoi(argv[2]), std::atoi(argv[3]), std::atoi(argv[4]), std::atoi(argv[5]),
                std::atoi(argv[6]), std::atoi(argv[7]), std::atoi(argv[8]), unsigned(std::atoi(argv[9])),
                argv[10]);
  if (argc >= 3) {
    const int W = std::atoi(argv[1]), H = std::atoi(argv[2]);
    const int sw = argc >= 5 ? std::atoi(argv[3]) : W, sh = argc >= 5 ? std::atoi(argv[4]) : H;
    return Check(W, H, sw, sh, sw, sh, argc >= 6 ? std::atoi(argv[5]) : 0, 1);
  }
  // small frames: every code path (one / two / three row passes, single-pass and four-step
  // columns, odd offsets, partial source and destination, the three epilogues)
  const int frames[][2] = {{2, 1},   {4, 3},    {16, 16},   {42, 30},  {64, 64},  {128, 18}, {36, 50},
                           {250, 14}, {96, 240}, {360, 336}, {512, 28}, {28, 1024}, {70, 1296}};
  unsigned seed = 1;
  for (const auto& f : frames) {
    failures += Check(f[0], f[1], f[0], f[1], f[0], f[1], 0, seed++);
    if (f[0] >= 8 && f[1] >= 6) {
      failures += Check(f[0], f[1], f[0] - 3, f[1] - 2, f[0] - 4, f[1] - 5, 1, seed++);
      failures += Check(f[0], f[1], f[0] - 4, f[1] - 4, f[0] - 4, f[1] - 4, 2, seed++);
    }
  }
  // long rows and long columns of the BASELINE sizes (the other axis kept short)
  for (int n : {4096, 4536, 4608, 4704, 4860, 8192, 9072, 9216, 9408, 16384}) {
    failures += Check(n, 2, n, 2, n, 2, 0, seed++);
    failures += Check(32, n, 32, n, 32, n, 0, seed++);
  }
  std::printf("%s (%d failures)\n", failures ? "FAILED" : "ALL OK", failures);
  return failures ? 1 : 0;
}
