// Host-side planning for the register-radix FFT passes (fft_fast.cuh): radix triples, the
// butterfly-pair table of the split pass, twiddle tables. Pure host code, shared by fft.cu and
// tests/fft_emulation.cu.
#ifndef RDL_FFT_FAST_PLAN_H_
#define RDL_FFT_FAST_PLAN_H_

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <vector>

#include "fft_fast.cuh"

namespace rdl {
namespace fast {

inline bool IsRadix(int r) {
  for (int k : kRadices)
    if (k == r) return true;
  return false;
}

struct RowChoice {
  int ra = 0, rb = 0, rc = 0;
  bool ok = false;
};

// n = ra * rb * rc with rc <= kMaxLastRadix; fewest passes, then the smallest largest radix
// (registers), then the largest first radix (pass 1 holds a single butterfly).
inline RowChoice ChooseRowRadices(int n) {
  RowChoice best;
  long best_cost = -1;
  for (int rc : kRadices) {
    if (rc > kMaxLastRadix || n % rc != 0) continue;
    for (int rb : kRadices) {
      if ((n / rc) % rb != 0) continue;
      const int ra = n / rc / rb;
      if (!IsRadix(ra)) continue;
      if (rb == 1 && rc != 1 && ra < rc) continue;  // two passes: keep (ra, 1, rc) with ra >= rc
      if (rb != 1 && (ra == 1 || rc == 1)) continue;  // a unit radix only in the middle or at the end
      if (rb == 1 && rc == 1 && ra != n) continue;
      const int passes = (ra > 1) + (rb > 1) + (rc > 1);
      const int mx = std::max(ra, std::max(rb, rc));
      const long cost = long(passes) * 1000000 + long(mx) * 10000 + long(rc) * 100 + (100 - ra);
      if (best_cost < 0 || cost < best_cost) {
        best_cost = cost;
        best = RowChoice{ra, rb, rc, true};
      }
    }
  }
  return best;
}

struct ColChoice {
  int h2 = 0, ra = 0, rb = 0;
  bool ok = false;
};

// H = h2 * ra * rb; h2 = 1 (one pass over the whole column) when H <= 1024 allows it, else the
// smallest largest radix, the larger factors on the strided (register-only) side.
inline ColChoice ChooseColRadices(int H) {
  ColChoice best;
  long best_cost = -1;
  for (int h2 : kRadices) {
    if (H % h2 != 0) continue;
    const int h1 = H / h2;
    if (size_t(h1) * kTile * sizeof(float2) > size_t(160) * 1024) continue;
    for (int rb : kRadices) {
      if (h1 % rb != 0) continue;
      const int ra = h1 / rb;
      if (!IsRadix(ra) || ra < rb) continue;
      if (rb == 1 && ra != h1) continue;
      const int mx = std::max(h2, ra);
      const long cost = (h2 > 1 ? 1000000 : 0) + long(mx) * 10000 + (100 - h2) * 100 + ra;
      if (best_cost < 0 || cost < best_cost) {
        best_cost = cost;
        best = ColChoice{h2, ra, rb, true};
      }
    }
  }
  return best;
}

inline std::vector<float2> TwiddleTable(int n, int count, double denominator) {
  std::vector<float2> t(count);
  for (int i = 0; i < count; ++i) {
    const double a = -2.0 * M_PI * double(i) / denominator;
    t[i] = make_float2(float(std::cos(a)), float(std::sin(a)));
  }
  (void)n;
  return t;
}

// The split pass works on pairs of last-pass butterflies that hold each other's mirror
// frequencies: block (ka, kb) and block ((ra - ka) % ra, ka ? rb - 1 - kb : (rb - kb) % rb).
// Entry 0 is block (0, 0), whose pairing runs inside the block. Ordered by k_low = ka + ra * kb so
// that neighbouring threads write neighbouring frequencies.
inline std::vector<uint4> MakePairTable(int ra, int rb, int rc, int pitch_l2) {
  std::vector<uint4> pairs;
  auto offset = [&](int ka, int kb) { return uint32_t(ka * pitch_l2 + kb * rc); };
  for (int klow = 0; klow < ra * rb; ++klow) {
    const int ka = klow % ra, kb = klow / ra;
    const int pa = (ra - ka) % ra, pb = ka ? rb - 1 - kb : (rb - kb) % rb;
    const int plow = pa + ra * pb;
    if (klow > plow) continue;
    pairs.push_back(make_uint4(offset(ka, kb), offset(pa, pb), uint32_t(klow), uint32_t(plow)));
  }
  return pairs;
}

inline int RoundUp32(int v) { return (v + 31) / 32 * 32; }

// Everything a W x H frame needs, on the host: radices, launch geometry, tables.
struct HostFastPlan {
  bool ok = false;
  int W = 0, H = 0, n = 0, pitch = 0;
  RowChoice row;
  ColChoice col;
  int row_threads = 128, rows_per_cta = 1, row_stride = 0, pitch_l2 = 0;
  size_t row_smem = 0;
  int block_threads = 128;
  size_t block_smem = 0;
  std::vector<float2> tw_row1, tw_row2, tw_split, tw_h1, tw_h;
  std::vector<uint4> pairs;
  // whole columns in shared memory (WholePlan): radices of H, tile width, tables
  bool whole_ok = false;
  RowChoice whole;
  int whole_tw = 0, whole_threads = 512;
  size_t whole_smem = 0;
  std::vector<float2> tw_whole_a, tw_whole_b;

  void Build(int width, int height) {
    ok = false;
    W = width;
    H = height;
    if (W < 2 || W % 2 != 0 || H < 1) return;
    n = W / 2;
    pitch = (n + 1 + kTile - 1) / kTile * kTile;
    row = ChooseRowRadices(n);
    col = ChooseColRadices(H);
    if (!row.ok || !col.ok) return;
    const int l2 = row.rb * row.rc;
    pitch_l2 = l2 + ((l2 % 2 == 0 && row.ra > 1) ? 1 : 0);
    row_stride = row.ra * pitch_l2;
    pairs = MakePairTable(row.ra, row.rb, row.rc, pitch_l2);
    // short rows: several per CTA so that every pass has about 128 tasks
    const int min_tasks = std::max(1, std::min(l2, std::min(row.ra * row.rc, int(pairs.size()))));
    rows_per_cta = std::clamp(128 / min_tasks, 1, 16);
    while (rows_per_cta > 1 && size_t(rows_per_cta) * row_stride * sizeof(float2) > size_t(48) * 1024)
      --rows_per_cta;
    row_smem = size_t(rows_per_cta) * row_stride * sizeof(float2);
    if (row_smem > size_t(200) * 1024) return;
    const int max_tasks = rows_per_cta * std::max(l2, std::max(row.ra * row.rc, int(pairs.size())));
    row_threads = std::clamp(RoundUp32(max_tasks), 64, 128);
    // stage twiddles by (q, j): see RowPlan
    auto stage_table = [](int radix, int m, int length) {
      std::vector<float2> t(size_t(std::max(radix - 1, 0)) * m);
      for (int q = 1; q < radix; ++q)
        for (int j = 0; j < m; ++j) {
          const double a = -2.0 * M_PI * double(j) * double(q) / double(length);
          t[size_t(q - 1) * m + j] = make_float2(float(std::cos(a)), float(std::sin(a)));
        }
      if (t.empty()) t.push_back(make_float2(1.0f, 0.0f));
      return t;
    };
    tw_row1 = stage_table(row.ra, l2, n);
    tw_row2 = stage_table(row.rb, row.rc, l2);
    tw_split = TwiddleTable(n, n + 1, double(W));
    const int h1 = H / col.h2;
    tw_h1 = TwiddleTable(h1, h1, double(h1));
    tw_h = TwiddleTable(H, H, double(H));
    block_smem = size_t(h1) * kTile * sizeof(float2);
    block_threads = std::clamp(RoundUp32(std::max(col.ra, col.rb) * kTile), 64, 256);
    ok = true;
    // one kernel for the whole column transform where a 4- or 2-column tile of the full height
    // fits in shared memory (three passes worth it from H = 2048 on)
    whole = ChooseRowRadices(H);
    whole_ok = false;
    if (whole.ok && whole.ra > 1 && whole.rb > 1 && whole.rc > 1 && H >= 2048) {
      for (int tw : {4, 2}) {
        const size_t bytes = size_t(H + H / whole.rc) * tw * sizeof(float2);
        if (bytes <= size_t(200) * 1024 && pitch % tw == 0) {
          whole_tw = tw;
          whole_smem = bytes;
          whole_ok = true;
          break;
        }
      }
    }
    if (whole_ok) {
      tw_whole_a = stage_table(whole.ra, whole.rb * whole.rc, H);
      tw_whole_b = stage_table(whole.rb, whole.rc, whole.rb * whole.rc);
    }
  }

  WholePlan MakeWholePlan(const float2* d_twa, const float2* d_twb) const {
    WholePlan p{};
    p.H = H;
    p.ra = whole.ra;
    p.rb = whole.rb;
    p.rc = whole.rc;
    p.l2 = whole.rb * whole.rc;
    p.tw = whole_tw;
    p.twa = d_twa;
    p.twb = d_twb;
    return p;
  }

  RowPlan MakeRowPlan(const float2* d_tw1, const float2* d_tw2, const float2* d_split,
                      const uint4* d_pairs) const {
    RowPlan p{};
    p.n = n;
    p.ra = row.ra;
    p.rb = row.rb;
    p.rc = row.rc;
    p.l2 = row.rb * row.rc;
    p.pitch_l2 = pitch_l2;
    p.n_pairs = int(pairs.size());
    p.rows_per_cta = rows_per_cta;
    p.row_stride = row_stride;
    p.tw1 = d_tw1;
    p.tw2 = d_tw2;
    p.split_tw = d_split;
    p.pairs = d_pairs;
    return p;
  }

  ColPlan MakeColPlan(const float2* d_tw1, const float2* d_tw_h) const {
    ColPlan p{};
    p.H = H;
    p.H2 = col.h2;
    p.H1 = H / col.h2;
    p.ra = col.ra;
    p.rb = col.rb;
    p.tw1 = d_tw1;
    p.tw_h = d_tw_h;
    return p;
  }
};

}  // namespace fast
}  // namespace rdl
#endif
