// See components.h. The reference merges through one W x H image per frequency; sorting the
// components by position gives the same sums (same order of additions) and the same output
// order without those images.
#include "components.h"

#include <algorithm>
#include <numeric>

namespace rdl {

void ComponentList::Add(size_t x, size_t y, size_t scale_index, const float* v) {
  ComponentScaleList& list = scales[scale_index];
  list.x.push_back(uint32_t(x));
  list.y.push_back(uint32_t(y));
  list.values.insert(list.values.end(), v, v + n_values);
  merged = false;
}

void ComponentList::MergeDuplicates() {
  if (merged) return;
  for (ComponentScaleList& list : scales) {
    const size_t n = list.x.size();
    std::vector<size_t> order(n);
    std::iota(order.begin(), order.end(), size_t(0));
    auto position = [&](size_t i) { return size_t(list.y[i]) * width + list.x[i]; };
    std::stable_sort(order.begin(), order.end(),
                     [&](size_t a, size_t b) { return position(a) < position(b); });
    // sums per position, row-major
    std::vector<uint32_t> sx, sy;
    std::vector<float> sums;
    for (size_t k = 0; k != n;) {
      const size_t pos = position(order[k]);
      const size_t at = sums.size();
      sums.insert(sums.end(), n_values, 0.0f);
      for (; k != n && position(order[k]) == pos; ++k)
        for (size_t f = 0; f != n_values; ++f) sums[at + f] += list.values[order[k] * n_values + f];
      sx.push_back(uint32_t(pos % width));
      sy.push_back(uint32_t(pos / width));
    }
    // output: for every image in turn, the positions whose sum in that image is non-zero and
    // that were not emitted for an earlier image
    ComponentScaleList out;
    std::vector<uint8_t> emitted(sx.size(), 0);
    for (size_t image = 0; image != n_values; ++image)
      for (size_t p = 0; p != sx.size(); ++p)
        if (!emitted[p] && sums[p * n_values + image] != 0.0f) {
          emitted[p] = 1;
          out.x.push_back(sx[p]);
          out.y.push_back(sy[p]);
          out.values.insert(out.values.end(), sums.begin() + p * n_values,
                            sums.begin() + (p + 1) * n_values);
        }
    list = std::move(out);
  }
  merged = true;
}

}  // namespace rdl
