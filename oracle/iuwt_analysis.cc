// ORACLE — TEST INFRASTRUCTURE ONLY (never linked into or called by the product).
// CPU restatement of the reference's IUWT structure selection and of the scalar measurements
// the IUWT minor loop takes from a coefficient plane:
//   cpp/algorithms/iuwt/image_analysis.cc:9-19    ExceedsThreshold / ExceedsThresholdAbs
//   cpp/algorithms/iuwt/image_analysis.cc:86-149  Floodfill          (6-connected over scale,y,x)
//   cpp/algorithms/iuwt/image_analysis.cc:151-206 MaskedFloodfill    (same, inside a prior mask)
//   cpp/algorithms/iuwt/image_analysis.cc:208-239 SelectStructures   (fill from every seed)
//   cpp/algorithms/iuwt/image_analysis.cc:241-333 FloodFill2D        (signed and abs flavours)
//   cpp/algorithms/iuwt_deconvolution_algorithm.cc:104-110 Mad
//   cpp/algorithms/iuwt_deconvolution_algorithm.cc:112-167 GetMaxAbsWithoutMask / WithMask
// The walk is the reference's: a LIFO of voxels, a voxel is marked when it is pushed, the moves
// out of a voxel are bounded by the position of the voxel we LEAVE (x > min_x to go left,
// x < max_x - 1 to go right, ...), so a seed outside the border can still spread along the other
// axes. Pinned against the reference's own image_analysis.cc (oracle/_ref) by
// tests/test_iuwt_analysis_oracle.py.
#include <algorithm>
#include <cmath>
#include <limits>
#include <vector>

#include "oracle.h"

namespace oracle {
namespace {

bool Exceeds(float v, float threshold) {
  if (threshold >= 0.0f) return v > threshold;
  return v < threshold || v > -threshold;
}

struct Voxel {
  size_t x, y;
  int scale;
};

struct Border {
  size_t min_x, max_x, min_y, max_y;
  Border(float clean_border, size_t width, size_t height) {
    const size_t bx = clean_border * width, by = clean_border * height;
    min_x = bx;
    max_x = width - bx;
    min_y = by;
    max_y = height - by;
  }
};

}  // namespace

size_t IuwtFloodfill(const float* coefficients, uint8_t* masks, size_t width, size_t height,
                     int n_scales, const float* thresholds, int min_scale, int end_scale,
                     size_t seed_x, size_t seed_y, int seed_scale, float clean_border,
                     const uint8_t* prior_mask) {
  const size_t area = width * height;
  const Border b(clean_border, width, height);
  end_scale = std::min(end_scale, n_scales);
  std::vector<Voxel> todo{Voxel{seed_x, seed_y, seed_scale}};
  masks[size_t(seed_scale) * area + seed_x + seed_y * width] = 1;
  size_t filled = 0;
  auto visit = [&](size_t x, size_t y, int scale) {
    const size_t px = x + y * width, vox = size_t(scale) * area + px;
    if (Exceeds(coefficients[vox], thresholds[scale]) && !masks[vox] &&
        (!prior_mask || prior_mask[px])) {
      masks[vox] = 1;
      todo.push_back(Voxel{x, y, scale});
    }
  };
  while (!todo.empty()) {
    const Voxel c = todo.back();
    todo.pop_back();
    ++filled;
    if (c.x > b.min_x) visit(c.x - 1, c.y, c.scale);
    if (c.x < b.max_x - 1) visit(c.x + 1, c.y, c.scale);
    if (c.y > b.min_y) visit(c.x, c.y - 1, c.scale);
    if (c.y < b.max_y - 1) visit(c.x, c.y + 1, c.scale);
    if (c.scale > min_scale) visit(c.x, c.y, c.scale - 1);
    if (c.scale < end_scale - 1) visit(c.x, c.y, c.scale + 1);
  }
  return filled;
}

size_t IuwtSelectStructures(const float* coefficients, uint8_t* masks, size_t width, size_t height,
                            int n_scales, const float* thresholds, int min_scale, int end_scale,
                            float clean_border, const uint8_t* prior_mask) {
  const size_t area = width * height;
  const Border b(clean_border, width, height);
  size_t total = 0;
  for (int scale = min_scale; scale != end_scale; ++scale)
    for (size_t y = b.min_y; y != b.max_y; ++y)
      for (size_t x = b.min_x; x != b.max_x; ++x) {
        const size_t px = x + y * width, vox = size_t(scale) * area + px;
        if (Exceeds(coefficients[vox], thresholds[scale]) && !masks[vox] &&
            (!prior_mask || prior_mask[px]))
          total += IuwtFloodfill(coefficients, masks, width, height, n_scales, thresholds,
                                 min_scale, end_scale, x, y, scale, clean_border, prior_mask);
      }
  return total;
}

size_t IuwtFloodfill2D(const float* image, uint8_t* mask, size_t width, size_t height,
                       float threshold, bool use_abs, size_t seed_x, size_t seed_y) {
  std::vector<Voxel> todo{Voxel{seed_x, seed_y, 0}};
  mask[seed_x + seed_y * width] = 1;
  size_t filled = 0;
  auto visit = [&](size_t x, size_t y) {
    const size_t px = x + y * width;
    const bool over = use_abs ? std::fabs(image[px]) > threshold : Exceeds(image[px], threshold);
    if (over && !mask[px]) {
      mask[px] = 1;
      todo.push_back(Voxel{x, y, 0});
    }
  };
  while (!todo.empty()) {
    const Voxel c = todo.back();
    todo.pop_back();
    ++filled;
    if (c.x > 0) visit(c.x - 1, c.y);
    if (c.x < width - 1) visit(c.x + 1, c.y);
    if (c.y > 0) visit(c.x, c.y - 1);
    if (c.y < height - 1) visit(c.x, c.y + 1);
  }
  return filled;
}

float IuwtMad(const float* image, size_t n) {
  std::vector<float> v(n);
  for (size_t i = 0; i != n; ++i) v[i] = std::fabs(image[i]);
  const size_t mid = n / 2;
  std::nth_element(v.begin(), v.begin() + mid, v.end());
  return v[mid] / 0.674559;
}

float IuwtMaxAbs(const float* image, size_t width, size_t height, float clean_border,
                 bool allow_negative, const uint8_t* mask, size_t& x, size_t& y) {
  const Border b(clean_border, width, height);
  x = width;
  y = height;
  float best = std::numeric_limits<float>::lowest();
  for (size_t yi = b.min_y; yi != b.max_y; ++yi)
    for (size_t xi = b.min_x; xi != b.max_x; ++xi) {
      if (mask && !mask[yi * width + xi]) continue;
      const float v = image[yi * width + xi];
      const float val = allow_negative ? std::fabs(v) : v;
      if (val > best) {
        best = val;
        x = xi;
        y = yi;
      }
    }
  return best;
}

}  // namespace oracle
