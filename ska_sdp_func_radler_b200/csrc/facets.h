// Facet geometry of ParallelDeconvolution (host side): minimum-power dividing paths between
// facets and the per-facet bounding boxes / masks.
//   cpp/math/dijkstra_splitter.{h,cc}                      shortest dividing paths, area flooding
//   cpp/algorithms/parallel_deconvolution.cc:69-168        MakeSubImages
// This stays on the host like in the reference ("Dijkstra splitting stays on host");
// the dividers are independent and run on separate threads.
#ifndef RDL_FACETS_H_
#define RDL_FACETS_H_

#include <cstddef>
#include <cstdint>
#include <functional>
#include <vector>

namespace rdl {

struct Facet {  // parallel_deconvolution.h:22-35 (SubImage)
  size_t index = 0;
  size_t x = 0, y = 0, width = 0, height = 0;
  std::vector<uint8_t> mask;           // boundary mask AND user mask
  std::vector<uint8_t> boundary_mask;  // pixels this facet owns
  double peak = 0.0;
  bool reached_major_threshold = false;
  size_t psf_index = 0;  // NearestPsfIndex of the facet centre (direction-dependent PSFs)
};

// Marks with 1.0 the cheapest top-to-bottom (vertical) / left-to-right (horizontal) path of
// |image| whose lateral coordinate stays inside [lo, hi); the rest of that strip of `output` is
// zeroed (DijkstraSplitter::DivideVertically / DivideHorizontally).
void DividingPath(const float* image, float* output, size_t width, size_t height, size_t lo,
                  size_t hi, bool vertical);

// DijkstraSplitter::FloodVerticalArea / FloodHorizontalArea: mask = the area between the two
// dividers around `inside` (a coordinate between them); returns its start and extent.
void FloodArea(const float* dividers, size_t width, size_t height, size_t inside, bool vertical,
               uint8_t* mask, size_t& start, size_t& extent);

// DijkstraSplitter::GetBoundingMask.
void BoundingMask(const uint8_t* vertical_mask, size_t vertical_mask_x, size_t vertical_mask_width,
                  const uint8_t* horizontal_mask, uint8_t* mask, size_t width, size_t height,
                  size_t& sub_x, size_t& sub_y, size_t& sub_width, size_t& sub_height);

// Several processes that hold the same frame can split the divider searches (the expensive part
// of the geometry): rank r searches dividers k with k % world == r into otherwise zero divider
// images, and `merge` sums the two images over the ranks in place (every pixel belongs to at
// most one divider, so the sum is exact).
struct DividerSharing {
  int rank = 0, world = 1;
  std::function<void(float* vertical, float* horizontal, size_t n_pixels)> merge;
};

// MakeSubImages (parallel_deconvolution.cc:69-168). `image` = the integrated residual on the
// host; user_mask may be null. Facets are ordered row-major over the grid.
std::vector<Facet> MakeFacets(const float* image, size_t width, size_t height,
                              const uint8_t* user_mask, size_t grid_width, size_t grid_height,
                              const DividerSharing* sharing = nullptr);

}  // namespace rdl
#endif
