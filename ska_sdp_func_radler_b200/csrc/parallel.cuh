// ParallelDeconvolution (facets) on the device; see parallel.cu.
#ifndef RDL_PARALLEL_CUH_
#define RDL_PARALLEL_CUH_

#include "handle.h"

namespace rdl {

// ParallelDeconvolution::ExecuteMajorIteration for a grid_width x grid_height facet grid
// (parallel_deconvolution.cc:494-654). A 1 x 1 grid is the single-algorithm run.
rdl_result RunParallel(rdl_handle& h, DeviceCube& cube, size_t grid_width, size_t grid_height,
                       cudaStream_t stream);

struct FacetInfo {
  uint64_t x, y, width, height;
  double peak;
  int reached_major_threshold;
  uint64_t iteration_number;
};
// the one algorithm of a 1 x 1 grid run (nullptr otherwise)
rdl_handle* ParallelSingleAlgorithm(const rdl_handle& h);
size_t ParallelFacetCount(const rdl_handle& h);
FacetInfo ParallelFacet(const rdl_handle& h, size_t i);
// host copy of a facet's boundary mask (width * height bytes)
const std::vector<uint8_t>& ParallelFacetBoundaryMask(const rdl_handle& h, size_t i);
// wall-clock split of the last facet run: host divider search, whole call (milliseconds)
void ParallelTiming(const rdl_handle& h, double& geometry_ms, double& total_ms);

}  // namespace rdl
#endif
