// Kernel-side declarations for the Hogbom / GenericClean path.
#ifndef RDL_CLEAN_KERNELS_CUH_
#define RDL_CLEAN_KERNELS_CUH_

#include "common.cuh"

namespace rdl {

// Search window of a peak scan (peak_finder.cc:28-33), plus optional mask / RMS factor.
struct PeakWindow {
  int x0, x1, y0, y1;        // [x0,x1) x [y0,y1)
  const uint8_t* mask;       // device, W*H bools or null
  const float* rms_factor;   // device, W*H or null (generic_clean.cc:259-265)
  int allow_negative;
};

inline PeakWindow MakePeakWindow(size_t width, size_t height, size_t start_y, size_t end_y,
                                 size_t h_border, size_t v_border, const uint8_t* mask,
                                 const float* rms, bool allow_negative) {
  size_t xs = h_border, xe = width - h_border;
  size_t ys = start_y > v_border ? start_y : v_border;
  size_t ye = end_y < height - v_border ? end_y : height - v_border;
  if (h_border > width || xe < xs) xe = xs;
  if (v_border > height || ye < ys) ye = ys;
  PeakWindow w;
  w.x0 = int(xs);
  w.x1 = int(xe);
  w.y0 = int(ys);
  w.y1 = int(ye);
  w.mask = mask;
  w.rms_factor = rms;
  w.allow_negative = allow_negative ? 1 : 0;
  return w;
}

// Device-resident control block of a clean loop (one per handle).
struct CleanState {
  unsigned long long key[2];  // double-buffered arg-max accumulator
  unsigned int ticket;
  int slot;                   // which key[] the next launch reduces into
  int done;                   // 0 = loop still running
  int has_peak;
  float max_value;            // signed integrated value at the current peak
  int has_start;
  float start_value;
  float initial_max, first_threshold, major_iter_threshold;
  int diverging;
  int px, py;                 // position of the component being subtracted
  unsigned long long iteration;
  unsigned long long n_trace;
  float factors[kMaxImages];  // per image: fitted peak value * gain
};

// Row-sharded (multi-GPU) Hogbom: what one rank contributes to the joined arg-max — its best
// key (global row-major index, so the lowest index still wins ties across ranks) and the
// per-image residual values at that pixel, which only the owning rank holds.
struct DistCandidate {
  unsigned long long key;
  float values[kMaxImages];
};

struct TraceEntry {  // == rdl_trace_entry
  uint32_t iteration, scale_index, x, y;
  float value;
};

struct HogbomParams {
  PtrTable img;         // residual images
  PtrTable model;       // model images
  ConstPtrTable psf;    // psf of image i (already mapped through PsfIndex)
  IntegSpec first_spec; // GetLinearIntegrated (generic_clean.cc:71)
  IntegSpec step_spec;  // GetSquareIntegrated (generic_clean.cc:198)
  PeakWindow window;
  int n_images;
  int width, height;
  int simple_first, simple_step;  // every image is exactly one integration term, in order
  // settings (deconvolution_algorithm.h:189-202)
  float threshold, major_iteration_threshold, gain, mgain, divergence_limit;
  int stop_on_negative;
  unsigned long long max_iterations;
  // spectral fit (deconvolution_algorithm.cc:29-46)
  const double* projection;  // device, n_fit x n_fit row-major, or null
  int n_fit, n_pol;
  TraceEntry* trace;
  unsigned long long trace_capacity;
  // Component log (host-memory calls): instead of model[i][peak] += value, component k of the
  // call (k = iteration - log_first_iteration) is recorded as its frame index and its n_images
  // values; the host applies them to its model planes in order (the same fp32 adds), so the model
  // never crosses PCIe. Null: the model planes are updated on the device.
  uint32_t* log_index;
  float* log_values;
  unsigned long long log_first_iteration;
  // rows of a sharded cube: local row yl is frame row
  //   y = y_offset + ((yl / block_rows) * block_world + block_rank) * block_rows + yl % block_rows
  // of a width x global_height frame. Contiguous band: block_rows = height, block_world = 1;
  // cyclic row blocks (balances the PSF-window work over the ranks): y_offset = 0,
  // block_world = world. PSFs, mask, RMS image, borders and peak indices are the frame's.
  int y_offset, global_height;
  int block_rows, block_world, block_rank;
  int dist;                   // 1: the last block writes cand_local instead of running the control
  DistCandidate* cand_local;
};

void LaunchHogbomStep(const HogbomParams& p, CleanState* state, bool first, int vec,
                      cudaStream_t stream);

// The whole single-GPU loop in one cooperative launch, for cubes that stay in L2 (K1, facets).
// scratch: HogbomPersistentScratchBytes() of device memory (barrier words + candidate slots).
// Returns false when a cooperative launch is not possible on this device.
size_t HogbomPersistentScratchBytes();
bool LaunchHogbomPersistent(const HogbomParams& p, CleanState* state, void* scratch, int vec,
                            cudaStream_t stream);

// Joined control step of the row-sharded loop: picks the best of the gathered candidates and
// runs the same loop control as the single-GPU path (every rank redundantly, bit-identically).
// frame row of local row yl / local row of frame row y (false when another rank owns it)
__host__ __device__ inline int FrameRow(const HogbomParams& p, int yl) {
  const int b = yl / p.block_rows;
  return p.y_offset + (b * p.block_world + p.block_rank) * p.block_rows + (yl - b * p.block_rows);
}
__host__ __device__ inline bool LocalRow(const HogbomParams& p, int y, int& yl) {
  const int yy = y - p.y_offset;
  if (yy < 0) return false;
  const int b = yy / p.block_rows;
  if (b % p.block_world != p.block_rank) return false;
  yl = (b / p.block_world) * p.block_rows + (yy - b * p.block_rows);
  return yl < p.height;
}

void LaunchHogbomDistControl(const HogbomParams& p, CleanState* state, const DistCandidate* all,
                             int world, bool first, cudaStream_t stream);

void LaunchFindPeak(const IntegSpec& spec, const ConstPtrTable& images, int width, int height,
                    const PeakWindow& window, unsigned long long* key, int vec,
                    cudaStream_t stream);

void LaunchIntegrate(const IntegSpec& spec, const ConstPtrTable& images, size_t n, float* dest,
                     int vec, cudaStream_t stream);

void LaunchSubtract(const PtrTable& images, const ConstPtrTable& psfs, int n_images, int width,
                    int height, int x, int y, const float* factors_host, cudaStream_t stream);

}  // namespace rdl
#endif
