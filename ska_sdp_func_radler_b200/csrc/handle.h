// Host-side state of one algorithm instance (the analogue of a
// radler::algorithms::DeconvolutionAlgorithm object, deconvolution_algorithm.h:60-210).
#ifndef RDL_HANDLE_H_
#define RDL_HANDLE_H_

#include <array>
#include <map>
#include <memory>
#include <string>
#include <utility>
#include <vector>

#include "../../include/radler_b200.h"
#include "clean_kernels.cuh"
#include "common.cuh"
#include "components.h"
#include "dist.h"

namespace rdl {

// Device memory comes from a size-keyed cache (device_pool.cc): blocks of 1 MB and more are kept
// when a buffer is released and handed out again, smaller ones go to cudaMalloc / cudaFree.
void* DeviceAlloc(size_t bytes);  // throws CudaError("cudaMalloc(...)") when HBM is exhausted
void DeviceFree(void* p, size_t bytes);
void DeviceTrim();  // gives every cached block back to the driver
size_t DeviceAvailableBytes();  // free HBM of the current device, cached blocks included

// RAII device allocation.
template <typename T>
class DeviceBuffer {
 public:
  DeviceBuffer() = default;
  explicit DeviceBuffer(size_t n) { Resize(n); }
  DeviceBuffer(const DeviceBuffer&) = delete;
  DeviceBuffer& operator=(const DeviceBuffer&) = delete;
  DeviceBuffer(DeviceBuffer&& o) noexcept : ptr_(o.ptr_), n_(o.n_) { o.ptr_ = nullptr; o.n_ = 0; }
  DeviceBuffer& operator=(DeviceBuffer&& o) noexcept {
    if (this != &o) {
      Free();
      ptr_ = o.ptr_;
      n_ = o.n_;
      o.ptr_ = nullptr;
      o.n_ = 0;
    }
    return *this;
  }
  ~DeviceBuffer() { Free(); }
  void Resize(size_t n) {
    if (n == n_) return;
    Free();
    if (n) ptr_ = static_cast<T*>(DeviceAlloc(n * sizeof(T)));
    n_ = n;
  }
  // Grow-only variant for scratch space.
  void Reserve(size_t n) {
    if (n > n_) Resize(n);
  }
  void Free() {
    if (ptr_) DeviceFree(ptr_, n_ * sizeof(T));
    ptr_ = nullptr;
    n_ = 0;
  }
  T* get() const { return ptr_; }
  size_t size() const { return n_; }
  bool empty() const { return n_ == 0; }

 private:
  T* ptr_ = nullptr;
  size_t n_ = 0;
};

// The cube an ExecuteMajorIteration call works on, resident in HBM.
struct DeviceCube {
  size_t width = 0, height = 0, n_channels = 0, n_pol = 0, n_images = 0;
  std::vector<float> weights;
  std::vector<uint8_t> pol_linked;
  float pol_norm = 1.0f;
  bool squared_joins = false;
  PtrTable residual{}, model{};
  ConstPtrTable psf{};  // per channel
  // direction-dependent PSFs of a facet run (parallel_deconvolution.cc:34-55): n_dd sets of
  // n_channels planes ([dd][channel]) of psf_width x psf_height, with their positions
  std::vector<const float*> dd_psfs;
  size_t n_dd = 0, psf_width = 0, psf_height = 0;
  std::vector<int64_t> psf_offset_x, psf_offset_y;
  int vec = 1;          // 4 when width % 4 == 0 and all pointers are 16-byte aligned
  size_t PsfIndex(size_t image) const { return image / n_pol; }
  size_t Pixels() const { return width * height; }
};

// Builds the integration recipes of image_set.cc for a cube.
IntegSpec MakeLinearSpec(const DeviceCube& cube);  // GetLinearIntegrated
IntegSpec MakeSquareSpec(const DeviceCube& cube);  // GetSquareIntegrated
IntegSpec MakeIntegratedPsfSpec(const DeviceCube& cube);  // GetIntegratedPsf, over channels
bool IsSimpleSpec(const IntegSpec& spec, size_t n_images);

// Polynomial projection matrix (spectral fit restated).
std::vector<double> PolynomialProjection(size_t n_terms, const std::vector<double>& freqs,
                                         const std::vector<float>& weights);
// The t_used x n operator of the same fit that maps channel values to polynomial terms
// (projection = powers x this); t_used = min(n_terms, channels with weight).
std::vector<double> PolynomialTermSolve(size_t n_terms, const std::vector<double>& freqs,
                                        const std::vector<float>& weights, size_t& t_used);

struct Settings {  // deconvolution_algorithm.h:189-202
  float threshold = 0.0f;
  float major_iteration_threshold = 0.0f;
  float minor_loop_gain = 0.1f;
  float major_loop_gain = 1.0f;
  float clean_border_ratio = 0.05f;
  uint64_t max_iterations = 500;
  float divergence_limit = 4.0f;
  bool allow_negative_components = true;
  bool stop_on_negative_component = false;
};

class MultiScaleState;
class FftConvolver;
struct ParallelState;  // facets (parallel.cu)
struct ParallelStateDeleter {
  void operator()(ParallelState* p) const;
};

// Device buffers of the sub-minor loop (grow-only, reused between runs).
struct SubMinorBuffers {
  DeviceBuffer<uint32_t> flags, block_counts, block_offsets, positions;
  DeviceBuffer<float> residual, model, rms, candidates;
  DeviceBuffer<unsigned char> state;
  size_t n_selected = 0, capacity = 0, width = 0;
  int n_images = 0;
  uint64_t n_trace = 0;  // trace entries the last run produced
};

}  // namespace rdl

struct rdl_handle {
  rdl_config config{};
  std::vector<double> ms_scale_list;
  rdl::Settings settings;
  uint64_t iteration_number = 0;
  int device = 0;
  cudaStream_t own_stream = nullptr;
  // host-memory multiscale calls: the model planes are uploaded on a second stream behind the
  // residual and PSF planes, while the call already works; model_ready is set (non-null) for the
  // duration of such a call and the algorithm waits on it before it first touches the model
  cudaStream_t upload_stream = nullptr;
  cudaEvent_t upload_event = nullptr, model_ready = nullptr;
  bool model_pending = false;
  std::string error;
  uint64_t launches = 0;

  // clean mask / rms factor image (device copies) + host copies for Clone()
  rdl::DeviceBuffer<uint8_t> mask;
  std::vector<uint8_t> mask_host;
  size_t mask_w = 0, mask_h = 0;
  rdl::DeviceBuffer<float> rms;
  std::vector<float> rms_host;
  size_t rms_w = 0, rms_h = 0;
  // bumped by the setters; facet clones remember the parent's value they were last synced to
  uint64_t mask_generation = 0, rms_generation = 0, fit_generation = 0;
  uint64_t synced_mask = 0, synced_rms = 0, synced_fit = 0;

  // spectral fitter
  int32_t fit_mode = RDL_FIT_NONE;
  uint64_t fit_terms = 0, fit_n_pol = 1;
  std::vector<double> fit_freqs;
  std::vector<float> fit_weights;
  std::vector<double> projection_host;
  rdl::DeviceBuffer<double> projection;
  rdl::DeviceBuffer<double> interpolation;  // InterpolateAndStoreModel: term solve + powers

  // auto-mask mode (multiscale)
  bool track_per_scale_masks = false, use_per_scale_masks = false;

  // trace
  uint64_t trace_capacity = 0;
  rdl::DeviceBuffer<rdl::TraceEntry> trace_dev;
  std::vector<rdl::TraceEntry> trace_host;

  // loop control block + scratch
  rdl::DeviceBuffer<rdl::CleanState> state;
  rdl::DeviceBuffer<unsigned long long> key;
  uint64_t last_selected = 0;
  uint64_t minor_iterations = 0;
  // schaapcommon::math::Convolve calls the REFERENCE would make for the work of the last call,
  // by frame size (the device path caches spectra and twice-convolved PSFs, so it runs fewer):
  // bench.py prices the CPU side with it
  std::map<std::pair<size_t, size_t>, uint64_t> reference_convolutions;

  // staging for RDL_MEM_HOST calls
  rdl::DeviceBuffer<float> stage_residual, stage_model, stage_psf;
  // rdl_set_psfs_constant: the PSF planes staged by the previous host-memory call are reused when
  // the caller passes the same pointers again
  bool psfs_constant = false;
  std::vector<const float*> staged_psf_ptrs;
  size_t staged_psf_pixels = 0;
  // rdl_set_pin_host_buffers: caller planes registered with cudaHostRegister (pointer -> bytes)
  bool pin_host_buffers = false;
  std::map<const void*, size_t> pinned;

  rdl::SubMinorBuffers sub;
  // FFT scratch: one W x H image, one kernel spectrum, one uploaded scale kernel
  rdl::DeviceBuffer<float> fft_scratch, small_kernel;
  rdl::DeviceBuffer<float2> fft_spectrum;
  rdl::DeviceBuffer<float> iuwt_a, iuwt_b;  // IUWT ping-pong planes
  rdl::DeviceBuffer<uint8_t> iuwt_state;    // flood-fill state volume (iuwt_analysis.cu)
  rdl::DeviceBuffer<int> iuwt_due;          // flood fill: round in which a tile has to run
  rdl::DeviceBuffer<unsigned long long> iuwt_scalars;  // counters, histogram, arg-max key
  // facet run state: per-facet clones and geometry (not cloned)
  std::unique_ptr<rdl::ParallelState, rdl::ParallelStateDeleter> parallel;
  // row-sharded multi-GPU loop (not cloned)
  rdl::DistContext dist;
  rdl::DeviceBuffer<rdl::DistCandidate> cand_local, cand_all;
  rdl::DeviceBuffer<unsigned char> persist_scratch;  // HogbomPersistentKernel: barrier + candidates
  // local-RMS maps (rms_image.cu)
  rdl::DeviceBuffer<float> rms_temp, rms_kernel, rms_min;
  rdl::DeviceBuffer<unsigned> rms_scalar;
  // component log of a host-memory GenericClean call (clean_kernels.cuh, HogbomParams::log_*)
  bool model_log = false;        // set by the caller of RunGenericClean for this call
  uint64_t model_log_count = 0;  // components recorded by the call
  rdl::DeviceBuffer<uint32_t> log_index;
  rdl::DeviceBuffer<float> log_values;
  std::vector<uint32_t> log_index_host;
  std::vector<float> log_values_host;
  rdl::ComponentList components;     // MultiScaleAlgorithm::component_list_ (track_components)
  double beam_size_in_pixels = 1.0;  // multiscale_algorithm.cc:153-163
  std::shared_ptr<rdl::MultiScaleState> ms;  // deep-copied by rdl_clone
  // FFT engines by frame size (plans, twiddle tables, work spectrum); not cloned
  std::map<std::array<size_t, 4>, std::shared_ptr<rdl::FftConvolver>> convolvers;
};

namespace rdl {
rdl_result RunGenericClean(rdl_handle& h, DeviceCube& cube, cudaStream_t stream);
rdl_result RunMultiScale(rdl_handle& h, DeviceCube& cube, cudaStream_t stream);
rdl_result RunIuwt(rdl_handle& h, DeviceCube& cube, cudaStream_t stream);
}  // namespace rdl

#endif
