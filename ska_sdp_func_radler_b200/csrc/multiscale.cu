// MultiScaleAlgorithm on the device (algorithms/multiscale_algorithm.{h,cc},
// algorithms/multiscale/multiscale_transforms.{h,cc}, threaded_deconvolution_tools.cc:30-107).
//
// The host keeps the reference's scalar state machine (scale list, thresholds, scale
// selection, activation) and drives device-resident work: every image, PSF, convolved PSF
// and scale-kernel spectrum lives in HBM for the whole call; the only host traffic per
// minor iteration is a handful of 8-byte peak records and the sub-minor loop's result.
//
// Several GPUs (rdl_dist_init on the handle, no row band): the CHANNELS are sharded. Every rank
// holds the whole cube; rank r owns the contiguous channel block [r P / world, (r + 1) P / world)
// and runs all per-image FFT convolutions of those channels (convolved / twice-convolved PSFs,
// scale-convolved residuals, CorrectResidualDirty, the model convolution:
// multiscale_algorithm.cc:331-354, 432-461) -- 90+ % of a major iteration. What a joined step
// needs from the other channels travels as whole planes, broadcast by their owners over NCCL /
// NVLink (the scale-convolved residuals before the sub-minor loop, the corrected residuals after
// it, the twice-convolved PSFs once per scale, the model at the end). Everything joined --
// integration over the channels in the reference's sequential order, scale selection, pixel
// selection and the sub-minor loop on the compact sets -- is computed by every rank on identical
// data, so the ranks take identical decisions without exchanging them, and the result is
// bit-identical to the single-GPU run.
#include "multiscale.cuh"

#include <algorithm>
#include <cmath>
#include <map>
#include <set>

#include "fft.cuh"
#include "subminor.cuh"

namespace rdl {

// ---- scale kernels: multiscale_transforms.h:91-195 (host, same float/double mix) ------------
namespace {

float HannWindowFunction(float x, size_t n) {
  return (x * 2 <= n + 1) ? float(0.5 * (1.0 + std::cos(2.0 * M_PI * x / double(n + 1)))) : 0.0f;
}
float ShapeFunction(float x) {
  // `1.0 - x * x` with float x: the reference build (GCC, -ffp-contract=fast) evaluates the
  // product exactly inside a double fnmadd; verified against oracle/_ref bit for bit.
  return (x < 1.0) ? float(std::fma(-double(x), double(x), 1.0)) : 0.0f;
}

std::vector<float> MakeTaperedQuadratic(double scale, size_t& n) {
  n = size_t(std::ceil(scale * 0.5) * 2.0) + 1;
  std::vector<float> out(n * n, 0.0f);
  if (scale == 0.0) {
    out[0] = 1.0f;
    return out;
  }
  float sum = 0.0f;
  float* p = out.data();
  for (int y = 0; y != int(n); ++y) {
    const float dy = y - 0.5 * (n - 1);
    const float dydy = dy * dy;
    for (int x = 0; x != int(n); ++x) {
      const float dx = x - 0.5 * (n - 1);
      const float r = std::sqrt(dx * dx + dydy);
      *p = HannWindowFunction(r, n) * ShapeFunction(r / scale);
      sum += *p;
      ++p;
    }
  }
  const float norm = 1.0 / sum;
  for (float& v : out) v *= norm;
  return out;
}

std::vector<float> MakeGaussian(double scale, size_t& n, size_t max_n) {
  float sigma = float(scale) * float(3.0 / 16.0);
  n = int(std::ceil(sigma * 12.0 / 2.0)) * 2 + 1;
  if (n > max_n) {
    n = max_n;
    if ((n % 2) == 0 && n > 0) --n;
  }
  if (n < 1) n = 1;
  if (sigma == 0.0) {
    sigma = 1.0;
    n = 1;
  }
  std::vector<float> out(n * n), gaus(n);
  const float mu = int(n / 2);
  const float two_sigma_squared = 2.0 * sigma * sigma;
  for (int i = 0; i != int(n); ++i) {
    const float v = float(i) - mu;
    gaus[i] = std::exp(-v * v / two_sigma_squared);
  }
  float sum = 0.0f;
  float* p = out.data();
  for (int y = 0; y != int(n); ++y)
    for (int x = 0; x != int(n); ++x) {
      *p = gaus[x] * gaus[y];
      sum += *p;
      ++p;
    }
  const float norm = 1.0 / sum;
  for (float& v : out) v *= norm;
  return out;
}

std::vector<float> MakeShapeFunction(float scale, size_t& n, size_t max_n, int shape) {
  if (shape == RDL_SHAPE_GAUSSIAN) return MakeGaussian(scale, n, max_n);
  return MakeTaperedQuadratic(scale, n);
}

float KernelPeakValue(double scale, size_t max_n, int shape) {  // multiscale_transforms.h:56-60
  size_t n;
  const std::vector<float> k = MakeShapeFunction(scale, n, max_n, shape);
  return k[n / 2 + (n / 2) * n];
}

// InitializeScales, multiscale_algorithm.cc:90-131
void InitializeScales(std::vector<rdl_scale_info>& scales, double beam_size_in_pixels,
                      size_t min_width_height, int shape, size_t max_scales,
                      const std::vector<double>& scale_list) {
  if (scale_list.empty()) {
    if (scales.empty()) {
      size_t scale_index = 0;
      double scale = beam_size_in_pixels * 2.0;
      do {
        scales.emplace_back();
        rdl_scale_info& e = scales.back();
        e = rdl_scale_info{};
        e.scale = scale_index == 0 ? 0.0f : float(scale);
        e.kernel_peak = KernelPeakValue(scale, min_width_height, shape);
        scale *= 2.0;
        ++scale_index;
      } while (scale < min_width_height * 0.5 && (max_scales == 0 || scale_index < max_scales));
    } else {
      while (!scales.empty() && scales.back().scale >= min_width_height * 0.5) scales.pop_back();
    }
  } else if (scales.empty()) {
    std::multiset<double> sorted(scale_list.begin(), scale_list.end());
    for (double s : sorted) {
      scales.emplace_back();
      rdl_scale_info& e = scales.back();
      e = rdl_scale_info{};
      e.scale = float(s);
      e.kernel_peak = KernelPeakValue(e.scale, min_width_height, shape);
    }
  }
}

// sum of squares in double; the reference's sequential fp32 sum (threaded_deconvolution_tools.h
// :40-44) only feeds a log line, so this value is tolerance-level, not bit-level.
__global__ void SumSquaresKernel(const float* image, size_t n, double* out) {
  __shared__ double red[32];
  double acc = 0.0;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    acc += double(image[i]) * double(image[i]);
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    acc = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.0;
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (threadIdx.x == 0) atomicAdd(out, acc);
  }
}

}  // namespace

// Everything a call derives from the PSFs alone: kept from one ExecuteMajorIteration to the next
// when the caller has promised that the PSFs do not change (rdl_set_psfs_constant; radler.cc:175-268
// passes the same psf_images in every major iteration, and the reference recomputes all of this
// each time). The key is what the contents depend on besides the PSF pixels.
struct MultiScalePsfCache {
  bool valid = false;
  std::vector<const float*> psf_ptrs;
  std::vector<float> weights, scale_sizes;
  size_t width = 0, height = 0, n_pol = 0;
  int shape = 0;
  double padding = 0.0;
  bool sharded = false;
  std::vector<float> psf_peak;  // of the integrated PSF per scale
  DeviceBuffer<float> integrated_psf;
  std::vector<DeviceBuffer<float>> convolved_psfs;  // [psf][scale], empty for scale 0
  std::vector<DeviceBuffer<float2>> scale_spectra;
  std::map<std::pair<size_t, size_t>, DeviceBuffer<float2>> psf_spectra;
  std::map<std::pair<size_t, size_t>, DeviceBuffer<float>> twice;  // [psf, scale]
  void Clear() {
    valid = false;
    psf_peak.clear();
    convolved_psfs.clear();
    scale_spectra.clear();
    psf_spectra.clear();
    twice.clear();
  }
};

class MultiScaleState {
 public:
  std::vector<rdl_scale_info> scales;
  std::vector<DeviceBuffer<uint8_t>> scale_masks;
  size_t mask_pixels = 0;
  MultiScalePsfCache psf_cache;  // not cloned: a clone fills its own
};

void MultiScaleInvalidatePsfCache(MultiScaleState& s) { s.psf_cache.Clear(); }

std::shared_ptr<MultiScaleState> CloneMultiScaleState(const MultiScaleState& src, cudaStream_t stream) {
  auto dst = std::make_shared<MultiScaleState>();
  dst->scales = src.scales;
  dst->mask_pixels = src.mask_pixels;
  for (const auto& m : src.scale_masks) {
    dst->scale_masks.emplace_back(m.size());
    if (m.size())
      RDL_CUDA(cudaMemcpyAsync(dst->scale_masks.back().get(), m.get(), m.size(), cudaMemcpyDeviceToDevice, stream));
  }
  RDL_CUDA(cudaStreamSynchronize(stream));
  return dst;
}

uint64_t MultiScaleScaleCount(const MultiScaleState& s) { return s.scales.size(); }
rdl_scale_info MultiScaleScaleInfo(const MultiScaleState& s, uint64_t i) { return s.scales[i]; }

std::vector<DeviceBuffer<uint8_t>>& MultiScaleMasks(rdl_handle& h, size_t mask_pixels) {
  if (!h.ms) h.ms = std::make_shared<MultiScaleState>();
  h.ms->mask_pixels = mask_pixels;
  return h.ms->scale_masks;
}

void MultiScaleCopyScaleMask(const MultiScaleState& s, uint64_t i, uint8_t* out_host) {
  if (i >= s.scale_masks.size()) throw std::invalid_argument("scale mask index out of range");
  RDL_CUDA(cudaMemcpy(out_host, s.scale_masks[i].get(), s.scale_masks[i].size(), cudaMemcpyDeviceToHost));
}

namespace {

// Uploads the n x n scale kernel and builds its spectrum for the convolver's frame.
void MakeScaleSpectrum(rdl_handle& h, FftConvolver& conv, float scale, int shape, float2* spectrum,
                       cudaStream_t stream) {
  size_t n = 0;
  const std::vector<float> kernel =
      MakeShapeFunction(scale, n, std::min(conv.Width(), conv.Height()), shape);
  h.small_kernel.Reserve(kernel.size());
  RDL_CUDA(cudaMemcpyAsync(h.small_kernel.get(), kernel.data(), kernel.size() * sizeof(float),
                           cudaMemcpyHostToDevice, stream));
  RDL_CUDA(cudaStreamSynchronize(stream));  // `kernel` is pageable host memory
  conv.MakeSmallKernelSpectrum(h.small_kernel.get(), n, spectrum, stream);
}

// values[i] = images[i][index]  (MeasureComponentValues, multiscale_algorithm.cc:658-674)
__global__ void GatherPixelKernel(ConstPtrTable images, int n_images, size_t index, float* values) {
  const int i = threadIdx.x;
  if (i < n_images) values[i] = images.p[i][index];
}

struct ComponentValues {
  float v[kMaxImages];
};

// AddComponentToModel for a non-zero scale (multiscale_transforms.h:62-89): the n x n shape
// function centred on (x, y), clipped at the frame, row = fma(kernel, gain, row).
__global__ void AddShapeKernel(PtrTable model, int n_images, size_t width, size_t height,
                               const float* __restrict__ shape, size_t n, size_t x, size_t y,
                               ComponentValues gains) {
  const size_t left = x > n / 2 ? x - n / 2 : 0, top = y > n / 2 ? y - n / 2 : 0;
  const size_t right = min(x + (n + 1) / 2, width), bottom = min(y + (n + 1) / 2, height);
  const size_t bw = right - left, bh = bottom - top;
  const int image = blockIdx.y;
  float* data = model.p[image];
  const float gain = gains.v[image];
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < bw * bh;
       i += size_t(gridDim.x) * blockDim.x) {
    const size_t yi = top + i / bw, xi = left + i % bw;
    const float k = shape[(yi + n / 2 - y) * n + xi + n / 2 - x];
    float* d = data + yi * width + xi;
    *d = __fmaf_rn(k, gain, *d);
  }
}

// model[i][index] += values[i]  (scale 0)
__global__ void AddPointKernel(PtrTable model, int n_images, size_t index, ComponentValues gains) {
  const int i = threadIdx.x;
  if (i < n_images) model.p[i][index] = __fadd_rn(model.p[i][index], gains.v[i]);
}

struct Peak {
  bool has = false;
  float value = 0.0f;
  size_t x = 0, y = 0;
};

}  // namespace

// MultiScaleTransforms::Transform (multiscale_transforms.cc:11-23): circular, unpadded.
void MultiScaleTransform(rdl_handle& h, float* const* images, uint64_t n_images, uint64_t width,
                         uint64_t height, float scale, int32_t shape, cudaStream_t stream) {
  FftConvolver& conv = GetConvolver(h, width, height);
  h.fft_spectrum.Reserve(conv.SpectrumSize());
  MakeScaleSpectrum(h, conv, scale, shape, h.fft_spectrum.get(), stream);
  for (uint64_t i = 0; i != n_images; ++i)
    conv.ConvolveWith(images[i], width, height, h.fft_spectrum.get(), images[i], width, height, 0, stream);
  h.launches += conv.TakeLaunchCount();
}

namespace {

class MultiScaleRun {
 public:
  MultiScaleRun(rdl_handle& h, DeviceCube& cube, cudaStream_t stream)
      : h_(h), cube_(cube), stream_(stream), st_(*h.ms), width_(cube.width), height_(cube.height),
        n_(cube.Pixels()), conv_(GetConvolver(h, cube.width, cube.height)), cache_(h.ms->psf_cache),
        integrated_psf_(cache_.integrated_psf), convolved_psfs_(cache_.convolved_psfs),
        scale_spectra_(cache_.scale_spectra), psf_spectra_(cache_.psf_spectra),
        twice_cache_(cache_.twice) {
    if (h.dist.comm && h.dist.world > 1 && !h.dist.band_set && cube.n_channels > 1) {
      sharded_ = true;
      world_ = h.dist.world;
      rank_ = h.dist.rank;
    }
  }

  rdl_result Execute();
  ~MultiScaleRun() {
    if (comm_stream_) {
      cudaStreamSynchronize(comm_stream_);
      cudaStreamDestroy(comm_stream_);
    }
    for (cudaEvent_t e : events_) cudaEventDestroy(e);
  }

 private:
  std::vector<rdl_scale_info>& Scales() { return st_.scales; }
  const uint8_t* MaskFor(size_t si) const {
    return h_.use_per_scale_masks ? st_.scale_masks[si].get() : h_.mask.get();
  }
  Peak FindPeakOn(const float* image, size_t hb, size_t vb, const uint8_t* mask);
  float RmsOf(const float* image);
  float ReadPixel(const float* image, size_t index);
  const float2* ScaleSpectrum(size_t si);
  void ConvolvePsfs(size_t psf_slot, const float* psf, bool is_integrated);
  void FindPeakDirect(const float* image, size_t si);
  void FindActiveScaleConvolvedMaxima(bool report_rms);
  bool SelectMaximumScale(size_t& scale);
  void ActivateScales(size_t last);
  bool RunWithoutSubMinorLoop(size_t swp, float first_sub_iteration_threshold,
                              float initial_peak_value, PtrTable& conv_residual,
                              const ConstPtrTable& psf_per_image);
  // PSF of slot p convolved with the kernel of scale si. The per-channel ones are made when a
  // scale is first selected (the reference convolves every PSF with every scale up front,
  // multiscale_algorithm.cc:29-88; scales that are never selected never need theirs).
  const float* ConvolvedPsf(size_t p, size_t si) {
    if (Scales2()[si].scale == 0.0f) return base_psf_[p];
    DeviceBuffer<float>& dst = convolved_psfs_[p * st_.scales.size() + si];
    if (dst.empty()) {
      dst.Reserve(n_);
      conv_.ConvolveWith(base_psf_[p], width_, height_, ScaleSpectrum(si), dst.get(), width_, height_, 0,
                         stream_);
    }
    return dst.get();
  }
  const std::vector<rdl_scale_info>& Scales2() const { return st_.scales; }
  const float2* PsfSpectrum(size_t p, size_t si, FftConvolver& padded);
  const float* TwiceConvolvedPsf(size_t p, size_t si, const float2* scale_spectrum);
  void Sync() { RDL_CUDA(cudaStreamSynchronize(stream_)); }
  // components of image i at scale si that still have to be transformed into the model; null
  // when HBM is short (then the caller transforms at once, like the reference)
  float* PendingModel(size_t image, size_t si) {
    const auto key = std::make_pair(image, si);
    auto it = pending_model_.find(key);
    if (it != pending_model_.end()) return it->second.get();
    if (DeviceAvailableBytes() < n_ * sizeof(float) * 4 + (size_t(16) << 30)) return nullptr;
    DeviceBuffer<float> buf(n_);
    RDL_CUDA(cudaMemsetAsync(buf.get(), 0, n_ * sizeof(float), stream_));
    float* p = buf.get();
    pending_model_.emplace(key, std::move(buf));
    return p;
  }
  void FlushModelTransforms() {
    for (auto& [key, buf] : pending_model_)
      conv_.ConvolveWith(buf.get(), width_, height_, ScaleSpectrum(key.second), cube_.model.p[key.first],
                         width_, height_, 2, stream_);
    h_.launches += conv_.TakeLaunchCount();
    if (!pending_model_.empty()) Sync();  // the buffers go back to the pool below
    pending_model_.clear();
  }
  std::map<std::pair<size_t, size_t>, DeviceBuffer<float>> pending_model_;
  // host-memory calls upload the model planes on a second stream (abi.cu): wait for them before
  // the first kernel that touches the model
  void WaitForModel() {
    if (!h_.model_pending) return;
    RDL_CUDA(cudaStreamWaitEvent(stream_, h_.model_ready, 0));
    h_.model_pending = false;
  }
  // channel sharding over the ranks
  int OwnerOfChannel(size_t channel) const {
    return sharded_ ? int(channel * size_t(world_) / cube_.n_channels) : 0;
  }
  bool MineChannel(size_t channel) const { return !sharded_ || OwnerOfChannel(channel) == rank_; }
  bool MineImage(size_t image) const { return MineChannel(cube_.PsfIndex(image)); }
  // every plane goes from the rank that owns its channel to all the others
  void ShareImagePlanes(float* const* planes) {  // one plane per image
    if (!sharded_) return;
    std::vector<int> roots(cube_.n_images);
    for (size_t i = 0; i != cube_.n_images; ++i) roots[i] = OwnerOfChannel(cube_.PsfIndex(i));
    NcclBroadcastPlanes(h_.dist, planes, roots.data(), cube_.n_images, n_, stream_);
  }

  // Pipelined exchange: planes travel on a second stream while the next channel's convolution
  // runs on the main one. BeginShare: the exchange stream waits for everything queued so far
  // (receive buffers are free); SharePlane: the owner's exchange waits for the producing kernels
  // of that plane only; EndShare: the main stream waits for the exchange. Every rank calls
  // SharePlane in the same order (one communicator, one issue order).
  void BeginShare() {
    if (!sharded_) return;
    if (!comm_stream_) RDL_CUDA(cudaStreamCreateWithFlags(&comm_stream_, cudaStreamNonBlocking));
    WaitOn(comm_stream_, stream_);
  }
  void SharePlane(float* plane, int root) {
    if (!sharded_) return;
    if (root == rank_) WaitOn(comm_stream_, stream_);
    NcclBroadcastPlanes(h_.dist, &plane, &root, 1, n_, comm_stream_);
  }
  void EndShare() {
    if (!sharded_) return;
    WaitOn(stream_, comm_stream_);
  }
  void WaitOn(cudaStream_t waiter, cudaStream_t producer) {
    // a stream wait binds to the record that precedes it, so events can be recycled
    if (events_.size() < 32) {
      cudaEvent_t fresh = nullptr;
      RDL_CUDA(cudaEventCreateWithFlags(&fresh, cudaEventDisableTiming));
      events_.push_back(fresh);
    }
    cudaEvent_t e = events_[next_event_++ % events_.size()];
    RDL_CUDA(cudaEventRecord(e, producer));
    RDL_CUDA(cudaStreamWaitEvent(waiter, e, 0));
  }
  cudaStream_t comm_stream_ = nullptr;
  std::vector<cudaEvent_t> events_;
  size_t next_event_ = 0;

  bool sharded_ = false;
  int world_ = 1, rank_ = 0;
  rdl_handle& h_;
  DeviceCube& cube_;
  cudaStream_t stream_;
  MultiScaleState& st_;
  size_t width_, height_, n_;
  FftConvolver& conv_;
  // PSF-derived planes and spectra live in the state's cache (see MultiScalePsfCache)
  MultiScalePsfCache& cache_;
  bool cache_hit_ = false;
  bool CacheMatches() const;
  void StampCache();
  DeviceBuffer<float> integrated_, scratch_;
  DeviceBuffer<float>& integrated_psf_;
  std::vector<DeviceBuffer<float>>& convolved_psfs_;
  std::vector<const float*> base_psf_;               // scale-0 "convolved" PSF of each slot
  std::vector<DeviceBuffer<float>> twice_convolved_, individual_convolved_;
  std::vector<DeviceBuffer<float2>>& scale_spectra_;
  std::map<std::pair<size_t, size_t>, DeviceBuffer<float2>>& psf_spectra_;
  std::map<std::pair<size_t, size_t>, DeviceBuffer<float>>& twice_cache_;
  DeviceBuffer<float2> psf_spectrum_tmp_;
  DeviceBuffer<float> shape_dev_, values_dev_;  // loop without the sub-minor optimisation
  DeviceBuffer<double> sumsq_;
};

bool MultiScaleRun::CacheMatches() const {
  const MultiScalePsfCache& c = cache_;
  if (!c.valid || !h_.psfs_constant) return false;
  if (c.width != width_ || c.height != height_ || c.n_pol != cube_.n_pol ||
      c.shape != h_.config.ms_shape || c.padding != double(h_.config.ms_convolution_padding) ||
      c.sharded != sharded_ || c.weights != cube_.weights)
    return false;
  if (c.psf_ptrs.size() != cube_.n_channels || c.scale_sizes.size() != st_.scales.size() ||
      c.psf_peak.size() != st_.scales.size())
    return false;
  for (size_t p = 0; p != cube_.n_channels; ++p)
    if (c.psf_ptrs[p] != cube_.psf.p[p]) return false;
  for (size_t si = 0; si != st_.scales.size(); ++si)
    if (c.scale_sizes[si] != st_.scales[si].scale) return false;
  return true;
}

void MultiScaleRun::StampCache() {
  MultiScalePsfCache& c = cache_;
  c.width = width_;
  c.height = height_;
  c.n_pol = cube_.n_pol;
  c.shape = h_.config.ms_shape;
  c.padding = double(h_.config.ms_convolution_padding);
  c.sharded = sharded_;
  c.weights = cube_.weights;
  c.psf_ptrs.assign(cube_.psf.p, cube_.psf.p + cube_.n_channels);
  c.scale_sizes.clear();
  for (const rdl_scale_info& e : st_.scales) c.scale_sizes.push_back(e.scale);
}

float MultiScaleRun::ReadPixel(const float* image, size_t index) {
  float v = 0.0f;
  RDL_CUDA(cudaMemcpyAsync(&v, image + index, sizeof(float), cudaMemcpyDeviceToHost, stream_));
  Sync();
  return v;
}

Peak MultiScaleRun::FindPeakOn(const float* image, size_t hb, size_t vb, const uint8_t* mask) {
  IntegSpec spec{};
  spec.mode = kIntegCopy;
  spec.n_terms = 1;
  ConstPtrTable t{};
  t.p[0] = image;
  const PeakWindow w = MakePeakWindow(width_, height_, 0, height_, hb, vb, mask, h_.rms.get(),
                                      h_.settings.allow_negative_components);
  h_.key.Reserve(1);
  const bool aligned = (reinterpret_cast<uintptr_t>(image) & 15u) == 0 && width_ % 4 == 0;
  LaunchFindPeak(spec, t, int(width_), int(height_), w, h_.key.get(), aligned ? 4 : 1, stream_);
  ++h_.launches;
  unsigned long long key = 0;
  RDL_CUDA(cudaMemcpyAsync(&key, h_.key.get(), sizeof(key), cudaMemcpyDeviceToHost, stream_));
  Sync();
  Peak p;
  uint32_t index = 0;
  p.has = DecodePeakKey(key, p.value, index);
  if (p.has) {
    p.x = index % width_;
    p.y = index / width_;
  }
  return p;
}

float MultiScaleRun::RmsOf(const float* image) {
  sumsq_.Reserve(1);
  RDL_CUDA(cudaMemsetAsync(sumsq_.get(), 0, sizeof(double), stream_));
  SumSquaresKernel<<<kSmCount * 4, 256, 0, stream_>>>(image, n_, sumsq_.get());
  ++h_.launches;
  double s = 0.0;
  RDL_CUDA(cudaMemcpyAsync(&s, sumsq_.get(), sizeof(double), cudaMemcpyDeviceToHost, stream_));
  Sync();
  return float(std::sqrt(s / double(n_)));
}

const float2* MultiScaleRun::ScaleSpectrum(size_t si) {
  if (scale_spectra_.size() < Scales().size()) scale_spectra_.resize(Scales().size());
  if (scale_spectra_[si].empty()) {
    scale_spectra_[si].Resize(conv_.SpectrumSize());
    MakeScaleSpectrum(h_, conv_, Scales()[si].scale, h_.config.ms_shape, scale_spectra_[si].get(), stream_);
  }
  return scale_spectra_[si].get();
}

// ConvolvePsfs, multiscale_algorithm.cc:29-88
void MultiScaleRun::ConvolvePsfs(size_t slot, const float* psf, bool is_integrated) {
  const size_t n_scales = Scales().size();
  const double first_auto_scale_size = h_.beam_size_in_pixels * 2.0;
  base_psf_[slot] = psf;
  for (size_t si = 0; si != n_scales; ++si) {
    rdl_scale_info& e = Scales()[si];
    const float* convolved = psf;
    if (e.scale != 0.0f) {
      ++h_.reference_convolutions[{width_, height_}];
      // per-channel PSFs: only the owner of the channel ever uses its convolved PSFs
      if (!is_integrated && !MineChannel(slot)) continue;
      DeviceBuffer<float>& dst = convolved_psfs_[slot * n_scales + si];
      if (!is_integrated) {
        // per-channel PSFs: on demand (ConvolvedPsf); slot 0 held the integrated PSF's until now
        if (!cache_hit_) dst = DeviceBuffer<float>();
        continue;
      }
      if (!cache_hit_) {
        dst.Reserve(n_);
        conv_.ConvolveWith(psf, width_, height_, ScaleSpectrum(si), dst.get(), width_, height_, 0, stream_);
      }
      convolved = dst.get();
    }
    if (is_integrated) {
      if (!cache_hit_) cache_.psf_peak[si] = ReadPixel(convolved, width_ / 2 + (height_ / 2) * width_);
      e.psf_peak = cache_.psf_peak[si];
      double exp_term;
      if (e.scale == 0.0f || n_scales < 2)
        exp_term = 0.0;
      else
        exp_term = std::log2(e.scale / first_auto_scale_size);
      e.bias_factor = float(std::pow(h_.config.ms_scale_bias, -exp_term));
      e.gain = float(h_.settings.minor_loop_gain / e.psf_peak);
      e.is_active = 1;
    }
  }
}

// FindPeakDirect, multiscale_algorithm.cc:700-748
void MultiScaleRun::FindPeakDirect(const float* image, size_t si) {
  rdl_scale_info& info = Scales()[si];
  const size_t hb = size_t(std::round(width_ * h_.settings.clean_border_ratio));
  const size_t vb = size_t(std::round(height_ * h_.settings.clean_border_ratio));
  const Peak p = FindPeakOn(image, hb, vb, MaskFor(si));
  if (p.has) {
    info.max_x = p.x;
    info.max_y = p.y;
    info.max_unnormalized_image_value = p.value;
    if (!h_.rms.get())
      info.max_normalized_image_value = p.value;
    else
      info.max_normalized_image_value = p.value / ReadPixel(h_.rms.get(), p.x + p.y * width_);
  } else {
    // peak_finder leaves x, y at width, height when nothing is found
    info.max_x = width_;
    info.max_y = height_;
    info.max_unnormalized_image_value = 0.0f;
    info.max_normalized_image_value = 0.0f;
  }
}

// FindActiveScaleConvolvedMaxima (multiscale_algorithm.cc:578-634) with
// ThreadedDeconvolutionTools::FindMultiScalePeak (threaded_deconvolution_tools.cc:30-107)
void MultiScaleRun::FindActiveScaleConvolvedMaxima(bool report_rms) {
  ConstPtrTable t{};
  for (size_t i = 0; i != cube_.n_images; ++i) t.p[i] = cube_.residual.p[i];
  LaunchIntegrate(MakeLinearSpec(cube_), t, n_, integrated_.get(), cube_.vec, stream_);
  ++h_.launches;
  for (size_t si = 0; si != Scales().size(); ++si) {
    rdl_scale_info& e = Scales()[si];
    if (!e.is_active) continue;
    if (e.scale == 0.0f) {
      FindPeakDirect(integrated_.get(), si);
      if (report_rms) e.rms = RmsOf(integrated_.get());
      continue;
    }
    ++h_.reference_convolutions[{width_, height_}];
    conv_.ConvolveWith(integrated_.get(), width_, height_, ScaleSpectrum(si), scratch_.get(), width_,
                       height_, 0, stream_);
    const size_t border_scale = size_t(std::ceil(e.scale * 0.5));
    const size_t hb = std::max<size_t>(size_t(std::round(width_ * h_.settings.clean_border_ratio)), border_scale);
    const size_t vb = std::max<size_t>(size_t(std::round(height_ * h_.settings.clean_border_ratio)), border_scale);
    if (report_rms) e.rms = RmsOf(scratch_.get());
    const Peak p = FindPeakOn(scratch_.get(), hb, vb, MaskFor(si));
    e.max_unnormalized_image_value = p.has ? p.value : 0.0f;
    e.max_normalized_image_value = e.max_unnormalized_image_value;
    if (p.has && h_.rms.get())
      e.max_normalized_image_value = p.value / ReadPixel(h_.rms.get(), p.x + p.y * width_);
    e.max_x = p.has ? p.x : width_;
    e.max_y = p.has ? p.y : height_;
  }
  h_.launches += conv_.TakeLaunchCount();
}

// SelectMaximumScale, multiscale_algorithm.cc:133-151: std::map keeps the first of equal keys
bool MultiScaleRun::SelectMaximumScale(size_t& scale) {
  std::map<float, size_t> peak_to_scale;
  for (size_t i = 0; i != Scales().size(); ++i)
    if (Scales()[i].is_active) {
      const float max_val = std::fabs(Scales()[i].max_unnormalized_image_value * Scales()[i].bias_factor);
      peak_to_scale.insert({max_val, i});
    }
  if (peak_to_scale.empty()) return false;
  scale = peak_to_scale.rbegin()->second;
  return true;
}

void MultiScaleRun::ActivateScales(size_t last) {  // multiscale_algorithm.cc:636-656
  auto& s = Scales();
  for (size_t i = 0; i != s.size(); ++i) {
    const bool do_activate =
        i == last || std::fabs(s[i].max_unnormalized_image_value) * s[i].bias_factor >
                         std::fabs(s[last].max_unnormalized_image_value) *
                             (1.0 - h_.settings.minor_loop_gain) * s[last].bias_factor;
    s[i].is_active = do_activate;
  }
}

// Spectrum of PrepareConvolutionKernel(Untrim(convolved PSF)) for CorrectResidualDirty. The
// reference recomputes it per image and per minor iteration (subminor_loop.cc:199-202); it only
// depends on (psf, scale), so it is cached for the duration of the call when memory allows.
// The minor loop without the sub-minor optimisation (multiscale_algorithm.cc:462-520): one
// component per iteration at the peak of the selected scale; it is subtracted from the data
// images with the scale-convolved PSF and from the scale-convolved images with the twice-convolved
// PSF, added to the model as a scaled shape function, and the peak of that scale is searched
// again. The host reads one peak per iteration, like the reference.
bool MultiScaleRun::RunWithoutSubMinorLoop(size_t swp, float first_sub_iteration_threshold,
                                           float initial_peak_value, PtrTable& convolved,
                                           const ConstPtrTable& twice_convolved) {
  rdl_handle& h = h_;
  const Settings& set = h.settings;
  rdl_scale_info& info = Scales()[swp];
  const int n_images = int(cube_.n_images);
  size_t shape_n = 0;
  if (info.scale != 0.0f) {
    const std::vector<float> shape =
        MakeShapeFunction(info.scale, shape_n, std::min(width_, height_), h.config.ms_shape);
    shape_dev_.Reserve(shape.size());
    RDL_CUDA(cudaMemcpyAsync(shape_dev_.get(), shape.data(), shape.size() * sizeof(float),
                             cudaMemcpyHostToDevice, stream_));
    Sync();  // `shape` is pageable host memory
  }
  ConstPtrTable convolved_const{}, psf_at_scale{};
  PtrTable residual{}, model{};
  for (int i = 0; i != n_images; ++i) {
    convolved_const.p[i] = convolved.p[i];
    psf_at_scale.p[i] = ConvolvedPsf(cube_.PsfIndex(i), swp);
    residual.p[i] = cube_.residual.p[i];
    model.p[i] = cube_.model.p[i];
  }
  values_dev_.Reserve(kMaxImages);
  const IntegSpec linear = MakeLinearSpec(cube_);
  const int n_fit = int(h.fit_freqs.size()), n_pol = int(cube_.n_pol);
  bool diverging = false;
  while (h.iteration_number < set.max_iterations &&
         std::fabs(info.max_unnormalized_image_value * info.bias_factor) >
             first_sub_iteration_threshold &&
         (!set.stop_on_negative_component || info.max_unnormalized_image_value >= 0.0f) &&
         !diverging) {
    const size_t x = info.max_x, y = info.max_y, index = x + y * width_;
    // MeasureComponentValues + PerformSpectralFit (deconvolution_algorithm.cc:29-46)
    float values[kMaxImages];
    GatherPixelKernel<<<1, kMaxImages, 0, stream_>>>(convolved_const, n_images, index,
                                                     values_dev_.get());
    RDL_CUDA(cudaMemcpyAsync(values, values_dev_.get(), n_images * sizeof(float),
                             cudaMemcpyDeviceToHost, stream_));
    Sync();
    if (h.fit_mode == RDL_FIT_POLYNOMIAL) {
      float fitted[kMaxImages];
      for (int i = 0; i != n_images; ++i) {
        const int ch = i / n_pol, pol = i - ch * n_pol;
        double sum = 0.0;
        for (int j = 0; j != n_fit; ++j)
          sum = std::fma(h.projection_host[ch * n_fit + j], double(values[j * n_pol + pol]), sum);
        fitted[i] = float(sum);
      }
      std::copy_n(fitted, n_images, values);
    }
    ComponentValues gains{};
    for (int i = 0; i != n_images; ++i) {
      values[i] = values[i] * info.gain;
      gains.v[i] = values[i];
    }
    LaunchSubtract(residual, psf_at_scale, n_images, int(width_), int(height_), int(x), int(y),
                   values, stream_);
    LaunchSubtract(convolved, twice_convolved, n_images, int(width_), int(height_), int(x), int(y),
                   values, stream_);
    // AddComponentToModel (:676-698), once per image
    if (info.scale == 0.0f) {
      AddPointKernel<<<1, kMaxImages, 0, stream_>>>(model, n_images, index, gains);
    } else {
      const dim3 grid(unsigned((shape_n * shape_n + 255) / 256), unsigned(n_images));
      AddShapeKernel<<<grid, 256, 0, stream_>>>(model, n_images, width_, height_, shape_dev_.get(),
                                                shape_n, x, y, gains);
    }
    for (int i = 0; i != n_images; ++i) {
      info.n_components_cleaned++;
      info.total_flux_cleaned += values[i];
    }
    if (h.track_per_scale_masks)
      RDL_CUDA(cudaMemsetAsync(st_.scale_masks[swp].get() + index, 1, 1, stream_));
    if (h.config.track_components) h.components.Add(x, y, swp, values);  // :502-504
    // the peak of this scale again
    LaunchIntegrate(linear, convolved_const, n_, integrated_.get(), cube_.vec, stream_);
    h.launches += 5;
    FindPeakDirect(integrated_.get(), swp);
    const float abs_peak_value =
        std::fabs(info.max_unnormalized_image_value * info.bias_factor);
    if (set.divergence_limit != 0.0f)
      diverging = abs_peak_value > initial_peak_value * set.divergence_limit;
    ++h.iteration_number;
    if (h.trace_host.size() < h.trace_capacity)
      h.trace_host.push_back(TraceEntry{uint32_t(h.iteration_number), uint32_t(swp),
                                        uint32_t(info.max_x), uint32_t(info.max_y),
                                        info.max_unnormalized_image_value});
  }
  RDL_CUDA(cudaGetLastError());
  return diverging;
}

// The PSF of slot p convolved twice with the kernel of scale si (multiscale_algorithm.cc:331-343).
// The reference recomputes it in every minor iteration that selects the scale; the PSFs do not
// change during a call, so it is kept while HBM allows (K4: 16 PSFs x 4 scales x 268 MB).
const float* MultiScaleRun::TwiceConvolvedPsf(size_t p, size_t si, const float2* scale_spectrum) {
  ++h_.reference_convolutions[{width_, height_}];  // the reference transforms it again every time
  const auto key = std::make_pair(p, si);
  auto it = twice_cache_.find(key);
  if (it != twice_cache_.end()) return it->second.get();
  if (sharded_) {
    // the sub-minor loop of every rank reads the twice-convolved PSFs of ALL channels: each is
    // computed by its owner once per call and scale, and sent to the others
    std::vector<float*> planes(cube_.n_channels);
    std::vector<int> roots(cube_.n_channels);
    for (size_t q = 0; q != cube_.n_channels; ++q) {
      DeviceBuffer<float> buf(n_);
      planes[q] = buf.get();
      roots[q] = OwnerOfChannel(q);
      if (MineChannel(q))
        conv_.ConvolveWith(ConvolvedPsf(q, si), width_, height_, scale_spectrum, planes[q], width_,
                           height_, 0, stream_);
      twice_cache_.emplace(std::make_pair(q, si), std::move(buf));
    }
    NcclBroadcastPlanes(h_.dist, planes.data(), roots.data(), planes.size(), n_, stream_);
    return twice_cache_.find(key)->second.get();
  }
  const size_t free_bytes = DeviceAvailableBytes();
  float* out = nullptr;
  if (free_bytes > n_ * sizeof(float) * 4 + (size_t(16) << 30)) {
    DeviceBuffer<float> buf(n_);
    out = buf.get();
    twice_cache_.emplace(key, std::move(buf));
  } else {
    twice_convolved_[p].Reserve(n_);
    out = twice_convolved_[p].get();
  }
  conv_.ConvolveWith(ConvolvedPsf(p, si), width_, height_, scale_spectrum, out, width_, height_, 0,
                     stream_);
  return out;
}

const float2* MultiScaleRun::PsfSpectrum(size_t p, size_t si, FftConvolver& padded) {
  const auto key = std::make_pair(p, si);
  auto it = psf_spectra_.find(key);
  if (it != psf_spectra_.end()) return it->second.get();
  const size_t free_bytes = DeviceAvailableBytes();
  const size_t need = padded.SpectrumSize() * sizeof(float2);
  float2* out = nullptr;
  if (free_bytes > need * 4 + (size_t(8) << 30)) {
    DeviceBuffer<float2> buf(padded.SpectrumSize());
    out = buf.get();
    psf_spectra_.emplace(key, std::move(buf));
  } else {
    psf_spectrum_tmp_.Reserve(padded.SpectrumSize());
    out = psf_spectrum_tmp_.get();
  }
  padded.MakePsfKernelSpectrum(ConvolvedPsf(p, si), width_, height_, out, stream_);
  h_.launches += padded.TakeLaunchCount();
  return out;
}

rdl_result MultiScaleRun::Execute() {  // multiscale_algorithm.cc:183-576
  rdl_handle& h = h_;
  Settings& set = h.settings;
  if (set.stop_on_negative_component) set.allow_negative_components = true;
  if (!h.config.ms_fast_sub_minor_loop && h.fit_mode != RDL_FIT_NONE && h.fit_mode != RDL_FIT_POLYNOMIAL)
    throw std::domain_error("multiscale without the sub-minor loop supports the polynomial fit only");
  InitializeScales(Scales(), h.beam_size_in_pixels, std::min(width_, height_), h.config.ms_shape,
                   h.config.ms_max_scales, h.ms_scale_list);
  const size_t n_scales = Scales().size();
  if (h.config.track_components) {  // multiscale_algorithm.cc:228-240
    if (h.components.Empty()) {
      h.components.width = width_;
      h.components.height = height_;
      h.components.n_values = cube_.n_images;
      h.components.scales.resize(n_scales);
    } else if (h.components.width != width_ || h.components.height != height_) {
      throw std::invalid_argument("Error in component list dimensions!");
    }
  }
  if (h.track_per_scale_masks) {
    // multiscale_algorithm.cc:214-229
    for (const auto& m : st_.scale_masks)
      if (m.size() != n_) throw std::invalid_argument("Invalid automask size in multiscale algorithm");
    while (st_.scale_masks.size() < n_scales) {
      st_.scale_masks.emplace_back(n_);
      RDL_CUDA(cudaMemsetAsync(st_.scale_masks.back().get(), 0, n_, stream_));
    }
    st_.mask_pixels = n_;
  }
  if (h.use_per_scale_masks && st_.scale_masks.size() < n_scales)
    throw std::invalid_argument("per-scale masks requested before they were tracked");

  size_t threshold_countdown = std::max<size_t>(8, n_scales * 3 / 2);
  integrated_.Resize(n_);
  scratch_.Resize(n_);
  const size_t n_psfs = cube_.n_channels;
  cache_hit_ = CacheMatches();
  if (!cache_hit_) {
    cache_.Clear();
    integrated_psf_.Resize(n_);
    convolved_psfs_.resize(n_psfs * n_scales);
    cache_.psf_peak.assign(n_scales, 0.0f);
    StampCache();
  }
  base_psf_.assign(n_psfs, nullptr);

  // GetIntegratedPsf (image_set.cc:499-530) + ConvolvePsfs
  if (!cache_hit_) {
    ConstPtrTable t{};
    for (size_t p = 0; p != n_psfs; ++p) t.p[p] = cube_.psf.p[p];
    const bool aligned = cube_.vec == 4;
    LaunchIntegrate(MakeIntegratedPsfSpec(cube_), t, n_, integrated_psf_.get(), aligned ? 4 : 1, stream_);
    ++h.launches;
  }
  ConvolvePsfs(0, integrated_psf_.get(), true);
  if (n_psfs > 1)
    for (size_t p = 0; p != n_psfs; ++p) ConvolvePsfs(p, cube_.psf.p[p], false);
  h.launches += conv_.TakeLaunchCount();
  cache_.valid = h.psfs_constant;  // the lazily filled maps (spectra, twice-convolved) only grow

  FindActiveScaleConvolvedMaxima(true);
  rdl_result result{};
  result.iteration_number = h.iteration_number;
  size_t swp = 0;
  if (!SelectMaximumScale(swp)) return result;

  auto& S = Scales();
  bool is_final_threshold = false;
  const float initial_peak_value = std::fabs(S[swp].max_unnormalized_image_value * S[swp].bias_factor);
  float mgain_threshold = initial_peak_value * (1.0 - set.major_loop_gain);
  mgain_threshold = std::max(mgain_threshold, set.major_iteration_threshold);
  float first_threshold = mgain_threshold;
  if (set.threshold > first_threshold) {
    first_threshold = set.threshold;
    is_final_threshold = true;
  }

  twice_convolved_.resize(n_psfs);
  individual_convolved_.resize(cube_.n_images);
  bool diverging = false;
  const IntegSpec linear = MakeLinearSpec(cube_);
  h.sub.width = width_;

  while (h.iteration_number < set.max_iterations &&
         std::fabs(S[swp].max_unnormalized_image_value * S[swp].bias_factor) > first_threshold &&
         (!set.stop_on_negative_component || S[swp].max_unnormalized_image_value >= 0.0f) &&
         threshold_countdown > 0 && !diverging) {
    ++h.minor_iterations;
    if (h.trace_host.size() < h.trace_capacity)
      h.trace_host.push_back(TraceEntry{uint32_t(h.iteration_number), uint32_t(swp), uint32_t(S[swp].max_x),
                                        uint32_t(S[swp].max_y), S[swp].max_unnormalized_image_value});
    // transformList: convolved PSFs and the data images at the selected scale (:331-354)
    ConstPtrTable psf_per_image{}, conv_residual{};
    if (S[swp].scale != 0.0f) {
      const float2* spectrum = ScaleSpectrum(swp);
      std::vector<const float*> twice(n_psfs);
      for (size_t p = 0; p != n_psfs; ++p) twice[p] = TwiceConvolvedPsf(p, swp, spectrum);
      for (size_t i = 0; i != cube_.n_images; ++i) individual_convolved_[i].Reserve(n_);
      BeginShare();
      for (size_t i = 0; i != cube_.n_images; ++i) {
        ++h.reference_convolutions[{width_, height_}];
        if (MineImage(i))
          conv_.ConvolveWith(cube_.residual.p[i], width_, height_, spectrum,
                             individual_convolved_[i].get(), width_, height_, 0, stream_);
        SharePlane(individual_convolved_[i].get(), OwnerOfChannel(cube_.PsfIndex(i)));
        conv_residual.p[i] = individual_convolved_[i].get();
        psf_per_image.p[i] = twice[cube_.PsfIndex(i)];
      }
      EndShare();
      h.launches += conv_.TakeLaunchCount();
    } else {
      // scale 0: the "convolved" images are the images themselves; the sub-minor loop only
      // reads them (it works on gathered copies), so no copy is needed.
      for (size_t i = 0; i != cube_.n_images; ++i) {
        conv_residual.p[i] = cube_.residual.p[i];
        psf_per_image.p[i] = ConvolvedPsf(cube_.PsfIndex(i), swp);
        if (!h.config.ms_fast_sub_minor_loop) {
          // without the sub-minor loop the convolved images are subtracted from as well (:489):
          // they have to be the copies the reference makes (:346-349)
          individual_convolved_[i].Reserve(n_);
          RDL_CUDA(cudaMemcpyAsync(individual_convolved_[i].get(), cube_.residual.p[i],
                                   n_ * sizeof(float), cudaMemcpyDeviceToDevice, stream_));
          conv_residual.p[i] = individual_convolved_[i].get();
        }
      }
    }

    const float sub_iteration_gain_threshold =
        std::fabs(S[swp].max_unnormalized_image_value * S[swp].bias_factor) *
        (1.0 - h.config.ms_sub_minor_loop_gain);
    float first_sub_iteration_threshold = sub_iteration_gain_threshold;
    if (first_threshold > first_sub_iteration_threshold) {
      first_sub_iteration_threshold = first_threshold;
      threshold_countdown--;
    }

    if (!h.config.ms_fast_sub_minor_loop) {
      if (sharded_)
        throw std::domain_error("channel-sharded multiscale needs the fast sub-minor loop");
      PtrTable convolved{};
      for (size_t i = 0; i != cube_.n_images; ++i) convolved.p[i] = individual_convolved_[i].get();
      WaitForModel();
      diverging = RunWithoutSubMinorLoop(swp, first_sub_iteration_threshold, initial_peak_value,
                                         convolved, psf_per_image);
      ActivateScales(swp);
      FindActiveScaleConvolvedMaxima(false);
      if (!SelectMaximumScale(swp)) {
        result.iteration_number = h.iteration_number;
        result.another_iteration_required = 0;
        return result;
      }
      continue;
    }

    const uint64_t sub_minor_start_iteration = h.iteration_number;
    const size_t conv_w = GetConvolutionSize(S[swp].scale, width_, h.config.ms_convolution_padding);
    const size_t conv_h = GetConvolutionSize(S[swp].scale, height_, h.config.ms_convolution_padding);
    SubMinorSetup sub;
    sub.width = width_;
    sub.height = height_;
    sub.current_iteration = h.iteration_number;
    sub.max_iterations = set.max_iterations;
    sub.threshold = first_sub_iteration_threshold / S[swp].bias_factor;
    sub.gain = S[swp].gain;
    sub.divergence_limit = set.divergence_limit;
    sub.allow_negative = set.allow_negative_components;
    sub.stop_on_negative = set.stop_on_negative_component;
    const size_t scale_border = size_t(std::ceil(S[swp].scale * 0.5));
    sub.h_border = std::max<size_t>(size_t(std::round(width_ * set.clean_border_ratio)), scale_border);
    sub.v_border = std::max<size_t>(size_t(std::round(height_ * set.clean_border_ratio)), scale_border);
    sub.rms_factor = h.rms.get();
    sub.mask = MaskFor(swp);
    sub.trace_scale = uint32_t(swp);

    const SubMinorOutcome o = RunSubMinorLoop(h, linear, conv_residual, psf_per_image, int(cube_.n_images),
                                              int(cube_.n_pol), sub, stream_);
    SubMinorCollectTrace(h, stream_);
    diverging = o.diverging;
    if (set.divergence_limit != 0.0f && o.has_value)
      diverging = diverging || std::fabs(o.max_value) > initial_peak_value * set.divergence_limit;
    if (!o.has_value) break;

    h.iteration_number = o.iteration;
    S[swp].n_components_cleaned += h.iteration_number - sub_minor_start_iteration;
    S[swp].total_flux_cleaned += o.flux_cleaned;

    FftConvolver& padded = GetConvolver(h, conv_w, conv_h);
    // the reference updates the mask / component list inside the loop at image 0; both only read
    // the compact model of the sub-minor run, which every rank holds
    if (h.track_per_scale_masks)
      SubMinorUpdateAutoMask(h, int(cube_.n_images), st_.scale_masks[swp].get(), stream_);
    if (h.config.track_components)
      SubMinorUpdateComponentList(h, int(cube_.n_images), swp, stream_);
    WaitForModel();
    BeginShare();
    for (size_t i = 0; i != cube_.n_images; ++i) {
      ++h.reference_convolutions[{conv_w, conv_h}];
      if (S[swp].scale != 0.0f) ++h.reference_convolutions[{width_, height_}];
      if (!MineImage(i)) {
        SharePlane(cube_.residual.p[i], OwnerOfChannel(cube_.PsfIndex(i)));
        continue;
      }
      const float2* psf_spectrum = PsfSpectrum(cube_.PsfIndex(i), swp, padded);
      SubMinorScatterModel(h, int(i), scratch_.get(), n_, stream_);
      // CorrectResidualDirty (subminor_loop.cc:195-218): residual -= Trim(Convolve(Untrim(model)))
      padded.ConvolveWith(scratch_.get(), width_, height_, psf_spectrum, cube_.residual.p[i], width_,
                          height_, 1, stream_);
      // the corrected residual of this channel, for the joined peak search of all ranks: on its
      // way while the model transform below runs
      SharePlane(cube_.residual.p[i], rank_);
      // model += Transform(full model, scale)  (multiscale_algorithm.cc:442-460). The transform
      // is linear: the components of every minor iteration that picked this scale are collected
      // per (image, scale) and transformed ONCE when the call ends (FlushModelTransforms),
      // memory permitting -- one convolution per image and scale used instead of one per image and
      // minor iteration. The model differs from the iteration-by-iteration sum by fp32 round-off
      // only and feeds back into nothing (peaks come from the residual).
      if (S[swp].scale == 0.0f) {
        SubMinorAddModelTo(h, int(i), cube_.model.p[i], stream_);
      } else if (float* pending = PendingModel(i, swp)) {
        SubMinorAddModelTo(h, int(i), pending, stream_);
      } else {
        conv_.ConvolveWith(scratch_.get(), width_, height_, ScaleSpectrum(swp), cube_.model.p[i], width_,
                           height_, 2, stream_);
      }
    }
    EndShare();
    h.launches += padded.TakeLaunchCount() + conv_.TakeLaunchCount();

    ActivateScales(swp);
    FindActiveScaleConvolvedMaxima(false);
    if (!SelectMaximumScale(swp)) {
      WaitForModel();
      FlushModelTransforms();
      ShareImagePlanes(cube_.model.p);
      Sync();
      result.iteration_number = h.iteration_number;
      result.another_iteration_required = 0;
      return result;
    }
  }
  WaitForModel();
  FlushModelTransforms();
  ShareImagePlanes(cube_.model.p);

  const bool max_iter_reached = h.iteration_number >= set.max_iterations;
  const bool negative_reached =
      set.stop_on_negative_component && S[swp].max_unnormalized_image_value < 0.0f;
  result.is_diverging = diverging;
  result.another_iteration_required =
      !max_iter_reached && !is_final_threshold && !negative_reached && !diverging;
  result.final_peak_value = S[swp].max_unnormalized_image_value * S[swp].bias_factor;
  result.iteration_number = h.iteration_number;
  Sync();
  return result;
}

}  // namespace

rdl_result RunMultiScale(rdl_handle& h, DeviceCube& cube, cudaStream_t stream) {
  if (!h.ms) h.ms = std::make_shared<MultiScaleState>();
  h.trace_host.clear();
  h.minor_iterations = 0;
  h.reference_convolutions.clear();
  try {
    MultiScaleRun run(h, cube, stream);
    const rdl_result r = run.Execute();
    if (!h.psfs_constant) {  // nothing is kept between calls: give the HBM back
      h.ms->psf_cache.Clear();
      h.ms->psf_cache.integrated_psf = DeviceBuffer<float>();
    }
    return r;
  } catch (...) {
    h.ms->psf_cache.Clear();
    throw;
  }
}

}  // namespace rdl
