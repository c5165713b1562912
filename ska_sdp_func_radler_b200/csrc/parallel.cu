// ParallelDeconvolution on the device (cpp/algorithms/parallel_deconvolution.cc):
//   SetAlgorithm :227-242 (one clone per facet), RunSubImage :294-492 (trim, masked model trim,
//   centre-cropped PSFs, peak-only probe, deconvolve, divergence rollback, masked copy-back,
//   model add), ExecuteParallelRun :548-654 (two passes, global start peak, major-iteration
//   threshold, result bookkeeping). The facet geometry is computed on the host (facets.cc);
//   every pixel operation — trims, masked copies, model accumulation and the per-facet
//   ExecuteMajorIteration — runs on the device, so the cube never leaves HBM.
// Facets run one after the other on the handle's stream, in index order: the reference's result
// depends on the order in which facets copy their residual back (a facet's bounding box overlaps
// its neighbours' pixels); index order is what the reference does with one thread.
#include "parallel.cuh"

#include <algorithm>
#include <chrono>
#include <cmath>
#include <stdexcept>

#include "clean_kernels.cuh"
#include "facets.h"
#include "multiscale.cuh"

namespace rdl {
namespace {

constexpr int kThreads = 256;

// dst (w x h) = src box at (x, y), zeroed where mask == 0  (ImageSet::TrimMasked, image_set.h:230)
__global__ void TrimMaskedKernel(float* dst, const float* src, size_t src_width, size_t x, size_t y,
                                 size_t w, size_t h, const uint8_t* mask) {
  const size_t n = w * h;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    const size_t row = i / w, col = i - row * w;
    dst[i] = (mask == nullptr || mask[i]) ? src[(y + row) * src_width + x + col] : 0.0f;
  }
}

// frame box at (x, y) = sub where mask (Image::CopyMasked) / += sub (Image::AddSubImage)
template <bool ADD>
__global__ void PlaceKernel(float* frame, size_t frame_width, size_t x, size_t y, const float* sub,
                            size_t w, size_t h, const uint8_t* mask) {
  const size_t n = w * h;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    const size_t row = i / w, col = i - row * w;
    float* d = frame + (y + row) * frame_width + x + col;
    if (ADD)
      *d = __fadd_rn(*d, sub[i]);
    else if (mask[i])
      *d = sub[i];
  }
}

// owned[box at (x, y)] = value where the facet's boundary mask is set
__global__ void MarkOwnedKernel(uint8_t* owned, size_t frame_width, size_t x, size_t y, size_t w,
                                size_t h, const uint8_t* mask, uint8_t value) {
  const size_t n = w * h;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    const size_t row = i / w, col = i - row * w;
    if (mask[i]) owned[(y + row) * frame_width + x + col] = value;
  }
}

// zero what other ranks own, so that the sum over the ranks picks every pixel from its owner
__global__ void KeepOwnedKernel(float* image, const uint8_t* owned, size_t n) {
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    if (!owned[i]) image[i] = 0.0f;
}

// facet scale mask = frame scale mask box AND the facet's clean mask (:376-389)
__global__ void ScaleMaskInKernel(uint8_t* dst, const uint8_t* frame, size_t frame_width, size_t x,
                                  size_t y, size_t w, size_t h, const uint8_t* facet_mask) {
  const size_t n = w * h;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    const size_t row = i / w, col = i - row * w;
    dst[i] = (frame[(y + row) * frame_width + x + col] && facet_mask[i]) ? 1 : 0;
  }
}

// frame scale mask box = facet scale mask where the facet owns the pixel (:438-452)
__global__ void ScaleMaskOutKernel(uint8_t* frame, size_t frame_width, size_t x, size_t y, size_t w,
                                   size_t h, const uint8_t* facet_scale_mask,
                                   const uint8_t* boundary_mask) {
  const size_t n = w * h;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x) {
    const size_t row = i / w, col = i - row * w;
    if (boundary_mask[i]) frame[(y + row) * frame_width + x + col] = facet_scale_mask[i] ? 1 : 0;
  }
}

double MsSince(std::chrono::steady_clock::time_point t0) {
  return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
}

int GridFor(size_t n) { return int(std::min<size_t>((n + kThreads - 1) / kThreads, kSmCount * 8)); }

}  // namespace

struct ParallelState {
  size_t grid_width = 0, grid_height = 0;
  std::vector<rdl_handle*> algorithms;  // one clone per facet
  std::vector<Facet> facets;            // geometry + results of the last call
  DeviceBuffer<float> integrated, result_model, sub_residual, sub_model, sub_model_start, sub_psf;
  DeviceBuffer<uint8_t> boundary_mask;
  // facet-sharded run over several GPUs (facet i belongs to rank i % world)
  std::vector<uint64_t> iterations;  // IterationNumber() of every facet's algorithm, all ranks
  double geometry_ms = 0.0, total_ms = 0.0;  // host divider search / whole last call
  DeviceBuffer<uint8_t> owned, wire_send, wire_recv;
  DeviceBuffer<uint8_t> facet_mask;  // a facet's clean mask on the device (scale-mask trims)
  ~ParallelState() {
    for (rdl_handle* a : algorithms) rdl_destroy(a);
  }
};

void ParallelStateDeleter::operator()(ParallelState* p) const { delete p; }

namespace {

void EnsureAlgorithms(rdl_handle& h, ParallelState& st, size_t n) {
  if (st.algorithms.size() == n) return;
  for (rdl_handle* a : st.algorithms) rdl_destroy(a);
  st.algorithms.clear();
  for (size_t i = 0; i != n; ++i) {  // SetAlgorithm: algorithms_[i] = front()->Clone()
    rdl_handle* c = nullptr;
    if (rdl_clone(&h, &c) != RDL_OK) throw std::runtime_error("cloning the facet algorithm failed");
    c->synced_mask = h.mask_generation;  // rdl_clone copied mask, RMS image and fitter
    c->synced_rms = h.rms_generation;
    c->synced_fit = h.fit_generation;
    st.algorithms.push_back(c);
  }
}

rdl_result RunAlgorithm(rdl_handle& a, DeviceCube& cube, cudaStream_t stream) {
  switch (a.config.algorithm) {
    case RDL_GENERIC_CLEAN: return RunGenericClean(a, cube, stream);
    case RDL_MULTISCALE: return RunMultiScale(a, cube, stream);
    default: return RunIuwt(a, cube, stream);
  }
}

// aocommon::Image::Resize: dst (w x h) from src (sw x sh), per dimension centred Trim or Untrim
// with start = (outer - inner) / 2
__global__ void ResizeKernel(float* dst, size_t w, size_t h, const float* src, size_t sw, size_t sh) {
  const size_t mw = min(w, sw), mh = min(h, sh);
  const size_t dx0 = (w - mw) / 2, dy0 = (h - mh) / 2, sx0 = (sw - mw) / 2, sy0 = (sh - mh) / 2;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < w * h;
       i += size_t(gridDim.x) * blockDim.x) {
    const size_t y = i / w, x = i - y * w;
    const bool in = x >= dx0 && x < dx0 + mw && y >= dy0 && y < dy0 + mh;
    dst[i] = in ? src[(y - dy0 + sy0) * sw + (x - dx0 + sx0)] : 0.0f;
  }
}

// NearestPsfIndex (parallel_deconvolution.cc:34-55): the first PSF set with the smallest squared
// distance to (x, y); 0 without direction-dependent PSFs
size_t NearestPsfIndex(const DeviceCube& cube, size_t x, size_t y) {
  size_t best = 0;
  uint64_t best_distance = 0;
  for (size_t i = 0; i != cube.n_dd; ++i) {
    const int64_t dx = cube.psf_offset_x[i] - int64_t(x), dy = cube.psf_offset_y[i] - int64_t(y);
    const uint64_t d = uint64_t(dx * dx) + uint64_t(dy * dy);
    if (i == 0 || d < best_distance) {
      best = i;
      best_distance = d;
    }
  }
  return best;
}

// RunSubImage (parallel_deconvolution.cc:294-492) for one facet.
void RunFacet(rdl_handle& h, ParallelState& st, Facet& f, DeviceCube& cube,
              double major_iteration_threshold, bool find_peak_only, cudaStream_t stream) {
  rdl_handle& a = *st.algorithms[f.index];
  const size_t W = cube.width, w = f.width, fh = f.height, n = w * fh;
  st.sub_residual.Reserve(cube.n_images * n);
  st.sub_model.Reserve(cube.n_images * n);
  st.sub_model_start.Reserve(cube.n_images * n);
  st.sub_psf.Reserve(cube.n_channels * n);
  st.boundary_mask.Reserve(n);
  RDL_CUDA(cudaMemcpyAsync(st.boundary_mask.get(), f.boundary_mask.data(), n, cudaMemcpyHostToDevice,
                           stream));
  DeviceCube sub = cube;
  sub.width = w;
  sub.height = fh;
  for (size_t i = 0; i != cube.n_images; ++i) {
    sub.residual.p[i] = st.sub_residual.get() + i * n;
    sub.model.p[i] = st.sub_model.get() + i * n;
    // data_image.Trim(...)
    RDL_CUDA(cudaMemcpy2DAsync(sub.residual.p[i], w * sizeof(float),
                               cube.residual.p[i] + f.y * W + f.x, W * sizeof(float),
                               w * sizeof(float), fh, cudaMemcpyDeviceToDevice, stream));
    // model_image.TrimMasked(..., boundary_mask): pixels other facets own start at zero
    TrimMaskedKernel<<<GridFor(n), kThreads, 0, stream>>>(sub.model.p[i], cube.model.p[i], W, f.x,
                                                          f.y, w, fh, st.boundary_mask.get());
    ++h.launches;
  }
  // the model as it was, to undo a diverging facet
  RDL_CUDA(cudaMemcpyAsync(st.sub_model_start.get(), st.sub_model.get(),
                           cube.n_images * n * sizeof(float), cudaMemcpyDeviceToDevice, stream));
  // psf_image.Resize(sub width, sub height) of the PSF set of this facet (:329): per dimension a
  // centred crop when the PSF is larger than the facet, centred zero padding when it is smaller
  for (size_t c = 0; c != cube.n_channels; ++c) {
    float* dst = st.sub_psf.get() + c * n;
    const float* src = cube.n_dd ? cube.dd_psfs[f.psf_index * cube.n_channels + c] : cube.psf.p[c];
    const size_t pw = cube.n_dd ? cube.psf_width : W, ph = cube.n_dd ? cube.psf_height : cube.height;
    ResizeKernel<<<GridFor(n), kThreads, 0, stream>>>(dst, w, fh, src, pw, ph);
    ++h.launches;
    sub.psf.p[c] = dst;
  }
  sub.vec = (w % 4 == 0) ? 4 : 1;  // the sub buffers are cudaMalloc-aligned, planes are n floats
  if ((n % 4) != 0) sub.vec = 1;

  if (rdl_set_clean_mask(&a, f.mask.data(), w, fh) != RDL_OK) throw std::runtime_error(a.error);
  if (!h.rms_host.empty()) {  // rms_image_.TrimBox(...)
    std::vector<float> crop(n);
    for (size_t row = 0; row != fh; ++row)
      std::copy_n(h.rms_host.data() + (f.y + row) * W + f.x, w, crop.data() + row * w);
    if (rdl_set_rms_factor_image(&a, crop.data(), w, fh) != RDL_OK) throw std::runtime_error(a.error);
  }
  const uint64_t max_iterations = a.settings.max_iterations;
  if (find_peak_only)
    a.settings.max_iterations = 0;
  else
    a.settings.major_iteration_threshold = float(major_iteration_threshold);

  const bool scale_masks_on = (h.use_per_scale_masks || h.track_per_scale_masks) &&
                              a.config.algorithm == RDL_MULTISCALE;
  if (scale_masks_on) {
    // the frame's masks of the previous major iteration, trimmed to this facet (:359-391);
    // they live in the parent handle, where rdl_scale_mask_get finds them
    std::vector<DeviceBuffer<uint8_t>>& frame_masks = MultiScaleMasks(h, cube.Pixels());
    if (!frame_masks.empty()) {
      std::vector<DeviceBuffer<uint8_t>>& masks = MultiScaleMasks(a, n);
      const size_t count = std::max(masks.size(), frame_masks.size());
      while (masks.size() < count) masks.emplace_back();
      st.facet_mask.Reserve(n);
      RDL_CUDA(cudaMemcpyAsync(st.facet_mask.get(), f.mask.data(), n, cudaMemcpyHostToDevice,
                               stream));
      for (size_t i = 0; i != count; ++i) {
        masks[i].Resize(n);
        if (i < frame_masks.size()) {
          ScaleMaskInKernel<<<GridFor(n), kThreads, 0, stream>>>(
              masks[i].get(), frame_masks[i].get(), W, f.x, f.y, w, fh, st.facet_mask.get());
          ++h.launches;
        } else {
          RDL_CUDA(cudaMemsetAsync(masks[i].get(), 0, n, stream));
        }
      }
      RDL_CUDA(cudaStreamSynchronize(stream));  // f.mask is pageable host memory
    }
  }

  const double peak_at_start = std::fabs(f.peak);
  const rdl_result r = RunAlgorithm(a, sub, stream);
  h.launches += a.launches;
  a.launches = 0;
  f.peak = r.final_peak_value;
  f.reached_major_threshold = r.another_iteration_required != 0;
  const double limit = h.settings.divergence_limit;
  const bool converging = (limit == 0.0 || std::fabs(f.peak) <= peak_at_start * limit) &&
                          std::isfinite(f.peak) && !r.is_diverging;
  if (!converging && !find_peak_only) f.reached_major_threshold = false;
  if (!h.rms_host.empty()) rdl_set_rms_factor_image(&a, nullptr, 0, 0);

  if (h.config.track_components && a.config.algorithm == RDL_MULTISCALE) {
    // the facet's components join the frame's list at the facet's offset (:465-479)
    if (converging && !a.components.Empty()) {
      ComponentList& frame = h.components;
      if (frame.Empty()) {
        frame.width = W;
        frame.height = cube.height;
        frame.n_values = cube.n_images;
      }
      if (a.components.scales.size() > frame.scales.size())
        frame.scales.resize(a.components.scales.size());
      for (size_t s = 0; s != a.components.scales.size(); ++s) {
        const ComponentScaleList& list = a.components.scales[s];
        for (size_t k = 0; k != list.x.size(); ++k)
          frame.Add(list.x[k] + f.x, list.y[k] + f.y, s, &list.values[k * frame.n_values]);
      }
    }
    a.components = ComponentList{};
  }
  if (find_peak_only) {
    a.settings.max_iterations = max_iterations;
    return;
  }
  if (scale_masks_on && h.track_per_scale_masks && converging) {
    // the facet's new masks go back into the frame's, inside the pixels it owns (:421-463)
    std::vector<DeviceBuffer<uint8_t>>& frame_masks = MultiScaleMasks(h, cube.Pixels());
    std::vector<DeviceBuffer<uint8_t>>& masks = MultiScaleMasks(a, n);
    const size_t scale_count = rdl_scale_count(&a);
    if (frame_masks.empty()) {
      for (size_t i = 0; i != scale_count; ++i) {
        frame_masks.emplace_back(cube.Pixels());
        RDL_CUDA(cudaMemsetAsync(frame_masks.back().get(), 0, cube.Pixels(), stream));
      }
    }
    for (size_t i = 0; i != scale_count && i < frame_masks.size() && i < masks.size(); ++i) {
      ScaleMaskOutKernel<<<GridFor(n), kThreads, 0, stream>>>(
          frame_masks[i].get(), W, f.x, f.y, w, fh, masks[i].get(), st.boundary_mask.get());
      ++h.launches;
    }
  }
  for (size_t i = 0; i != cube.n_images; ++i) {
    if (converging) {
      // data_image.CopyMasked(*sub_data, x, y, boundary_mask)
      PlaceKernel<false><<<GridFor(n), kThreads, 0, stream>>>(
          cube.residual.p[i], W, f.x, f.y, sub.residual.p[i], w, fh, st.boundary_mask.get());
      ++h.launches;
    }
    // result_model.AddSubImage(*sub_model, x, y); a diverged facet contributes its old model
    const float* m = converging ? sub.model.p[i] : st.sub_model_start.get() + i * n;
    PlaceKernel<true><<<GridFor(n), kThreads, 0, stream>>>(
        st.result_model.get() + i * cube.Pixels(), W, f.x, f.y, m, w, fh, nullptr);
    ++h.launches;
  }
  RDL_CUDA(cudaGetLastError());
}

// What the ranks tell each other about a facet after a pass.
struct FacetWire {
  double peak;
  uint64_t iteration;
  int32_t reached_major_threshold;
  int32_t pad;
};

// Every rank learns the results of the facets the other ranks ran (a no-op on one GPU apart from
// recording the iteration numbers). Rank r sends its facets r, r + world, ... in that order.
void ExchangeFacetResults(rdl_handle& h, ParallelState& st, cudaStream_t stream) {
  const size_t n_facets = st.facets.size();
  const size_t world = h.dist.comm ? size_t(h.dist.world) : 1, rank = h.dist.comm ? size_t(h.dist.rank) : 0;
  for (const Facet& f : st.facets)
    if (f.index % world == rank) st.iterations[f.index] = st.algorithms[f.index]->iteration_number;
  if (world == 1) return;
  const size_t per = (n_facets + world - 1) / world;
  std::vector<FacetWire> send(per, FacetWire{}), recv(per * world);
  for (const Facet& f : st.facets)
    if (f.index % world == rank)
      send[f.index / world] = FacetWire{f.peak, st.iterations[f.index], f.reached_major_threshold ? 1 : 0, 0};
  st.wire_send.Reserve(per * sizeof(FacetWire));
  st.wire_recv.Reserve(per * world * sizeof(FacetWire));
  RDL_CUDA(cudaMemcpyAsync(st.wire_send.get(), send.data(), per * sizeof(FacetWire),
                           cudaMemcpyHostToDevice, stream));
  NcclAllGatherBytes(h.dist, st.wire_send.get(), st.wire_recv.get(), per * sizeof(FacetWire), stream);
  RDL_CUDA(cudaMemcpyAsync(recv.data(), st.wire_recv.get(), per * world * sizeof(FacetWire),
                           cudaMemcpyDeviceToHost, stream));
  RDL_CUDA(cudaStreamSynchronize(stream));
  for (Facet& f : st.facets) {
    const FacetWire& w = recv[(f.index % world) * per + f.index / world];
    f.peak = w.peak;
    f.reached_major_threshold = w.reached_major_threshold != 0;
    st.iterations[f.index] = w.iteration;
  }
}

// After pass 2 every rank holds the new residual in the pixels its facets own and the model of
// its facets in result_model. Each rank zeroes the residual pixels that belong to facets of other
// ranks (rank 0 keeps the pixels no facet owns), then both cubes are summed over the ranks: every
// residual pixel has exactly one non-zero contributor, so the sum reproduces it exactly.
void MergeOverRanks(rdl_handle& h, ParallelState& st, DeviceCube& cube, cudaStream_t stream) {
  const size_t W = cube.width, A = cube.Pixels();
  const size_t world = size_t(h.dist.world), rank = size_t(h.dist.rank);
  st.owned.Reserve(A);
  RDL_CUDA(cudaMemsetAsync(st.owned.get(), rank == 0 ? 1 : 0, A, stream));
  for (const Facet& f : st.facets) {
    const bool is_mine = f.index % world == rank;
    if (is_mine == (rank == 0)) continue;  // rank 0 clears the others' pixels, the rest set their own
    const size_t n = f.width * f.height;
    st.boundary_mask.Reserve(n);
    RDL_CUDA(cudaMemcpyAsync(st.boundary_mask.get(), f.boundary_mask.data(), n,
                             cudaMemcpyHostToDevice, stream));
    MarkOwnedKernel<<<GridFor(n), kThreads, 0, stream>>>(st.owned.get(), W, f.x, f.y, f.width,
                                                         f.height, st.boundary_mask.get(),
                                                         is_mine ? 1 : 0);
    // the staging buffer is reused by the next facet
    RDL_CUDA(cudaStreamSynchronize(stream));
    ++h.launches;
  }
  for (size_t i = 0; i != cube.n_images; ++i) {
    KeepOwnedKernel<<<GridFor(A), kThreads, 0, stream>>>(cube.residual.p[i], st.owned.get(), A);
    ++h.launches;
    NcclAllReduceSumFloat(h.dist, cube.residual.p[i], A, stream);
  }
  NcclAllReduceSumFloat(h.dist, st.result_model.get(), cube.n_images * A, stream);
  RDL_CUDA(cudaGetLastError());
}

}  // namespace

rdl_result RunParallel(rdl_handle& h, DeviceCube& cube, size_t grid_width, size_t grid_height,
                       cudaStream_t stream) {
  if (grid_width == 0 || grid_height == 0) throw std::invalid_argument("empty facet grid");
  const auto call_start = std::chrono::steady_clock::now();
  if (h.dist.band_set) throw std::domain_error("facets of a row-sharded cube are not supported");
  if (!h.parallel) h.parallel.reset(new ParallelState);
  ParallelState& st = *h.parallel;
  const size_t n_facets = grid_width * grid_height;
  if (st.grid_width != grid_width || st.grid_height != grid_height) {
    for (rdl_handle* a : st.algorithms) rdl_destroy(a);
    st.algorithms.clear();
    st.grid_width = grid_width;
    st.grid_height = grid_height;
  }
  EnsureAlgorithms(h, st, n_facets);
  // ParallelDeconvolution::SetThreshold / SetMinorLoopGain / SetAutoMaskMode / ... reach every
  // facet algorithm (parallel_deconvolution.cc:244-292); the iteration counters stay per facet
  for (rdl_handle* a : st.algorithms) {
    a->settings = h.settings;
    a->track_per_scale_masks = h.track_per_scale_masks;
    a->use_per_scale_masks = h.use_per_scale_masks;
    a->trace_capacity = 0;
    // SetSpectralFitter set or changed on the parent after the clones were made
    if (a->synced_fit != h.fit_generation) {
      if (rdl_set_spectral_fitter(a, h.fit_mode, h.fit_terms, h.fit_freqs.data(),
                                  h.fit_weights.data(), h.fit_freqs.size(), h.fit_n_pol) != RDL_OK)
        throw std::runtime_error(a->error);
      a->synced_fit = h.fit_generation;
    }
  }
  if (n_facets == 1) {
    // ExecuteSingleThreadedRun (:505-546): the algorithm on the whole frame, user mask and all.
    // SetCleanMask / SetRmsFactorImage are forwarded on every call (:244-292): a mask or RMS
    // image set, replaced or CLEARED on the parent since the last call reaches the algorithm.
    rdl_handle& a = *st.algorithms.front();
    if (a.synced_mask != h.mask_generation) {
      if (rdl_set_clean_mask(&a, h.mask_host.empty() ? nullptr : h.mask_host.data(), h.mask_w,
                             h.mask_h) != RDL_OK)
        throw std::runtime_error(a.error);
      a.synced_mask = h.mask_generation;
    }
    if (a.synced_rms != h.rms_generation) {
      if (rdl_set_rms_factor_image(&a, h.rms_host.empty() ? nullptr : h.rms_host.data(), h.rms_w,
                                   h.rms_h) != RDL_OK)
        throw std::runtime_error(a.error);
      a.synced_rms = h.rms_generation;
    }
    if (cube.n_dd) {
      // the PSF set nearest to the image centre, padded to the image size (:518-546)
      const size_t index = NearestPsfIndex(cube, cube.width / 2, cube.height / 2);
      const size_t A = cube.Pixels();
      st.sub_psf.Reserve(cube.n_channels * A);
      for (size_t c = 0; c != cube.n_channels; ++c) {
        float* dst = st.sub_psf.get() + c * A;
        ResizeKernel<<<GridFor(A), kThreads, 0, stream>>>(dst, cube.width, cube.height,
                                                          cube.dd_psfs[index * cube.n_channels + c],
                                                          cube.psf_width, cube.psf_height);
        ++h.launches;
        cube.psf.p[c] = dst;
      }
    }
    const rdl_result r = RunAlgorithm(a, cube, stream);
    h.launches += a.launches;
    a.launches = 0;
    h.iteration_number = a.iteration_number;
    st.facets.assign(1, Facet{});
    st.facets[0].width = cube.width;
    st.facets[0].height = cube.height;
    st.facets[0].peak = r.final_peak_value;
    st.iterations.assign(1, a.iteration_number);
    return r;
  }
  if ((h.track_per_scale_masks || h.use_per_scale_masks || h.config.track_components) &&
      h.dist.comm && h.dist.world > 1)
    throw std::domain_error(
        "per-scale masks / component lists under facets sharded over several GPUs are not built");

  const size_t W = cube.width, H = cube.height, A = cube.Pixels();
  // image = data_image.GetLinearIntegrated(); the dividers follow its minimum-power paths
  st.integrated.Reserve(A);
  const IntegSpec spec = MakeLinearSpec(cube);
  LaunchIntegrate(spec, reinterpret_cast<const ConstPtrTable&>(cube.residual), A,
                  st.integrated.get(), cube.vec, stream);
  ++h.launches;
  std::vector<float> integrated(A);
  RDL_CUDA(cudaMemcpyAsync(integrated.data(), st.integrated.get(), A * sizeof(float),
                           cudaMemcpyDeviceToHost, stream));
  RDL_CUDA(cudaStreamSynchronize(stream));
  const auto geometry_start = std::chrono::steady_clock::now();
  // several GPUs: every rank searches its share of the dividers, the divider images are summed
  // over the ranks (the integrated image on the device is free to serve as the buffer)
  DividerSharing sharing;
  if (h.dist.comm && h.dist.world > 1) {
    sharing.rank = h.dist.rank;
    sharing.world = h.dist.world;
    sharing.merge = [&](float* vertical, float* horizontal, size_t n_pixels) {
      for (float* image : {vertical, horizontal}) {
        RDL_CUDA(cudaMemcpyAsync(st.integrated.get(), image, n_pixels * sizeof(float),
                                 cudaMemcpyHostToDevice, stream));
        NcclAllReduceSumFloat(h.dist, st.integrated.get(), n_pixels, stream);
        RDL_CUDA(cudaMemcpyAsync(image, st.integrated.get(), n_pixels * sizeof(float),
                                 cudaMemcpyDeviceToHost, stream));
        RDL_CUDA(cudaStreamSynchronize(stream));
      }
    };
  }
  st.facets = MakeFacets(integrated.data(), W, H,
                         h.mask_host.empty() ? nullptr : h.mask_host.data(), grid_width,
                         grid_height, &sharing);
  for (Facet& f : st.facets)  // :163
    f.psf_index = NearestPsfIndex(cube, f.x + f.width / 2, f.y + f.height / 2);
  st.geometry_ms = MsSince(geometry_start);
  std::vector<float>().swap(integrated);

  // several GPUs: facet i is run by rank i % world on that rank's copy of the cube
  const int world = h.dist.comm ? h.dist.world : 1, rank = h.dist.comm ? h.dist.rank : 0;
  auto mine = [&](const Facet& f) { return int(f.index % size_t(world)) == rank; };
  st.iterations.assign(n_facets, 0);

  st.result_model.Reserve(cube.n_images * A);
  RDL_CUDA(cudaMemsetAsync(st.result_model.get(), 0, cube.n_images * A * sizeof(float), stream));
  // pass 1: every facet reports its peak (max_iterations = 0)
  for (Facet& f : st.facets)
    if (mine(f)) RunFacet(h, st, f, cube, 0.0, true, stream);
  ExchangeFacetResults(h, st, stream);
  double start_peak = 0.0;
  for (const Facet& f : st.facets)
    if (f.peak > start_peak) start_peak = f.peak;  // signed, like the reference (:592-599)
  const double threshold = start_peak * (1.0 - double(h.settings.major_loop_gain));
  // pass 2: deconvolve every facet down to the common major-iteration threshold
  for (Facet& f : st.facets)
    if (mine(f)) RunFacet(h, st, f, cube, threshold, false, stream);
  ExchangeFacetResults(h, st, stream);
  if (world > 1) MergeOverRanks(h, st, cube, stream);
  // model_image.SetImages(resultModel)
  for (size_t i = 0; i != cube.n_images; ++i)
    RDL_CUDA(cudaMemcpyAsync(cube.model.p[i], st.result_model.get() + i * A, A * sizeof(float),
                             cudaMemcpyDeviceToDevice, stream));
  RDL_CUDA(cudaStreamSynchronize(stream));

  size_t finished = 0;
  bool reached_max_iterations = false;
  double end_peak = 0.0;
  for (const Facet& f : st.facets) {
    if (!f.reached_major_threshold) ++finished;
    if (st.iterations[f.index] >= h.settings.max_iterations) reached_max_iterations = true;
    end_peak = std::max(end_peak, f.peak);
  }
  rdl_result r{};
  r.has_starting_peak = 1;
  r.starting_peak_value = float(start_peak);
  r.final_peak_value = float(end_peak);
  r.another_iteration_required = (finished != st.facets.size()) && !reached_max_iterations;
  r.is_diverging = 0;
  h.iteration_number = st.iterations.front();
  r.iteration_number = h.iteration_number;
  st.total_ms = MsSince(call_start);
  return r;
}

void ParallelTiming(const rdl_handle& h, double& geometry_ms, double& total_ms) {
  geometry_ms = h.parallel ? h.parallel->geometry_ms : 0.0;
  total_ms = h.parallel ? h.parallel->total_ms : 0.0;
}

// ParallelDeconvolution::GetComponentList (:184-207): without facets the list is the single
// algorithm's own.
rdl_handle* ParallelSingleAlgorithm(const rdl_handle& h) {
  if (!h.parallel || h.parallel->algorithms.size() != 1) return nullptr;
  return h.parallel->algorithms.front();
}

size_t ParallelFacetCount(const rdl_handle& h) { return h.parallel ? h.parallel->facets.size() : 0; }

FacetInfo ParallelFacet(const rdl_handle& h, size_t i) {
  if (!h.parallel || i >= h.parallel->facets.size()) throw std::invalid_argument("no such facet");
  const Facet& f = h.parallel->facets[i];
  const uint64_t it = i < h.parallel->iterations.size() ? h.parallel->iterations[i] : 0;
  return FacetInfo{f.x, f.y, f.width, f.height, f.peak, f.reached_major_threshold ? 1 : 0, it};
}

const std::vector<uint8_t>& ParallelFacetBoundaryMask(const rdl_handle& h, size_t i) {
  if (!h.parallel || i >= h.parallel->facets.size()) throw std::invalid_argument("no such facet");
  return h.parallel->facets[i].boundary_mask;
}

}  // namespace rdl
