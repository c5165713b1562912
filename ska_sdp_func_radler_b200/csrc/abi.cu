// extern "C" boundary (include/radler_b200.h). Exceptions never cross it: they are
// turned into status codes + rdl_last_error().
#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstring>
#include <new>

#include "fft.cuh"
#include "iuwt.cuh"
#include "facets.h"
#include "parallel.cuh"
#include "handle.h"
#include "multiscale.cuh"
#include "image_statistics.cuh"
#include "rms_image.cuh"

using namespace rdl;

namespace {

thread_local std::string g_create_error;
std::atomic<long> g_live_handles{0};

template <typename F>
rdl_status Guard(rdl_handle* h, F&& f) {
  try {
    f();
    return RDL_OK;
  } catch (const CudaError& e) {
    if (h) h->error = e.what();
    cudaGetLastError();
    return std::strstr(e.what(), "cudaMalloc") ? RDL_ERR_OUT_OF_MEMORY : RDL_ERR_CUDA;
  } catch (const std::invalid_argument& e) {
    if (h) h->error = e.what();
    return RDL_ERR_INVALID_ARGUMENT;
  } catch (const std::domain_error& e) {
    if (h) h->error = e.what();
    return RDL_ERR_UNSUPPORTED;
  } catch (const std::bad_alloc&) {
    if (h) h->error = "host allocation failed";
    return RDL_ERR_OUT_OF_MEMORY;
  } catch (const std::exception& e) {
    if (h) h->error = e.what();
    return RDL_ERR_INVALID_ARGUMENT;
  }
}

bool DeviceUsable(int device, std::string& why) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess || n == 0) {
    cudaGetLastError();
    why = "no CUDA device available (this library has no CPU fallback)";
    return false;
  }
  if (device < 0 || device >= n) {
    why = "device ordinal out of range";
    return false;
  }
  cudaDeviceProp prop{};
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) {
    cudaGetLastError();
    why = "cudaGetDeviceProperties failed";
    return false;
  }
  if (prop.major != 10) {
    why = std::string("device '") + prop.name + "' is not sm_100 (kernels are built for sm_100a only)";
    return false;
  }
  return true;
}

void CheckCube(const rdl_cube_desc* c) {
  if (!c) throw std::invalid_argument("cube descriptor is null");
  if (c->width == 0 || c->height == 0) throw std::invalid_argument("empty image");
  if (c->n_channels == 0 || c->n_pol == 0) throw std::invalid_argument("cube has no images");
  if (c->n_channels * c->n_pol > size_t(kMaxImages))
    throw std::domain_error("more than 64 images per cube are not supported");
  if (c->width * c->height >= (1ull << 32))
    throw std::domain_error("images of 2^32 pixels or more are not supported");
}

void FillCubeMeta(DeviceCube& cube, const rdl_cube_desc* c) {
  cube.width = c->width;
  cube.height = c->height;
  cube.n_channels = c->n_channels;
  cube.n_pol = c->n_pol;
  cube.n_images = c->n_channels * c->n_pol;
  cube.weights.assign(c->n_channels, 1.0f);
  if (c->channel_weights) cube.weights.assign(c->channel_weights, c->channel_weights + c->n_channels);
  cube.pol_linked.assign(c->n_pol, 1);
  if (c->pol_linked) cube.pol_linked.assign(c->pol_linked, c->pol_linked + c->n_pol);
  cube.pol_norm = c->pol_norm;
  cube.squared_joins = c->squared_joins != 0;
}

bool Aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

void SetVec(DeviceCube& cube) {
  bool ok = cube.width % 4 == 0;
  for (size_t i = 0; ok && i != cube.n_images; ++i)
    ok = Aligned16(cube.residual.p[i]) && (!cube.model.p[i] || Aligned16(cube.model.p[i]));
  for (size_t i = 0; ok && i != cube.n_channels; ++i)
    ok = !cube.psf.p[i] || Aligned16(cube.psf.p[i]);
  cube.vec = ok ? 4 : 1;
}

// cudaHostRegister a caller's plane once, so that its copies run at PCIe rate and overlap; a plane
// that cannot be registered (already pinned by the caller, not mappable) is simply copied as it is
void PinPlane(rdl_handle* h, const void* p, size_t bytes) {
  if (!h->pin_host_buffers || !p) return;
  auto it = h->pinned.find(p);
  if (it != h->pinned.end() && it->second >= bytes) return;
  if (it != h->pinned.end()) {
    cudaHostUnregister(const_cast<void*>(p));
    h->pinned.erase(it);
  }
  if (cudaHostRegister(const_cast<void*>(p), bytes, cudaHostRegisterDefault) == cudaSuccess)
    h->pinned[p] = bytes;
  else
    cudaGetLastError();
}

void UnpinAll(rdl_handle* h) {
  for (const auto& kv : h->pinned) cudaHostUnregister(const_cast<void*>(kv.first));
  cudaGetLastError();
  h->pinned.clear();
}

// Device (or managed) memory? Host pointers CUDA does not know about are host memory.
bool IsDevicePointer(const void* p) {
  cudaPointerAttributes attributes{};
  if (cudaPointerGetAttributes(&attributes, p) != cudaSuccess) {
    cudaGetLastError();
    return false;
  }
  return attributes.type == cudaMemoryTypeDevice || attributes.type == cudaMemoryTypeManaged;
}

cudaStream_t StreamOf(rdl_handle* h, void* stream) {
  return stream ? static_cast<cudaStream_t>(stream) : h->own_stream;
}

}  // namespace

extern "C" {

void rdl_config_default(rdl_config* cfg) {
  std::memset(cfg, 0, sizeof(*cfg));
  cfg->algorithm = RDL_GENERIC_CLEAN;
  cfg->device = 0;
  cfg->use_sub_minor_optimization = 1;
  cfg->ms_fast_sub_minor_loop = 1;
  cfg->ms_sub_minor_loop_gain = 0.2;
  cfg->ms_scale_bias = 0.6;
  cfg->ms_max_scales = 0;
  cfg->ms_convolution_padding = 1.1;
  cfg->ms_shape = RDL_SHAPE_TAPERED_QUADRATIC;
  cfg->beam_size = 0.0;
  cfg->pixel_scale_x = cfg->pixel_scale_y = 1.0;
  cfg->track_components = 0;
}

rdl_status rdl_create(const rdl_config* cfg, rdl_handle** out) {
  if (!cfg || !out) return RDL_ERR_INVALID_ARGUMENT;
  *out = nullptr;
  std::string why;
  if (!DeviceUsable(cfg->device, why)) {
    g_create_error = why;
    return RDL_ERR_NO_DEVICE;
  }
  if (cfg->algorithm != RDL_GENERIC_CLEAN && cfg->algorithm != RDL_MULTISCALE &&
      cfg->algorithm != RDL_IUWT) {
    g_create_error = "algorithm not supported by this build";
    return RDL_ERR_UNSUPPORTED;
  }
  rdl_handle* h = new (std::nothrow) rdl_handle;
  if (!h) return RDL_ERR_OUT_OF_MEMORY;
  h->config = *cfg;
  if (cfg->ms_scale_list && cfg->ms_n_scale_list)
    h->ms_scale_list.assign(cfg->ms_scale_list, cfg->ms_scale_list + cfg->ms_n_scale_list);
  h->config.ms_scale_list = nullptr;
  h->device = cfg->device;
  // MultiScaleAlgorithm ctor (multiscale_algorithm.cc:153-163)
  h->beam_size_in_pixels = cfg->beam_size / std::max(cfg->pixel_scale_x, cfg->pixel_scale_y);
  if (!(h->beam_size_in_pixels > 0.0)) h->beam_size_in_pixels = 1.0;
  const rdl_status st = Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    // a BLOCKING stream: with stream == NULL the calls are ordered after the caller's
    // legacy-default-stream work that produced RDL_MEM_DEVICE planes (torch's default stream)
    // and after the synchronous cudaMemcpy uploads of mask / RMS image / projection
    RDL_CUDA(cudaStreamCreateWithFlags(&h->own_stream, cudaStreamDefault));
  });
  if (st != RDL_OK) {
    g_create_error = h->error;
    delete h;
    return st;
  }
  g_live_handles.fetch_add(1);
  *out = h;
  return RDL_OK;
}

rdl_status rdl_clone(const rdl_handle* src, rdl_handle** out) {
  if (!src || !out) return RDL_ERR_INVALID_ARGUMENT;
  rdl_config cfg = src->config;
  cfg.ms_scale_list = src->ms_scale_list.data();
  cfg.ms_n_scale_list = src->ms_scale_list.size();
  rdl_status st = rdl_create(&cfg, out);
  if (st != RDL_OK) return st;
  rdl_handle* h = *out;
  st = Guard(h, [&] {
    h->settings = src->settings;
    h->iteration_number = src->iteration_number;
    h->track_per_scale_masks = src->track_per_scale_masks;
    h->use_per_scale_masks = src->use_per_scale_masks;
    h->trace_capacity = src->trace_capacity;
    if (!src->mask_host.empty()) {
      st = rdl_set_clean_mask(h, src->mask_host.data(), src->mask_w, src->mask_h);
      if (st != RDL_OK) throw std::runtime_error(h->error);
    }
    if (!src->rms_host.empty()) {
      st = rdl_set_rms_factor_image(h, src->rms_host.data(), src->rms_w, src->rms_h);
      if (st != RDL_OK) throw std::runtime_error(h->error);
    }
    if (src->fit_mode != RDL_FIT_NONE) {
      st = rdl_set_spectral_fitter(h, src->fit_mode, src->fit_terms, src->fit_freqs.data(),
                                   src->fit_weights.data(), src->fit_freqs.size(), src->fit_n_pol);
      if (st != RDL_OK) throw std::runtime_error(h->error);
    } else {
      h->fit_n_pol = src->fit_n_pol;
    }
    if (src->ms) h->ms = CloneMultiScaleState(*src->ms, h->own_stream);
    h->components = src->components;  // cloned_ptr<ComponentList>
  });
  if (st != RDL_OK) {
    rdl_destroy(h);
    *out = nullptr;
  }
  return st;
}

void rdl_destroy(rdl_handle* h) {
  if (!h) return;
  cudaSetDevice(h->device);
  h->ms.reset();
  try {
    NcclDestroy(h->dist);
  } catch (...) {
  }
  UnpinAll(h);
  if (h->own_stream) cudaStreamDestroy(h->own_stream);
  if (h->upload_stream) cudaStreamDestroy(h->upload_stream);
  if (h->upload_event) cudaEventDestroy(h->upload_event);
  if (h->model_ready) cudaEventDestroy(h->model_ready);
  delete h;
  // the last handle takes the cached device memory with it
  if (g_live_handles.fetch_sub(1) == 1) DeviceTrim();
}

void rdl_release_cached_memory(void) { DeviceTrim(); }

const char* rdl_last_error(const rdl_handle* h) {
  return h ? h->error.c_str() : g_create_error.c_str();
}

#define RDL_SETTER(name, expr)            \
  rdl_status name {                       \
    if (!h) return RDL_ERR_INVALID_ARGUMENT; \
    expr;                                 \
    return RDL_OK;                        \
  }
RDL_SETTER(rdl_set_max_iterations(rdl_handle* h, uint64_t n), h->settings.max_iterations = n)
RDL_SETTER(rdl_set_threshold(rdl_handle* h, float v), h->settings.threshold = v)
RDL_SETTER(rdl_set_major_iteration_threshold(rdl_handle* h, float v),
           h->settings.major_iteration_threshold = v)
RDL_SETTER(rdl_set_minor_loop_gain(rdl_handle* h, float v), h->settings.minor_loop_gain = v)
RDL_SETTER(rdl_set_major_loop_gain(rdl_handle* h, float v), h->settings.major_loop_gain = v)
RDL_SETTER(rdl_set_allow_negative_components(rdl_handle* h, int32_t v),
           h->settings.allow_negative_components = v != 0)
RDL_SETTER(rdl_set_stop_on_negative_components(rdl_handle* h, int32_t v),
           h->settings.stop_on_negative_component = v != 0)
RDL_SETTER(rdl_set_clean_border_ratio(rdl_handle* h, float v), h->settings.clean_border_ratio = v)
RDL_SETTER(rdl_set_divergence_limit(rdl_handle* h, float v), h->settings.divergence_limit = v)
RDL_SETTER(rdl_set_iteration_number(rdl_handle* h, uint64_t n), h->iteration_number = n)
RDL_SETTER(rdl_set_trace_capacity(rdl_handle* h, uint64_t n), h->trace_capacity = n)
RDL_SETTER(rdl_set_psfs_constant(rdl_handle* h, int32_t v),
           (h->psfs_constant = v != 0, h->staged_psf_ptrs.clear(),
            h->ms ? MultiScaleInvalidatePsfCache(*h->ms) : void()))
RDL_SETTER(rdl_set_pin_host_buffers(rdl_handle* h, int32_t v),
           (h->pin_host_buffers = v != 0, (v == 0 ? UnpinAll(h) : void())))
RDL_SETTER(rdl_set_auto_mask_mode(rdl_handle* h, int32_t track, int32_t use),
           (h->track_per_scale_masks = track != 0, h->use_per_scale_masks = use != 0))
#undef RDL_SETTER

rdl_status rdl_set_clean_mask(rdl_handle* h, const uint8_t* mask, uint64_t width, uint64_t height) {
  if (!h) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    ++h->mask_generation;
    if (!mask) {
      h->mask.Free();
      h->mask_host.clear();
      h->mask_w = h->mask_h = 0;
      return;
    }
    h->mask.Resize(width * height);
    if (IsDevicePointer(mask)) {  // e.g. the output of rdl_auto_mask_from_model
      h->mask_host.resize(width * height);
      RDL_CUDA(cudaMemcpy(h->mask_host.data(), mask, width * height, cudaMemcpyDeviceToHost));
      RDL_CUDA(cudaMemcpy(h->mask.get(), mask, width * height, cudaMemcpyDeviceToDevice));
    } else {
      h->mask_host.assign(mask, mask + width * height);
      RDL_CUDA(cudaMemcpy(h->mask.get(), mask, width * height, cudaMemcpyHostToDevice));
    }
    h->mask_w = width;
    h->mask_h = height;
  });
}

rdl_status rdl_set_rms_factor_image(rdl_handle* h, const float* image, uint64_t width,
                                    uint64_t height) {
  if (!h) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    ++h->rms_generation;
    if (!image) {
      h->rms.Free();
      h->rms_host.clear();
      h->rms_w = h->rms_h = 0;
      return;
    }
    h->rms.Resize(width * height);
    if (IsDevicePointer(image)) {  // e.g. the output of rdl_rms_make_factor_image
      h->rms_host.resize(width * height);
      RDL_CUDA(cudaMemcpy(h->rms_host.data(), image, width * height * sizeof(float),
                          cudaMemcpyDeviceToHost));
      RDL_CUDA(cudaMemcpy(h->rms.get(), image, width * height * sizeof(float),
                          cudaMemcpyDeviceToDevice));
    } else {
      h->rms_host.assign(image, image + width * height);
      RDL_CUDA(cudaMemcpy(h->rms.get(), image, width * height * sizeof(float),
                          cudaMemcpyHostToDevice));
    }
    h->rms_w = width;
    h->rms_h = height;
  });
}

rdl_status rdl_set_spectral_fitter(rdl_handle* h, int32_t mode, uint64_t n_terms,
                                   const double* frequencies, const float* weights,
                                   uint64_t n_channels, uint64_t n_polarizations) {
  if (!h) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    if (mode != RDL_FIT_NONE && mode != RDL_FIT_POLYNOMIAL)
      throw std::domain_error("only kNoFitting and kPolynomial spectral fitting are supported");
    if (n_polarizations == 0) throw std::invalid_argument("n_polarizations must be >= 1");
    ++h->fit_generation;
    h->fit_mode = mode;
    h->fit_terms = n_terms;
    h->fit_n_pol = n_polarizations;
    h->fit_freqs.clear();
    h->fit_weights.clear();
    h->projection_host.clear();
    if (mode == RDL_FIT_NONE) return;
    if (!frequencies || n_channels == 0 || n_channels > size_t(kMaxImages))
      throw std::invalid_argument("spectral fitter needs 1..64 channel frequencies");
    h->fit_freqs.assign(frequencies, frequencies + n_channels);
    h->fit_weights.assign(n_channels, 1.0f);
    if (weights) h->fit_weights.assign(weights, weights + n_channels);
    h->projection_host = PolynomialProjection(n_terms, h->fit_freqs, h->fit_weights);
    RDL_CUDA(cudaSetDevice(h->device));
    h->projection.Resize(h->projection_host.size());
    RDL_CUDA(cudaMemcpy(h->projection.get(), h->projection_host.data(),
                        h->projection_host.size() * sizeof(double), cudaMemcpyHostToDevice));
  });
}

uint64_t rdl_iteration_number(const rdl_handle* h) { return h ? h->iteration_number : 0; }
uint64_t rdl_trace_size(const rdl_handle* h) { return h ? h->trace_host.size() : 0; }
uint64_t rdl_trace_get(const rdl_handle* h, rdl_trace_entry* out, uint64_t capacity) {
  if (!h || !out) return 0;
  const uint64_t n = std::min<uint64_t>(capacity, h->trace_host.size());
  static_assert(sizeof(rdl_trace_entry) == sizeof(TraceEntry), "trace layout");
  std::memcpy(out, h->trace_host.data(), n * sizeof(TraceEntry));
  return n;
}
uint64_t rdl_last_selected_count(const rdl_handle* h) { return h ? h->last_selected : 0; }
uint64_t rdl_minor_iteration_count(const rdl_handle* h) { return h ? h->minor_iterations : 0; }
uint64_t rdl_kernel_launch_count(const rdl_handle* h) { return h ? h->launches : 0; }
uint64_t rdl_reference_convolutions(const rdl_handle* h, uint64_t* triples, uint64_t capacity) {
  if (!h) return 0;
  uint64_t n = 0;
  for (const auto& kv : h->reference_convolutions) {
    if (triples && n < capacity) {
      triples[3 * n] = kv.first.first;
      triples[3 * n + 1] = kv.first.second;
      triples[3 * n + 2] = kv.second;
    }
    ++n;
  }
  return n;
}

uint64_t rdl_scale_count(const rdl_handle* h) { return h && h->ms ? MultiScaleScaleCount(*h->ms) : 0; }
rdl_status rdl_scale_info_get(const rdl_handle* h, uint64_t i, rdl_scale_info* out) {
  if (!h || !out || !h->ms || i >= MultiScaleScaleCount(*h->ms)) return RDL_ERR_INVALID_ARGUMENT;
  *out = MultiScaleScaleInfo(*h->ms, i);
  return RDL_OK;
}
}  // extern "C"
namespace {
// the handle's own list, or the list of the single algorithm behind a 1 x 1 facet run
rdl::ComponentList& ComponentsOf(rdl_handle* h) {
  if (h->components.Empty())
    if (rdl_handle* single = ParallelSingleAlgorithm(*h)) return single->components;
  return h->components;
}
}  // namespace
extern "C" {

uint64_t rdl_component_scale_count(const rdl_handle* h) {
  return h ? ComponentsOf(const_cast<rdl_handle*>(h)).scales.size() : 0;
}

rdl_status rdl_component_count(rdl_handle* h, uint64_t scale_index, uint64_t* count) {
  if (!h || !count || scale_index >= ComponentsOf(h).scales.size()) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    ComponentsOf(h).MergeDuplicates();
    *count = ComponentsOf(h).scales[scale_index].x.size();
  });
}

rdl_status rdl_component_get(rdl_handle* h, uint64_t scale_index, uint64_t first, uint64_t count,
                             uint64_t* x, uint64_t* y, float* values) {
  if (!h || !x || !y || !values || scale_index >= ComponentsOf(h).scales.size())
    return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    rdl::ComponentList& components = ComponentsOf(h);
    components.MergeDuplicates();
    const rdl::ComponentScaleList& list = components.scales[scale_index];
    if (first + count > list.x.size()) throw std::invalid_argument("component range out of bounds");
    const size_t n_values = components.n_values;
    for (uint64_t k = 0; k != count; ++k) {
      x[k] = list.x[first + k];
      y[k] = list.y[first + k];
    }
    std::copy_n(list.values.begin() + first * n_values, count * n_values, values);
  });
}

void rdl_component_list_clear(rdl_handle* h) {
  if (!h) return;
  if (rdl_handle* single = ParallelSingleAlgorithm(*h)) single->components = rdl::ComponentList{};
  h->components = rdl::ComponentList{};
}

rdl_status rdl_scale_mask_get(const rdl_handle* h, uint64_t i, uint8_t* out) {
  if (!h || !out || !h->ms) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(const_cast<rdl_handle*>(h), [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    MultiScaleCopyScaleMask(*h->ms, i, out);
  });
}

namespace {
struct DdPsfArgs {
  uint64_t n_dd = 0, width = 0, height = 0;
  const int64_t* offset_x = nullptr;
  const int64_t* offset_y = nullptr;
};

// grid_width = 0: the algorithm itself; otherwise ParallelDeconvolution over a facet grid
rdl_status ExecuteImpl(rdl_handle* h, const rdl_cube_desc* c, float* const* residual,
                       float* const* model, const float* const* psfs, void* stream_v,
                       rdl_result* out, uint64_t grid_width, uint64_t grid_height,
                       const DdPsfArgs& dd = DdPsfArgs()) {
  if (!h || !out) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    CheckCube(c);
    if (dd.n_dd != 0 && (dd.width == 0 || dd.height == 0 || dd.width > c->width ||
                         dd.height > c->height || !dd.offset_x || !dd.offset_y))
      throw std::invalid_argument("direction-dependent PSFs need a size within the image and offsets");
    if (!residual || !model || !psfs) throw std::invalid_argument("null image table");
    const bool sharded = h->dist.band_set;
    const uint64_t frame_height = sharded ? h->dist.global_height : c->height;
    if (c->width * frame_height >= (1ull << 32))  // peak keys carry a 32-bit frame index
      throw std::domain_error("frames of 2^32 pixels or more are not supported");
    if (sharded && (h->config.algorithm != RDL_GENERIC_CLEAN ||
                    (h->dist.block_rows
                         ? c->height * uint64_t(h->dist.world) != h->dist.global_height
                         : h->dist.y_offset + c->height > h->dist.global_height)))
      throw std::invalid_argument(sharded && h->config.algorithm != RDL_GENERIC_CLEAN
                                      ? "row bands are supported by GenericClean only"
                                      : "row band exceeds the frame");
    if (h->mask.get() && (h->mask_w != c->width || h->mask_h != frame_height))
      throw std::invalid_argument("clean mask dimensions do not match the image");
    if (h->rms.get() && (h->rms_w != c->width || h->rms_h != frame_height))
      throw std::invalid_argument("Error in RMS factor image dimensions!");
    if (h->fit_mode == RDL_FIT_POLYNOMIAL &&
        (h->fit_freqs.size() != c->n_channels || h->fit_n_pol != c->n_pol))
      throw std::invalid_argument("spectral fitter channel/polarisation count does not match the cube");
    RDL_CUDA(cudaSetDevice(h->device));
    cudaStream_t stream = StreamOf(h, stream_v);
    DeviceCube cube;
    FillCubeMeta(cube, c);
    const size_t n = cube.Pixels(), bytes = n * sizeof(float);
    const bool host = c->memory_space == RDL_MEM_HOST;
    // GenericClean without the sub-minor loop touches one model pixel per image and component:
    // with host buffers the components are logged on the device and applied to the caller's
    // model planes here, in order, with the same fp32 adds -- the model planes never cross PCIe
    // (8192^2 x 16: 8.6 GB less per call). Bounded by the size of the log.
    const uint64_t log_capacity = h->settings.max_iterations > h->iteration_number
                                      ? h->settings.max_iterations - h->iteration_number
                                      : 0;
    h->model_log = host && grid_width == 0 && !sharded &&
                   h->config.algorithm == RDL_GENERIC_CLEAN &&
                   !h->config.use_sub_minor_optimization &&
                   log_capacity * (c->n_channels * c->n_pol + 1) * sizeof(float) <= (size_t(256) << 20);
    const bool model_on_host = h->model_log;
    // multiscale first touches the model after its first sub-minor loop: the model planes follow
    // the residual and PSF planes on a second stream while the call is already at work
    const bool model_behind = host && grid_width == 0 && h->config.algorithm == RDL_MULTISCALE;
    struct UploadGuard {  // a failing call must not leave copies from the caller's planes in flight
      rdl_handle* h;
      ~UploadGuard() {
        if (h->model_pending && h->upload_stream) cudaStreamSynchronize(h->upload_stream);
        h->model_pending = false;
      }
    } upload_guard{h};
    if (model_behind && !h->upload_stream) {
      RDL_CUDA(cudaStreamCreateWithFlags(&h->upload_stream, cudaStreamNonBlocking));
      RDL_CUDA(cudaEventCreateWithFlags(&h->upload_event, cudaEventDisableTiming));
      RDL_CUDA(cudaEventCreateWithFlags(&h->model_ready, cudaEventDisableTiming));
    }
    if (host) {
      h->stage_residual.Reserve(cube.n_images * n);
      if (!model_on_host) h->stage_model.Reserve(cube.n_images * n);
      const size_t n_psf_planes = dd.n_dd ? dd.n_dd * cube.n_channels : cube.n_channels;
      const size_t psf_n = dd.n_dd ? dd.width * dd.height : cube.width * frame_height;
      const size_t psf_bytes = psf_n * sizeof(float);
      h->stage_psf.Reserve(n_psf_planes * psf_n);
      for (size_t i = 0; i != cube.n_images; ++i) {
        cube.residual.p[i] = h->stage_residual.get() + i * n;
        cube.model.p[i] = model_on_host ? nullptr : h->stage_model.get() + i * n;
        PinPlane(h, residual[i], bytes);
        RDL_CUDA(cudaMemcpyAsync(cube.residual.p[i], residual[i], bytes, cudaMemcpyHostToDevice, stream));
        if (model_on_host) continue;
        PinPlane(h, model[i], bytes);
        if (!model_behind)
          RDL_CUDA(cudaMemcpyAsync(cube.model.p[i], model[i], bytes, cudaMemcpyHostToDevice, stream));
      }
      // the PSFs do not change between the major iterations of a run (radler.cc:175-268 passes the
      // same psf_images every time): with rdl_set_psfs_constant they cross PCIe once
      bool staged = h->psfs_constant && h->staged_psf_pixels == psf_n &&
                    h->staged_psf_ptrs.size() == n_psf_planes;
      for (size_t i = 0; staged && i != n_psf_planes; ++i) staged = h->staged_psf_ptrs[i] == psfs[i];
      // new pixels under the same staging addresses: what multiscale kept of the old ones goes
      if (!staged && h->ms) MultiScaleInvalidatePsfCache(*h->ms);
      // Row-sharded runs: every rank is handed the same full-frame PSFs. Each plane crosses PCIe
      // on ONE rank (plane i on rank i % world) and reaches the others over NVLink, so the PSF
      // upload of the job is one copy of the PSFs instead of one per GPU.
      const bool share_upload = sharded && h->dist.comm && h->dist.world > 1 && !dd.n_dd;
      std::vector<float*> shared_planes;
      std::vector<int> shared_roots;
      for (size_t i = 0; i != n_psf_planes; ++i) {
        float* dst = h->stage_psf.get() + i * psf_n;
        const int root = share_upload ? int(i % size_t(h->dist.world)) : h->dist.rank;
        if (!staged && root == h->dist.rank) {
          PinPlane(h, psfs[i], psf_bytes);
          RDL_CUDA(cudaMemcpyAsync(dst, psfs[i], psf_bytes, cudaMemcpyHostToDevice, stream));
        }
        if (!staged && share_upload) {
          shared_planes.push_back(dst);
          shared_roots.push_back(root);
        }
        if (dd.n_dd)
          cube.dd_psfs.push_back(dst);
        else
          cube.psf.p[i] = dst;
      }
      if (!shared_planes.empty())
        NcclBroadcastPlanes(h->dist, shared_planes.data(), shared_roots.data(), shared_planes.size(),
                            psf_n, stream);
      h->staged_psf_ptrs.assign(psfs, psfs + n_psf_planes);
      h->staged_psf_pixels = psf_n;
      if (model_behind) {
        // behind everything queued so far on the call's stream (earlier calls' reads of the
        // staging planes, this call's residual and PSF uploads)
        RDL_CUDA(cudaEventRecord(h->upload_event, stream));
        RDL_CUDA(cudaStreamWaitEvent(h->upload_stream, h->upload_event, 0));
        for (size_t i = 0; i != cube.n_images; ++i)
          RDL_CUDA(cudaMemcpyAsync(cube.model.p[i], model[i], bytes, cudaMemcpyHostToDevice,
                                   h->upload_stream));
        RDL_CUDA(cudaEventRecord(h->model_ready, h->upload_stream));
        h->model_pending = true;
      }
    } else {
      for (size_t i = 0; i != cube.n_images; ++i) {
        cube.residual.p[i] = residual[i];
        cube.model.p[i] = model[i];
      }
      if (dd.n_dd)
        cube.dd_psfs.assign(psfs, psfs + dd.n_dd * cube.n_channels);
      else
        for (size_t i = 0; i != cube.n_channels; ++i) cube.psf.p[i] = psfs[i];
    }
    if (dd.n_dd) {
      cube.n_dd = dd.n_dd;
      cube.psf_width = dd.width;
      cube.psf_height = dd.height;
      cube.psf_offset_x.assign(dd.offset_x, dd.offset_x + dd.n_dd);
      cube.psf_offset_y.assign(dd.offset_y, dd.offset_y + dd.n_dd);
    }
    SetVec(cube);
    rdl_result r{};
    if (grid_width != 0) {
      r = RunParallel(*h, cube, grid_width, grid_height, stream);
    } else {
    if (h->config.algorithm == RDL_GENERIC_CLEAN)
      r = RunGenericClean(*h, cube, stream);
    else if (h->config.algorithm == RDL_MULTISCALE)
      r = RunMultiScale(*h, cube, stream);
    else
      r = RunIuwt(*h, cube, stream);
    }
    if (h->model_pending) {  // the algorithm never needed the model: the download below does
      RDL_CUDA(cudaStreamWaitEvent(stream, h->model_ready, 0));
      h->model_pending = false;
    }
    if (host) {
      for (size_t i = 0; i != cube.n_images; ++i) {
        RDL_CUDA(cudaMemcpyAsync(residual[i], cube.residual.p[i], bytes, cudaMemcpyDeviceToHost, stream));
        if (!model_on_host)
          RDL_CUDA(cudaMemcpyAsync(model[i], cube.model.p[i], bytes, cudaMemcpyDeviceToHost, stream));
      }
      if (model_on_host) {  // while the residual planes are on their way back
        const uint64_t count = h->model_log_count;
        for (uint64_t k = 0; k != count; ++k) {
          const size_t index = h->log_index_host[k];
          const float* v = h->log_values_host.data() + k * cube.n_images;
          for (size_t i = 0; i != cube.n_images; ++i) {
            volatile float sum = model[i][index] + v[i];  // one rounded fp32 add, as on the device
            model[i][index] = sum;
          }
        }
      }
    }
    h->model_log = false;
    RDL_CUDA(cudaStreamSynchronize(stream));
    *out = r;
  });
}
}  // namespace

rdl_status rdl_execute_major_iteration(rdl_handle* h, const rdl_cube_desc* c,
                                       float* const* residual, float* const* model,
                                       const float* const* psfs, void* stream_v, rdl_result* out) {
  return ExecuteImpl(h, c, residual, model, psfs, stream_v, out, 0, 0);
}

rdl_status rdl_parallel_execute_major_iteration(rdl_handle* h, uint64_t grid_width,
                                                uint64_t grid_height, const rdl_cube_desc* c,
                                                float* const* residual, float* const* model,
                                                const float* const* psfs, void* stream_v,
                                                rdl_result* out) {
  if (grid_width == 0 || grid_height == 0) return RDL_ERR_INVALID_ARGUMENT;
  return ExecuteImpl(h, c, residual, model, psfs, stream_v, out, grid_width, grid_height);
}

rdl_status rdl_parallel_execute_major_iteration_dd(
    rdl_handle* h, uint64_t grid_width, uint64_t grid_height, const rdl_cube_desc* c,
    float* const* residual, float* const* model, const float* const* psfs, uint64_t n_dd,
    uint64_t psf_width, uint64_t psf_height, const int64_t* psf_offset_x,
    const int64_t* psf_offset_y, void* stream_v, rdl_result* out) {
  if (grid_width == 0 || grid_height == 0 || n_dd == 0) return RDL_ERR_INVALID_ARGUMENT;
  DdPsfArgs dd;
  dd.n_dd = n_dd;
  dd.width = psf_width;
  dd.height = psf_height;
  dd.offset_x = psf_offset_x;
  dd.offset_y = psf_offset_y;
  return ExecuteImpl(h, c, residual, model, psfs, stream_v, out, grid_width, grid_height, dd);
}

// ---- facet geometry (host only; math/dijkstra_splitter.h) -----------------------------------
rdl_status rdl_facet_dividing_path(const float* image, float* output, uint64_t width,
                                   uint64_t height, uint64_t lo, uint64_t hi, int32_t vertical) {
  if (!image || !output) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(nullptr, [&] { DividingPath(image, output, width, height, lo, hi, vertical != 0); });
}

rdl_status rdl_facet_flood_area(const float* dividers, uint64_t width, uint64_t height,
                                uint64_t inside, int32_t vertical, uint8_t* mask, uint64_t* start,
                                uint64_t* extent) {
  if (!dividers || !mask || !start || !extent) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(nullptr, [&] {
    size_t s = 0, e = 0;
    FloodArea(dividers, width, height, inside, vertical != 0, mask, s, e);
    *start = s;
    *extent = e;
  });
}

rdl_status rdl_facet_bounding_mask(const uint8_t* vertical_mask, uint64_t vertical_mask_x,
                                   uint64_t vertical_mask_width, const uint8_t* horizontal_mask,
                                   uint8_t* mask, uint64_t width, uint64_t height, uint64_t* box) {
  if (!vertical_mask || !horizontal_mask || !mask || !box) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(nullptr, [&] {
    size_t x, y, w, h;
    BoundingMask(vertical_mask, vertical_mask_x, vertical_mask_width, horizontal_mask, mask, width,
                 height, x, y, w, h);
    box[0] = x;
    box[1] = y;
    box[2] = w;
    box[3] = h;
  });
}

rdl_status rdl_facet_grid(const float* image, uint64_t width, uint64_t height,
                          const uint8_t* user_mask, uint64_t grid_width, uint64_t grid_height,
                          rdl_facet_info* out) {
  if (!image || !out) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(nullptr, [&] {
    const std::vector<Facet> f = MakeFacets(image, width, height, user_mask, grid_width, grid_height);
    for (size_t i = 0; i != f.size(); ++i) {
      out[i] = rdl_facet_info{};
      out[i].x = f[i].x;
      out[i].y = f[i].y;
      out[i].width = f[i].width;
      out[i].height = f[i].height;
    }
  });
}

uint64_t rdl_parallel_facet_count(const rdl_handle* h) { return h ? ParallelFacetCount(*h) : 0; }

rdl_status rdl_parallel_timing(const rdl_handle* h, double* host_geometry_ms, double* total_ms) {
  if (!h || !host_geometry_ms || !total_ms) return RDL_ERR_INVALID_ARGUMENT;
  ParallelTiming(*h, *host_geometry_ms, *total_ms);
  return RDL_OK;
}

rdl_status rdl_parallel_facet_get(const rdl_handle* h, uint64_t i, rdl_facet_info* out,
                                  uint8_t* boundary_mask) {
  if (!h || !out) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(const_cast<rdl_handle*>(h), [&] {
    const FacetInfo f = ParallelFacet(*h, i);
    out->x = f.x;
    out->y = f.y;
    out->width = f.width;
    out->height = f.height;
    out->peak = f.peak;
    out->reached_major_threshold = f.reached_major_threshold;
    out->iteration_number = f.iteration_number;
    if (boundary_mask) {
      const std::vector<uint8_t>& m = ParallelFacetBoundaryMask(*h, i);
      std::copy(m.begin(), m.end(), boundary_mask);
    }
  });
}

rdl_status rdl_find_peak(rdl_handle* h, const rdl_cube_desc* c, const float* const* images,
                         int32_t integration_mode, int32_t allow_negative, uint64_t start_y,
                         uint64_t end_y, uint64_t horizontal_border, uint64_t vertical_border,
                         const uint8_t* device_mask, const float* device_rms_factor,
                         void* stream_v, int32_t* found, uint64_t* x, uint64_t* y, float* value) {
  if (!h) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    CheckCube(c);
    RDL_CUDA(cudaSetDevice(h->device));
    cudaStream_t stream = StreamOf(h, stream_v);
    DeviceCube cube;
    FillCubeMeta(cube, c);
    ConstPtrTable t{};
    for (size_t i = 0; i != cube.n_images; ++i) {
      t.p[i] = images[i];
      cube.residual.p[i] = const_cast<float*>(images[i]);
    }
    SetVec(cube);
    // the 128-bit path also reads the mask (uchar4) and the RMS image (float4)
    if ((device_mask && (reinterpret_cast<uintptr_t>(device_mask) & 3u)) ||
        (device_rms_factor && !Aligned16(device_rms_factor)))
      cube.vec = 1;
    const IntegSpec spec = integration_mode == 0 ? MakeLinearSpec(cube) : MakeSquareSpec(cube);
    const PeakWindow w = MakePeakWindow(cube.width, cube.height, start_y, end_y, horizontal_border,
                                        vertical_border, device_mask, device_rms_factor,
                                        allow_negative != 0);
    h->key.Reserve(1);
    LaunchFindPeak(spec, t, int(cube.width), int(cube.height), w, h->key.get(), cube.vec, stream);
    ++h->launches;
    unsigned long long key = 0;
    RDL_CUDA(cudaMemcpyAsync(&key, h->key.get(), sizeof(key), cudaMemcpyDeviceToHost, stream));
    RDL_CUDA(cudaStreamSynchronize(stream));
    float v = 0.0f;
    uint32_t index = 0;
    const bool has = DecodePeakKey(key, v, index);
    if (found) *found = has;
    if (x) *x = has ? index % cube.width : cube.width;
    if (y) *y = has ? index / cube.width : cube.height;
    if (value) *value = has ? v : 0.0f;
  });
}

rdl_status rdl_subtract_image(rdl_handle* h, const rdl_cube_desc* c, float* const* images,
                              const float* const* psfs, uint64_t x, uint64_t y,
                              const float* factors, void* stream_v) {
  if (!h) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    CheckCube(c);
    RDL_CUDA(cudaSetDevice(h->device));
    cudaStream_t stream = StreamOf(h, stream_v);
    const size_t n_images = c->n_channels * c->n_pol;
    PtrTable img{};
    ConstPtrTable psf{};
    for (size_t i = 0; i != n_images; ++i) {
      img.p[i] = images[i];
      psf.p[i] = psfs[i / c->n_pol];
    }
    LaunchSubtract(img, psf, int(n_images), int(c->width), int(c->height), int(x), int(y), factors,
                   stream);
    ++h->launches;
    RDL_CUDA(cudaGetLastError());
  });
}

rdl_status rdl_integrate(rdl_handle* h, const rdl_cube_desc* c, const float* const* images,
                         int32_t integration_mode, float* device_dest, void* stream_v) {
  if (!h) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    CheckCube(c);
    RDL_CUDA(cudaSetDevice(h->device));
    cudaStream_t stream = StreamOf(h, stream_v);
    DeviceCube cube;
    FillCubeMeta(cube, c);
    ConstPtrTable t{};
    for (size_t i = 0; i != cube.n_images; ++i) {
      t.p[i] = images[i];
      cube.residual.p[i] = const_cast<float*>(images[i]);
    }
    SetVec(cube);
    const int vec = cube.vec == 4 && Aligned16(device_dest) ? 4 : 1;
    const IntegSpec spec = integration_mode == 0 ? MakeLinearSpec(cube) : MakeSquareSpec(cube);
    LaunchIntegrate(spec, t, cube.Pixels(), device_dest, vec, stream);
    ++h->launches;
    RDL_CUDA(cudaGetLastError());
  });
}

rdl_status rdl_convolve(rdl_handle* h, float* const* images, uint64_t n_images,
                        const float* device_kernel, uint64_t width, uint64_t height,
                        void* stream_v) {
  if (!h) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    cudaStream_t stream = StreamOf(h, stream_v);
    FftConvolver& conv = GetConvolver(*h, width, height);
    conv.SetKernel(device_kernel, stream);
    for (uint64_t i = 0; i != n_images; ++i) conv.Convolve(images[i], images[i], stream);
    h->launches += conv.TakeLaunchCount();
    RDL_CUDA(cudaGetLastError());
  });
}

rdl_status rdl_padded_convolution(rdl_handle* h, const float* device_image, const float* device_psf,
                                  float* device_out, uint64_t width, uint64_t height,
                                  uint64_t padded_width, uint64_t padded_height, void* stream_v) {
  if (!h || !device_image || !device_psf || !device_out) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    if (padded_width < width || padded_height < height)
      throw std::invalid_argument("the padded size is smaller than the image");
    RDL_CUDA(cudaSetDevice(h->device));
    cudaStream_t stream = StreamOf(h, stream_v);
    FftConvolver& conv = GetConvolver(*h, padded_width, padded_height, width, height);
    h->fft_spectrum.Reserve(conv.SpectrumSize());
    conv.MakePsfKernelSpectrum(device_psf, width, height, h->fft_spectrum.get(), stream);
    conv.ConvolveWith(device_image, width, height, h->fft_spectrum.get(), device_out, width, height,
                      0, stream);
    h->launches += conv.TakeLaunchCount();
    RDL_CUDA(cudaGetLastError());
  });
}

rdl_status rdl_multiscale_transform(rdl_handle* h, float* const* images, uint64_t n_images,
                                    uint64_t width, uint64_t height, float scale, int32_t shape,
                                    void* stream_v) {
  if (!h) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    cudaStream_t stream = StreamOf(h, stream_v);
    MultiScaleTransform(*h, images, n_images, width, height, scale, shape, stream);
    RDL_CUDA(cudaGetLastError());
  });
}

// ---- multi-GPU: row-sharded GenericClean ----------------------------------------------------
rdl_status rdl_dist_unique_id(uint8_t* out128) {
  if (!out128) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(nullptr, [&] { NcclGetUniqueId(out128); });
}

rdl_status rdl_dist_init(rdl_handle* h, const uint8_t* id128, int32_t rank, int32_t world) {
  if (!h || !id128) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    NcclInit(h->dist, id128, rank, world);
  });
}

rdl_status rdl_dist_set_row_band(rdl_handle* h, uint64_t y_offset, uint64_t global_height) {
  if (!h) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    if (global_height == 0) {  // back to whole frames
      h->dist.band_set = false;
      return;
    }
    if (y_offset >= global_height) throw std::invalid_argument("row band starts outside the frame");
    h->dist.band_set = true;
    h->dist.y_offset = y_offset;
    h->dist.global_height = global_height;
    h->dist.block_rows = 0;
  });
}

rdl_status rdl_dist_set_row_blocks(rdl_handle* h, uint64_t block_rows, uint64_t global_height) {
  if (!h) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    if (block_rows == 0 || global_height == 0 ||
        global_height % (block_rows * uint64_t(h->dist.world)) != 0)
      throw std::invalid_argument("the frame height must be a multiple of block_rows * world");
    h->dist.band_set = true;
    h->dist.y_offset = 0;
    h->dist.global_height = global_height;
    h->dist.block_rows = block_rows;
  });
}

rdl_status rdl_dist_finalize(rdl_handle* h) {
  if (!h) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    NcclDestroy(h->dist);
    h->dist = DistContext{};
  });
}

rdl_status rdl_iuwt_smooth(rdl_handle* h, const float* device_in, float* device_out,
                           uint64_t width, uint64_t height, int32_t scale, int32_t flavour,
                           void* stream_v) {
  if (!h || !device_in || !device_out) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    IuwtSmooth(*h, device_in, device_out, width, height, scale, flavour != 0, StreamOf(h, stream_v));
  });
}

rdl_status rdl_iuwt_decompose(rdl_handle* h, const float* device_input, uint64_t width,
                              uint64_t height, int32_t n_scales, int32_t include_largest,
                              float* device_coefficients, void* stream_v) {
  if (!h || !device_input || !device_coefficients) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    IuwtDecompose(*h, device_input, width, height, n_scales, include_largest != 0,
                  device_coefficients, StreamOf(h, stream_v));
  });
}

rdl_status rdl_iuwt_decompose_aliased(rdl_handle* h, float* device_x, uint64_t width,
                                      uint64_t height, int32_t n_scales,
                                      float* device_coefficients, void* stream_v) {
  if (!h || !device_x || !device_coefficients) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    IuwtDecomposeAliased(*h, device_x, width, height, n_scales, device_coefficients,
                         StreamOf(h, stream_v));
  });
}

rdl_status rdl_rms_sliding_minimum(rdl_handle* h, const float* device_in, float* device_out,
                                   uint64_t width, uint64_t height, uint64_t window_size,
                                   int32_t maximum, void* stream_v) {
  if (!h || !device_in || !device_out) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    RmsSlidingMinimum(*h, device_in, device_out, width, height, window_size, maximum != 0,
                      StreamOf(h, stream_v));
  });
}

rdl_status rdl_rms_make_factor_image(rdl_handle* h, float* device_image, uint64_t n_pixels,
                                     double local_rms_strength, double* stddev, void* stream_v) {
  if (!h || !device_image) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    const double s = RmsMakeFactorImage(*h, device_image, n_pixels, local_rms_strength,
                                        StreamOf(h, stream_v));
    if (stddev) *stddev = s;
  });
}

rdl_status rdl_rms_make(rdl_handle* h, const float* device_in, float* device_out, uint64_t width,
                        uint64_t height, double window_size, double beam_major, double beam_minor,
                        double beam_pa, double pixel_scale_l, double pixel_scale_m,
                        int32_t with_negativity_limit, void* stream_v) {
  if (!h || !device_in || !device_out) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    RmsMake(*h, device_in, device_out, width, height, window_size, beam_major, beam_minor, beam_pa,
            pixel_scale_l, pixel_scale_m, with_negativity_limit != 0, StreamOf(h, stream_v));
  });
}

rdl_status rdl_iuwt_recompose(rdl_handle* h, const float* device_coefficients, uint64_t width,
                              uint64_t height, int32_t n_scales, int32_t include_largest,
                              const uint8_t* device_masks, float* device_out, void* stream_v) {
  if (!h || !device_coefficients || !device_out) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    IuwtRecompose(*h, device_coefficients, width, height, n_scales, include_largest != 0,
                  device_masks, device_out, StreamOf(h, stream_v));
  });
}

rdl_status rdl_iuwt_select_structures(rdl_handle* h, const float* device_coefficients,
                                      uint8_t* device_masks, uint64_t width, uint64_t height,
                                      int32_t n_scales, const float* thresholds, int32_t min_scale,
                                      int32_t end_scale, float clean_border,
                                      const uint8_t* device_prior_mask, uint64_t* area_size,
                                      void* stream_v) {
  if (!h || !device_coefficients || !device_masks || !thresholds || !area_size)
    return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    *area_size = IuwtSelectStructures(*h, device_coefficients, device_masks, width, height,
                                      n_scales, thresholds, min_scale, end_scale, clean_border,
                                      device_prior_mask, StreamOf(h, stream_v));
  });
}

rdl_status rdl_iuwt_floodfill(rdl_handle* h, const float* device_coefficients,
                              uint8_t* device_masks, uint64_t width, uint64_t height,
                              int32_t n_scales, const float* thresholds, int32_t min_scale,
                              int32_t end_scale, uint64_t seed_x, uint64_t seed_y,
                              int32_t seed_scale, float clean_border,
                              const uint8_t* device_prior_mask, uint64_t* area_size,
                              void* stream_v) {
  if (!h || !device_coefficients || !device_masks || !thresholds || !area_size)
    return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    *area_size = IuwtFloodfill(*h, device_coefficients, device_masks, width, height, n_scales,
                               thresholds, min_scale, end_scale, seed_x, seed_y, seed_scale,
                               clean_border, device_prior_mask, StreamOf(h, stream_v));
  });
}

rdl_status rdl_iuwt_floodfill_2d(rdl_handle* h, const float* device_image, uint8_t* device_mask,
                                 uint64_t width, uint64_t height, float threshold, int32_t use_abs,
                                 uint64_t seed_x, uint64_t seed_y, uint64_t* area_size,
                                 void* stream_v) {
  if (!h || !device_image || !device_mask || !area_size) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    *area_size = IuwtFloodfill2D(*h, device_image, device_mask, width, height, threshold,
                                 use_abs != 0, seed_x, seed_y, StreamOf(h, stream_v));
  });
}

rdl_status rdl_median_and_stddev_from_mad(rdl_handle* h, const float* device_image,
                                         uint64_t n_pixels, float* median, float* stddev,
                                         uint64_t* n_finite, void* stream_v) {
  if (!h || (!device_image && n_pixels)) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    const MadStatistics s =
        ImageMedianAndStdDevFromMad(*h, device_image, n_pixels, StreamOf(h, stream_v));
    if (median) *median = s.median;
    if (stddev) *stddev = s.stddev;
    if (n_finite) *n_finite = s.n_finite;
  });
}

rdl_status rdl_load_and_average(rdl_handle* h, const float* const* device_inputs,
                                const double* weights, const uint32_t* output_index,
                                uint64_t n_inputs, float* const* device_outputs,
                                uint64_t n_outputs, uint64_t n_pixels, void* stream_v) {
  if (!h || !device_outputs || (n_inputs && (!device_inputs || !weights || !output_index)))
    return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    ImageLoadAndAverage(*h, device_inputs, weights, output_index, n_inputs, device_outputs,
                        n_outputs, n_pixels, false, StreamOf(h, stream_v));
  });
}

rdl_status rdl_load_and_average_psfs(rdl_handle* h, const float* const* device_inputs,
                                     const double* weights, const uint32_t* output_index,
                                     uint64_t n_inputs, float* const* device_outputs,
                                     uint64_t n_outputs, uint64_t n_pixels, void* stream_v) {
  if (!h || !device_outputs || (n_inputs && (!device_inputs || !weights || !output_index)))
    return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    ImageLoadAndAverage(*h, device_inputs, weights, output_index, n_inputs, device_outputs,
                        n_outputs, n_pixels, true, StreamOf(h, stream_v));
  });
}

rdl_status rdl_auto_mask_from_model(rdl_handle* h, const float* const* device_models,
                                    uint64_t n_models, uint8_t* device_mask, uint64_t n_pixels,
                                    void* stream_v) {
  if (!h || !device_mask || (n_models && !device_models)) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    ImageMaskFromModel(*h, device_models, n_models, device_mask, n_pixels, StreamOf(h, stream_v));
  });
}

rdl_status rdl_interpolate_model(rdl_handle* h, const float* const* device_models,
                                 uint64_t n_models, const double* output_frequencies,
                                 uint64_t n_outputs, float* const* device_outputs,
                                 uint64_t n_pixels, void* stream_v) {
  if (!h || !device_models || (n_outputs && (!output_frequencies || !device_outputs)))
    return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    ImageInterpolateModel(*h, device_models, n_models, output_frequencies, n_outputs,
                          device_outputs, n_pixels, StreamOf(h, stream_v));
  });
}

rdl_status rdl_iuwt_mad(rdl_handle* h, const float* device_image, uint64_t n, float* mad,
                        void* stream_v) {
  if (!h || !device_image || !mad) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    *mad = IuwtMad(*h, device_image, n, StreamOf(h, stream_v));
  });
}

rdl_status rdl_iuwt_max_abs(rdl_handle* h, const float* device_image, uint64_t width,
                            uint64_t height, float clean_border, int32_t allow_negative,
                            const uint8_t* device_mask, float* value, uint64_t* x, uint64_t* y,
                            void* stream_v) {
  if (!h || !device_image || !value || !x || !y) return RDL_ERR_INVALID_ARGUMENT;
  return Guard(h, [&] {
    RDL_CUDA(cudaSetDevice(h->device));
    size_t px = 0, py = 0;
    *value = IuwtMaxAbs(*h, device_image, width, height, clean_border, allow_negative != 0,
                        device_mask, px, py, StreamOf(h, stream_v));
    *x = px;
    *y = py;
  });
}

uint64_t rdl_good_fft_size(uint64_t minimum_size) { return CalculateGoodFFTSize(minimum_size); }
uint64_t rdl_convolution_size(double convolution_scale, uint64_t original_size, double padding) {
  return GetConvolutionSize(convolution_scale, original_size, padding);
}

const char* rdl_version(void) { return "radler_b200 0.1 (sm_100a)"; }

int32_t rdl_device_count(void) {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}

}  // extern "C"
