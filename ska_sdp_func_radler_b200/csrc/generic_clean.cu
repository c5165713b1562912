// GenericClean::ExecuteMajorIteration on the device (generic_clean.cc:56-253).
// Path B (use_sub_minor_optimization = false): the FindPeak -> fit -> SubtractImage ->
// re-integrate loop runs as a chain of fused HogbomStepKernel launches; the host only
// polls the device-side control block once per batch of launches.
// Path A (sub-minor optimisation) lives in subminor.cu.
#include <algorithm>
#include <cmath>
#include <cstdlib>

#include "handle.h"
#include "subminor.cuh"

namespace rdl {

namespace {

struct PinnedState {
  CleanState* p = nullptr;
  PinnedState() { RDL_CUDA(cudaMallocHost(&p, sizeof(CleanState))); }
  ~PinnedState() { cudaFreeHost(p); }
};

struct Events {
  cudaEvent_t e[2] = {nullptr, nullptr};
  Events() {
    for (auto& v : e) RDL_CUDA(cudaEventCreateWithFlags(&v, cudaEventDisableTiming));
  }
  ~Events() {
    for (auto& v : e)
      if (v) cudaEventDestroy(v);
  }
};

rdl_result FinishResult(const rdl_handle& h, const CleanState& s, uint64_t iteration_at_start) {
  // generic_clean.cc:75-88, 208-252
  rdl_result r{};
  r.has_starting_peak = s.has_start;
  r.starting_peak_value = s.has_start ? s.start_value : 0.0f;
  r.final_peak_value = r.starting_peak_value;
  r.iteration_number = s.iteration;
  if (s.done == 2 || s.done == 3) return r;  // no peak / no iterations left
  if (s.diverging) {
    if (s.has_peak) r.final_peak_value = s.max_value;
    r.another_iteration_required = 0;
    r.is_diverging = 1;
  } else if (s.has_peak) {
    const float v = s.max_value;
    const bool final_threshold_reached = std::fabs(v) <= h.settings.threshold || v == 0.0f;
    const bool negative_reached = v < 0.0f && h.settings.stop_on_negative_component;
    const bool mgain_reached = std::fabs(v) <= s.major_iter_threshold;
    const bool did_work = (s.iteration - iteration_at_start) != 0;
    r.another_iteration_required =
        mgain_reached && did_work && !negative_reached && !final_threshold_reached;
    r.final_peak_value = v;
  } else {
    r.another_iteration_required = 0;
  }
  return r;
}

}  // namespace

rdl_result RunGenericClean(rdl_handle& h, DeviceCube& cube, cudaStream_t stream) {
  if (h.settings.stop_on_negative_component) h.settings.allow_negative_components = true;
  const size_t width = cube.width, height = cube.height;
  const uint64_t iteration_at_start = h.iteration_number;

  if (h.dist.band_set && h.config.use_sub_minor_optimization)
    throw std::domain_error("row-sharded cubes are supported by GenericClean without the sub-minor "
                            "optimisation only");
  if (h.config.use_sub_minor_optimization) return RunGenericCleanSubMinor(h, cube, stream);

  HogbomParams p{};
  p.img = cube.residual;
  p.model = cube.model;
  for (size_t i = 0; i != cube.n_images; ++i) p.psf.p[i] = cube.psf.p[cube.PsfIndex(i)];
  p.first_spec = MakeLinearSpec(cube);
  p.step_spec = MakeSquareSpec(cube);
  p.simple_first = IsSimpleSpec(p.first_spec, cube.n_images);
  p.simple_step = IsSimpleSpec(p.step_spec, cube.n_images);
  // peak_finder.h:99-107: borders = round(size * ratio)
  const size_t hb = size_t(std::round(width * h.settings.clean_border_ratio));
  const size_t vb = size_t(std::round(height * h.settings.clean_border_ratio));
  p.window = MakePeakWindow(width, height, 0, height, hb, vb, h.mask.get(), h.rms.get(),
                            h.settings.allow_negative_components);
  p.n_images = int(cube.n_images);
  p.width = int(width);
  p.height = int(height);
  p.threshold = h.settings.threshold;
  p.major_iteration_threshold = h.settings.major_iteration_threshold;
  p.gain = h.settings.minor_loop_gain;
  p.mgain = h.settings.major_loop_gain;
  p.divergence_limit = h.settings.divergence_limit;
  p.stop_on_negative = h.settings.stop_on_negative_component;
  p.max_iterations = h.settings.max_iterations;
  p.projection = h.fit_mode == RDL_FIT_POLYNOMIAL ? h.projection.get() : nullptr;
  p.n_fit = int(h.fit_freqs.size());
  p.n_pol = int(cube.n_pol);
  h.trace_dev.Reserve(std::max<uint64_t>(h.trace_capacity, 1));
  p.trace = h.trace_dev.get();
  p.trace_capacity = h.trace_capacity;
  p.y_offset = 0;
  p.global_height = int(height);
  p.block_rows = int(height);
  p.block_world = 1;
  p.block_rank = 0;
  const bool sharded = h.dist.band_set;
  const bool logging = h.model_log && !sharded;
  h.model_log_count = 0;
  if (logging) {
    const uint64_t capacity = h.settings.max_iterations > h.iteration_number
                                  ? h.settings.max_iterations - h.iteration_number
                                  : 0;
    h.log_index.Reserve(std::max<uint64_t>(capacity, 1));
    h.log_values.Reserve(std::max<uint64_t>(capacity, 1) * cube.n_images);
    p.log_index = h.log_index.get();
    p.log_values = h.log_values.get();
    p.log_first_iteration = h.iteration_number;
  }
  // the component log of the call, for the host to apply to its model planes
  auto collect_log = [&](const CleanState& s) {
    if (!logging) return;
    const uint64_t count = s.iteration - iteration_at_start;
    h.model_log_count = count;
    h.log_index_host.resize(count);
    h.log_values_host.resize(count * cube.n_images);
    if (count == 0) return;
    RDL_CUDA(cudaMemcpyAsync(h.log_index_host.data(), h.log_index.get(), count * sizeof(uint32_t),
                             cudaMemcpyDeviceToHost, stream));
    RDL_CUDA(cudaMemcpyAsync(h.log_values_host.data(), h.log_values.get(),
                             count * cube.n_images * sizeof(float), cudaMemcpyDeviceToHost, stream));
    RDL_CUDA(cudaStreamSynchronize(stream));
  };
  if (sharded) {
    // Row band of a width x global_height frame (rdl_dist_set_row_band): borders, window and
    // peak indices are the frame's; mask / RMS image are full-frame, offset to this band.
    const size_t gh = h.dist.global_height;
    p.y_offset = int(h.dist.y_offset);
    p.global_height = int(gh);
    const size_t gvb = size_t(std::round(gh * h.settings.clean_border_ratio));
    if (h.dist.block_rows) {
      p.y_offset = 0;
      p.block_rows = int(h.dist.block_rows);
      p.block_world = h.dist.world;
      p.block_rank = h.dist.rank;
    }
    p.window = MakePeakWindow(width, gh, 0, gh, hb, gvb, h.mask.get(), h.rms.get(),
                              h.settings.allow_negative_components);
    p.dist = 1;
    h.cand_local.Reserve(1);
    h.cand_all.Reserve(size_t(h.dist.world));
    p.cand_local = h.cand_local.get();
  }
  auto step = [&](bool first) {
    LaunchHogbomStep(p, h.state.get(), first, cube.vec, stream);
    ++h.launches;
    if (sharded) {
      if (h.dist.world > 1)
        NcclAllGatherBytes(h.dist, h.cand_local.get(), h.cand_all.get(), sizeof(DistCandidate),
                           stream);
      LaunchHogbomDistControl(p, h.state.get(),
                              h.dist.world > 1 ? h.cand_all.get() : h.cand_local.get(),
                              h.dist.world, first, stream);
      ++h.launches;
    }
  };

  h.state.Reserve(1);
  PinnedState host;
  *host.p = CleanState{};
  host.p->iteration = h.iteration_number;
  RDL_CUDA(cudaMemcpyAsync(h.state.get(), host.p, sizeof(CleanState), cudaMemcpyHostToDevice,
                           stream));

  // Cubes that stay in L2 together with their PSFs (K1, facets): the whole loop in one
  // cooperative launch, no launch + last-CTA control per iteration (clean_kernels.cu).
  // RDL_HOGBOM_PERSISTENT_MB moves the limit (0 switches the path off).
  const char* env_mb = getenv("RDL_HOGBOM_PERSISTENT_MB");
  const double limit_mb = env_mb ? atof(env_mb) : 40.0;
  const double cube_mb = double(cube.n_images) * double(cube.Pixels()) * sizeof(float) / 1048576.0;
  if (!sharded && cube_mb <= limit_mb) {
    h.persist_scratch.Reserve(HogbomPersistentScratchBytes());
    if (LaunchHogbomPersistent(p, h.state.get(), h.persist_scratch.get(), cube.vec, stream)) {
      ++h.launches;
      RDL_CUDA(cudaMemcpyAsync(host.p, h.state.get(), sizeof(CleanState), cudaMemcpyDeviceToHost,
                               stream));
      RDL_CUDA(cudaStreamSynchronize(stream));
      const CleanState s = *host.p;
      h.iteration_number = s.iteration;
      h.trace_host.resize(s.n_trace);
      if (s.n_trace)
        RDL_CUDA(cudaMemcpyAsync(h.trace_host.data(), h.trace_dev.get(),
                                 s.n_trace * sizeof(TraceEntry), cudaMemcpyDeviceToHost, stream));
      RDL_CUDA(cudaStreamSynchronize(stream));
      collect_log(s);
      return FinishResult(h, s, iteration_at_start);
    }
  }
  step(true);

  // The host runs AHEAD of the device: a launch whose loop has ended returns at once (st->done),
  // so batches are queued before it is known that they are needed -- two in flight, doubling from
  // about a millisecond of work up to about 50 ms (64 launches when every step carries a collective) --
  // and the control block of batch k is looked at while batch k + 1 runs. A host thread that is
  // descheduled for tens of milliseconds (measured on shared hosts: whole steps 2-3x slower with
  // a synchronisation every 8 launches) no longer idles the GPU; the price is at most one batch
  // of empty launches after the loop ends. Batch sizes are a fixed schedule, identical on every
  // rank of a sharded run (the ranks must issue the same number of exchanges).
  const double pixels = sharded ? double(width) * double(h.dist.global_height) / h.dist.world
                                : double(cube.Pixels());
  const double bytes_per_iteration = 12.0 * double(cube.n_images) * pixels;
  const double t_iteration = std::max(bytes_per_iteration / 6.0e12, 3.0e-6);
  uint64_t batch = uint64_t(std::clamp(1.0e-3 / t_iteration, 8.0, 256.0));
  // about 50 ms of queued work per batch (a queue of a thousand short launches measured slower
  // per launch than one of a few hundred)
  const uint64_t batch_cap =
      sharded ? 64 : uint64_t(std::clamp(50.0e-3 / t_iteration, double(batch), 256.0));
  // no more launches than iterations that can still happen
  uint64_t budget = h.settings.max_iterations > h.iteration_number
                        ? h.settings.max_iterations - h.iteration_number
                        : 0;
  PinnedState slots[2];
  Events events;
  auto issue = [&](int slot, uint64_t count) {
    for (uint64_t b = 0; b != count; ++b) step(false);
    RDL_CUDA(cudaGetLastError());
    RDL_CUDA(cudaMemcpyAsync(slots[slot].p, h.state.get(), sizeof(CleanState),
                             cudaMemcpyDeviceToHost, stream));
    RDL_CUDA(cudaEventRecord(events.e[slot], stream));
  };
  auto take = [&](uint64_t want) {
    const uint64_t n = std::min(want, budget);
    budget -= n;
    return n;
  };
  int cur = 0;
  issue(cur, take(batch));
  for (;;) {
    batch = std::min(batch * 2, batch_cap);
    issue(cur ^ 1, take(batch));
    RDL_CUDA(cudaEventSynchronize(events.e[cur]));
    if (slots[cur].p->done) break;
    cur ^= 1;
    if (budget == 0) {  // everything that can happen is queued: the last copy tells
      RDL_CUDA(cudaEventSynchronize(events.e[cur]));
      break;
    }
  }
  RDL_CUDA(cudaStreamSynchronize(stream));  // drains what was queued ahead
  *host.p = *slots[cur].p;
  const CleanState s = *host.p;
  h.iteration_number = s.iteration;
  h.trace_host.resize(s.n_trace);
  if (s.n_trace)
    RDL_CUDA(cudaMemcpyAsync(h.trace_host.data(), h.trace_dev.get(),
                             s.n_trace * sizeof(TraceEntry), cudaMemcpyDeviceToHost, stream));
  RDL_CUDA(cudaStreamSynchronize(stream));
  collect_log(s);
  return FinishResult(h, s, iteration_at_start);
}

}  // namespace rdl
