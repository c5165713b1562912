// SubMinorLoop on the device (algorithms/subminor_loop.{h,cc}) and GenericClean path A
// (generic_clean.cc:115-162).
//
//   SelectCountKernel / ScanKernel / GatherKernel
//        findPeakPositions + MakeSets (subminor_loop.cc:119-184): threshold selection that
//        keeps the reference's row-major order (ballot words + block offsets), then the
//        gather of every image into compact n_sel x 1 sets.
//   SubMinorLoopKernel
//        the component loop (subminor_loop.cc:62-115) as ONE persistent cooperative kernel:
//        the compact residuals, the model and the component list stay on the device; per
//        iteration every CTA subtracts the gathered PSF values of the previous component,
//        re-integrates its pixels over the channels, and the arg-max is joined through one
//        64-bit atomicMax and a single grid barrier. The winning pixel's per-image values
//        travel with the reduction (per-CTA candidate slots), so no second barrier is needed.
#include <algorithm>
#include <cmath>
#include <cstring>

#include "fft.cuh"
#include "subminor.cuh"

#include <vector>

namespace rdl {

namespace {

constexpr int kSelThreads = 256;
constexpr int kSelIters = 8;
constexpr int kSelChunk = kSelThreads * kSelIters;
constexpr int kLoopThreads = 256;
constexpr int kCandStride = kMaxImages + 1;

struct SelectParams {
  IntegSpec spec;
  ConstPtrTable img;
  const float* rms;
  const uint8_t* mask;
  int width, height, x0, x1, y0, y1;
  float threshold;
  int allow_negative;
};

__global__ void __launch_bounds__(kSelThreads)
    SelectCountKernel(const __grid_constant__ SelectParams p, uint32_t* flags,
                      uint32_t* block_counts) {
  __shared__ uint32_t s_count[kSelThreads / 32];
  const size_t n = size_t(p.width) * p.height;
  const size_t base = size_t(blockIdx.x) * kSelChunk;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t count = 0;
  for (int it = 0; it < kSelIters; ++it) {
    const size_t i = base + size_t(it) * kSelThreads + threadIdx.x;
    bool sel = false;
    if (i < n) {
      const int y = int(i / p.width), x = int(i - size_t(y) * p.width);
      if (x >= p.x0 && x < p.x1 && y >= p.y0 && y < p.y1) {
        float v = IntegratePixel(p.spec, p.img, i);
        if (p.rms) v = __fmul_rn(v, p.rms[i]);
        const float cmp = p.allow_negative ? fabsf(v) : v;
        sel = cmp >= p.threshold && (!p.mask || p.mask[i]);
      }
    }
    const uint32_t word = __ballot_sync(0xffffffffu, sel);
    if (lane == 0) {
      flags[(base + size_t(it) * kSelThreads) / 32 + warp] = word;
      count += __popc(word);
    }
  }
  if (lane == 0) s_count[warp] = count;
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t total = 0;
    for (int w = 0; w < kSelThreads / 32; ++w) total += s_count[w];
    block_counts[blockIdx.x] = total;
  }
}

__global__ void __launch_bounds__(1024)
    ScanKernel(const uint32_t* counts, uint32_t* offsets, int n, uint32_t* total) {
  __shared__ uint32_t s_warp[32];
  __shared__ uint32_t s_running;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) s_running = 0;
  __syncthreads();
  for (int base = 0; base < n; base += 1024) {
    const int i = base + threadIdx.x;
    const uint32_t v = i < n ? counts[i] : 0u;
    uint32_t incl = v;
    for (int o = 1; o < 32; o <<= 1) {
      const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    if (warp == 0) {
      uint32_t w = s_warp[lane];
      for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, w, o);
        if (lane >= o) w += t;
      }
      s_warp[lane] = w;
    }
    __syncthreads();
    const uint32_t before = s_running + (warp ? s_warp[warp - 1] : 0u);
    if (i < n) offsets[i] = before + incl - v;
    __syncthreads();
    if (threadIdx.x == 1023) s_running = before + incl;
    __syncthreads();
  }
  if (threadIdx.x == 0) *total = s_running;
}

struct GatherParams {
  ConstPtrTable img;
  PtrTable resid;
  const float* rms;
  float* rms_out;
  uint32_t* pos;
  int n_images, width;
  size_t n;
};

__global__ void __launch_bounds__(kSelThreads)
    GatherKernel(const __grid_constant__ GatherParams p, const uint32_t* flags,
                 const uint32_t* block_offsets) {
  __shared__ uint32_t s_off[kSelIters * (kSelThreads / 32)];
  const size_t base = size_t(blockIdx.x) * kSelChunk;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  constexpr int kWarps = kSelThreads / 32;
  if (threadIdx.x < kSelIters * kWarps) {
    const int it = threadIdx.x / kWarps, w = threadIdx.x - it * kWarps;
    s_off[threadIdx.x] = __popc(flags[(base + size_t(it) * kSelThreads) / 32 + w]);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t run = block_offsets[blockIdx.x];
    for (int k = 0; k < kSelIters * kWarps; ++k) {
      const uint32_t c = s_off[k];
      s_off[k] = run;
      run += c;
    }
  }
  __syncthreads();
  for (int it = 0; it < kSelIters; ++it) {
    const uint32_t word = flags[(base + size_t(it) * kSelThreads) / 32 + warp];
    if (!((word >> lane) & 1u)) continue;
    const size_t i = base + size_t(it) * kSelThreads + threadIdx.x;
    const uint32_t dst = s_off[it * kWarps + warp] + __popc(word & ((1u << lane) - 1u));
    const uint32_t y = uint32_t(i / p.width), x = uint32_t(i - size_t(y) * p.width);
    p.pos[dst] = x | (y << 16);
    for (int k = 0; k < p.n_images; ++k) p.resid.p[k][dst] = p.img.p[k][i];
    if (p.rms) p.rms_out[dst] = p.rms[i];
  }
}

struct LoopState {
  unsigned long long key[3];
  unsigned int bar_count, bar_gen;
  unsigned long long iteration, n_trace;
  float max_value, flux_cleaned;
  int diverging, pad;
};

struct LoopParams {
  PtrTable resid, model;  // compact sets, n_sel each
  ConstPtrTable psf;      // full-size PSF of image i
  const uint32_t* pos;    // x | y << 16
  const float* rms;       // compact RMS factors or null
  IntegSpec spec;
  int n_images, n_sel, width, height;
  float threshold, gain, divergence_limit;
  int allow_negative, stop_on_negative;
  unsigned long long max_iterations;
  const double* projection;
  int n_fit, n_pol;
  TraceEntry* trace;
  unsigned long long trace_capacity;
  uint32_t trace_scale;
  LoopState* st;
  float* cand;  // [grid][2][kCandStride]
};

__device__ __forceinline__ uint32_t Orderable(float v) {
  const uint32_t u = __float_as_uint(v);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

// GetMaxComponent (subminor_loop.cc:13-36): arg-max of |v| (or v), first index wins.
__device__ __forceinline__ unsigned long long ComponentKey(float v, bool allow_negative,
                                                           uint32_t index) {
  if (v != v) return 0ull;
  const float cmp = allow_negative ? fabsf(v) : v;
  return ((unsigned long long)Orderable(cmp) << 32) | (unsigned long long)(~index);
}

__device__ __forceinline__ void GridBarrier(LoopState* st) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    volatile unsigned int* gen = &st->bar_gen;
    const unsigned int g = *gen;
    if (atomicAdd(&st->bar_count, 1u) == gridDim.x - 1) {
      st->bar_count = 0u;
      __threadfence();
      atomicAdd(&st->bar_gen, 1u);
    } else {
      while (*gen == g) {
      }
    }
    __threadfence();
  }
  __syncthreads();
}

__global__ void __launch_bounds__(kLoopThreads)
    SubMinorLoopKernel(const __grid_constant__ LoopParams p) {
  __shared__ unsigned long long red[32];
  __shared__ float s_vals[kMaxImages];
  __shared__ float s_max;
  __shared__ unsigned long long s_key;
  const int total_threads = gridDim.x * blockDim.x;
  const int gtid = blockIdx.x * blockDim.x + threadIdx.x;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const ConstPtrTable& cres = reinterpret_cast<const ConstPtrTable&>(p.resid);
  LoopState* st = p.st;
  unsigned long long iteration = st->iteration, n_trace = st->n_trace;
  float flux = 0.0f, start = 0.0f, max_value = 0.0f;
  bool first = true, diverging = false;
  int slot = 0, parity = 0;
  const bool allow_negative = p.allow_negative != 0;

  // component of the previous iteration (none before the first arg-max)
  bool have_component = false;
  int comp_x = 0, comp_y = 0;

  for (;;) {
    // ---- subtract the previous component, re-integrate, local arg-max ----------------
    unsigned long long best = 0ull;
    for (int px = gtid; px < p.n_sel; px += total_threads) {
      if (have_component) {
        const uint32_t q = __ldg(p.pos + px);
        const int psf_x = int(q & 0xffffu) - comp_x + p.width / 2;
        const int psf_y = int(q >> 16) - comp_y + p.height / 2;
        if (psf_x >= 0 && psf_x < p.width && psf_y >= 0 && psf_y < p.height) {
          const size_t off = size_t(psf_y) * p.width + psf_x;
          for (int i = 0; i < p.n_images; ++i) {
            float* r = p.resid.p[i] + px;
            *r = __fmaf_rn(-__ldg(p.psf.p[i] + off), s_vals[i], *r);
          }
        }
      }
      float v = IntegratePixel(p.spec, cres, size_t(px));
      if (p.rms) v = __fmul_rn(v, __ldg(p.rms + px));
      const unsigned long long k = ComponentKey(v, allow_negative, uint32_t(px));
      best = k > best ? k : best;
    }
    best = BlockMaxU64(best, red);  // contains a __syncthreads: the block's stores are visible
    if (threadIdx.x == 0) s_key = best;
    __syncthreads();
    best = s_key;
    if (best && warp == 0) {
      // publish this block's candidate: per-image residuals + integrated value of its best pixel
      const uint32_t px = ~uint32_t(best);
      float* cand = p.cand + (size_t(blockIdx.x) * 2 + parity) * kCandStride;
      for (int i = lane; i < p.n_images; i += 32) __stcg(cand + i, p.resid.p[i][px]);
      if (lane == 0) {
        float v = IntegratePixel(p.spec, cres, size_t(px));
        if (p.rms) v = __fmul_rn(v, __ldg(p.rms + px));
        __stcg(cand + kMaxImages, v);
        atomicMax(&st->key[slot], best);
      }
    }
    GridBarrier(st);

    // ---- everyone reads the winner ------------------------------------------------------
    const unsigned long long key = __ldcg(&st->key[slot]);
    const uint32_t comp = key ? ~uint32_t(key) : 0u;
    const int owner = int((comp % uint32_t(total_threads)) / blockDim.x);
    const float* cand = p.cand + (size_t(owner) * 2 + parity) * kCandStride;
    __syncthreads();  // s_vals of the previous iteration are no longer needed
    if (threadIdx.x < p.n_images) s_vals[threadIdx.x] = __ldcg(cand + threadIdx.x);
    if (threadIdx.x == 0) s_max = key ? __ldcg(cand + kMaxImages) : __int_as_float(0x7fc00000);
    __syncthreads();
    max_value = s_max;
    if (first) {
      start = fabsf(max_value);
      first = false;
    } else {
      if (p.divergence_limit != 0.0f)
        diverging = fabsf(max_value) > __fmul_rn(start, p.divergence_limit);
      ++iteration;
    }
    const bool cont = fabsf(max_value) > p.threshold && iteration < p.max_iterations &&
                      (!p.stop_on_negative || max_value >= 0.0f) && !diverging;
    if (!cont) break;

    // component values: residual * gain, then the spectral fit (subminor_loop.cc:64-88)
    if (warp == 0) {
      for (int i = lane; i < p.n_images; i += 32) s_vals[i] = __fmul_rn(s_vals[i], p.gain);
      __syncwarp();
      if (p.projection) {
        float fitted[(kMaxImages + 31) / 32];
        int k = 0;
        for (int i = lane; i < p.n_images; i += 32, ++k) {
          const int ch = i / p.n_pol, pol = i - ch * p.n_pol;
          double s = 0.0;
          for (int j = 0; j < p.n_fit; ++j)
            s = fma(p.projection[ch * p.n_fit + j], double(s_vals[j * p.n_pol + pol]), s);
          fitted[k] = float(s);
        }
        __syncwarp();
        k = 0;
        for (int i = lane; i < p.n_images; i += 32, ++k) s_vals[i] = fitted[k];
        __syncwarp();
      }
    }
    __syncthreads();
    const uint32_t q = __ldg(p.pos + comp);
    comp_x = int(q & 0xffffu);
    comp_y = int(q >> 16);
    have_component = true;
    if (blockIdx.x == 0) {
      if (threadIdx.x < p.n_images) {
        float* m = p.model.p[threadIdx.x] + comp;
        *m = __fadd_rn(*m, s_vals[threadIdx.x]);
      }
      if (threadIdx.x == 0) {
        flux = __fadd_rn(flux, __fmul_rn(max_value, p.gain));
        if (n_trace < p.trace_capacity)
          p.trace[n_trace] = TraceEntry{uint32_t(iteration), p.trace_scale, uint32_t(comp_x),
                                        uint32_t(comp_y), max_value};
        ++n_trace;
        st->key[(slot + 2) % 3] = 0ull;  // the slot after next; a barrier separates its use
      }
    }
    slot = (slot + 1) % 3;
    parity ^= 1;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    st->iteration = iteration;
    st->n_trace = n_trace;
    st->max_value = max_value;
    st->flux_cleaned = flux;
    st->diverging = diverging ? 1 : 0;
  }
}

__global__ void ScatterKernel(const uint32_t* pos, const float* compact, float* full, int n_sel,
                              int width, int add) {
  for (int px = blockIdx.x * blockDim.x + threadIdx.x; px < n_sel; px += gridDim.x * blockDim.x) {
    const uint32_t q = pos[px];
    const size_t i = size_t(q >> 16) * width + (q & 0xffffu);
    full[i] = add ? __fadd_rn(full[i], compact[px]) : compact[px];
  }
}

__global__ void AutoMaskKernel(const uint32_t* pos, PtrTable model, int n_images, int n_sel,
                               int width, uint8_t* mask) {
  for (int px = blockIdx.x * blockDim.x + threadIdx.x; px < n_sel; px += gridDim.x * blockDim.x) {
    bool any = false;
    for (int i = 0; i < n_images; ++i) any = any || model.p[i][px] != 0.0f;
    if (any) {
      const uint32_t q = pos[px];
      mask[size_t(q >> 16) * width + (q & 0xffffu)] = 1;
    }
  }
}

int MaxCooperativeBlocks() {
  static int cached = 0;
  if (!cached) {
    int per_sm = 0, sms = 0, dev = 0;
    RDL_CUDA(cudaGetDevice(&dev));
    RDL_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    RDL_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, SubMinorLoopKernel,
                                                           kLoopThreads, 0));
    cached = std::max(1, std::min(per_sm, 2) * sms);
  }
  return cached;
}

int ScatterGrid(size_t n) { return int(std::clamp<size_t>((n + 255) / 256, 1, size_t(kSmCount) * 8)); }

}  // namespace

SubMinorOutcome RunSubMinorLoop(rdl_handle& h, const IntegSpec& linear_spec,
                                const ConstPtrTable& residual, const ConstPtrTable& psf_per_image,
                                int n_images, int n_pol, const SubMinorSetup& s,
                                cudaStream_t stream) {
  SubMinorBuffers& b = h.sub;
  const size_t n = s.width * s.height;
  if (s.width >= 65536 || s.height >= 65536)
    throw std::domain_error("sub-minor loop supports images up to 65535 pixels on a side");
  const int n_blocks = int((n + kSelChunk - 1) / kSelChunk);
  b.flags.Reserve(size_t(n_blocks) * (kSelChunk / 32));
  b.block_counts.Reserve(n_blocks);
  b.block_offsets.Reserve(n_blocks);
  b.state.Reserve(sizeof(LoopState) + 16);
  uint32_t* d_total = reinterpret_cast<uint32_t*>(b.state.get() + sizeof(LoopState));
  b.n_images = n_images;
  b.n_trace = 0;
  b.width = s.width;

  SelectParams sp{};
  sp.spec = linear_spec;
  sp.img = residual;
  sp.rms = s.rms_factor;
  sp.mask = s.mask;
  sp.width = int(s.width);
  sp.height = int(s.height);
  // subminor_loop.cc:150-157
  sp.x0 = int(s.h_border);
  sp.x1 = int(std::max<long>(long(s.h_border), long(s.width) - long(s.h_border)));
  sp.y0 = int(s.v_border);
  sp.y1 = int(std::max<long>(long(s.v_border), long(s.height) - long(s.v_border)));
  sp.threshold = s.threshold;
  sp.allow_negative = s.allow_negative;
  SelectCountKernel<<<n_blocks, kSelThreads, 0, stream>>>(sp, b.flags.get(), b.block_counts.get());
  ScanKernel<<<1, 1024, 0, stream>>>(b.block_counts.get(), b.block_offsets.get(), n_blocks, d_total);
  h.launches += 2;
  uint32_t n_sel = 0;
  RDL_CUDA(cudaMemcpyAsync(&n_sel, d_total, sizeof(n_sel), cudaMemcpyDeviceToHost, stream));
  RDL_CUDA(cudaStreamSynchronize(stream));
  b.n_selected = n_sel;
  h.last_selected = n_sel;

  SubMinorOutcome out;
  out.iteration = s.current_iteration;
  out.n_selected = n_sel;
  if (n_sel == 0) return out;  // Run returns {false, {}} (subminor_loop.cc:47-49)

  b.positions.Reserve(n_sel);
  if (b.capacity < size_t(n_sel) * n_images) {
    b.capacity = size_t(n_sel) * n_images * 5 / 4;
    b.residual.Resize(b.capacity);
    b.model.Resize(b.capacity);
  }
  RDL_CUDA(cudaMemsetAsync(b.model.get(), 0, size_t(n_sel) * n_images * sizeof(float), stream));
  if (s.rms_factor) b.rms.Reserve(n_sel);

  GatherParams gp{};
  gp.img = residual;
  for (int i = 0; i < n_images; ++i) gp.resid.p[i] = b.residual.get() + size_t(i) * n_sel;
  gp.rms = s.rms_factor;
  gp.rms_out = s.rms_factor ? b.rms.get() : nullptr;
  gp.pos = b.positions.get();
  gp.n_images = n_images;
  gp.width = int(s.width);
  gp.n = n;
  GatherKernel<<<n_blocks, kSelThreads, 0, stream>>>(gp, b.flags.get(), b.block_offsets.get());
  ++h.launches;

  const int grid = std::clamp(int((size_t(n_sel) + 1023) / 1024), 1, MaxCooperativeBlocks());
  b.candidates.Reserve(size_t(grid) * 2 * kCandStride);
  h.trace_dev.Reserve(std::max<uint64_t>(h.trace_capacity, 1));

  LoopState init{};
  init.iteration = s.current_iteration;
  RDL_CUDA(cudaMemcpyAsync(b.state.get(), &init, sizeof(init), cudaMemcpyHostToDevice, stream));

  LoopParams lp{};
  lp.resid = gp.resid;
  for (int i = 0; i < n_images; ++i) lp.model.p[i] = b.model.get() + size_t(i) * n_sel;
  lp.psf = psf_per_image;
  lp.pos = b.positions.get();
  lp.rms = gp.rms_out;
  lp.spec = linear_spec;
  lp.n_images = n_images;
  lp.n_sel = int(n_sel);
  lp.width = int(s.width);
  lp.height = int(s.height);
  lp.threshold = s.threshold;
  lp.gain = s.gain;
  lp.divergence_limit = s.divergence_limit;
  lp.allow_negative = s.allow_negative;
  lp.stop_on_negative = s.stop_on_negative;
  lp.max_iterations = s.max_iterations;
  lp.projection = h.fit_mode == RDL_FIT_POLYNOMIAL ? h.projection.get() : nullptr;
  lp.n_fit = int(h.fit_freqs.size());
  lp.n_pol = n_pol;
  lp.trace = h.trace_dev.get();
  lp.trace_capacity = h.trace_capacity;
  lp.trace_scale = s.trace_scale;
  lp.st = reinterpret_cast<LoopState*>(b.state.get());
  lp.cand = b.candidates.get();
  void* args[] = {&lp};
  RDL_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(SubMinorLoopKernel), dim3(grid),
                                       dim3(kLoopThreads), args, 0, stream));
  ++h.launches;
  LoopState fin{};
  RDL_CUDA(cudaMemcpyAsync(&fin, b.state.get(), sizeof(fin), cudaMemcpyDeviceToHost, stream));
  RDL_CUDA(cudaStreamSynchronize(stream));
  out.diverging = fin.diverging != 0;
  out.has_value = true;
  out.max_value = fin.max_value;
  out.iteration = fin.iteration;
  out.flux_cleaned = fin.flux_cleaned;
  b.n_trace = std::min<uint64_t>(fin.n_trace, h.trace_capacity);
  return out;
}

void SubMinorScatterModel(rdl_handle& h, int image_index, float* full, size_t n_pixels,
                          cudaStream_t stream) {
  SubMinorBuffers& b = h.sub;
  RDL_CUDA(cudaMemsetAsync(full, 0, n_pixels * sizeof(float), stream));
  if (b.n_selected == 0) return;
  ScatterKernel<<<ScatterGrid(b.n_selected), 256, 0, stream>>>(
      b.positions.get(), b.model.get() + size_t(image_index) * b.n_selected, full,
      int(b.n_selected), int(b.width), 0);
  ++h.launches;
}

void SubMinorAddModelTo(rdl_handle& h, int image_index, float* model_full, cudaStream_t stream) {
  SubMinorBuffers& b = h.sub;
  if (b.n_selected == 0) return;
  ScatterKernel<<<ScatterGrid(b.n_selected), 256, 0, stream>>>(
      b.positions.get(), b.model.get() + size_t(image_index) * b.n_selected, model_full,
      int(b.n_selected), int(b.width), 1);
  ++h.launches;
}

void SubMinorUpdateAutoMask(rdl_handle& h, int n_images, uint8_t* mask, cudaStream_t stream) {
  SubMinorBuffers& b = h.sub;
  if (b.n_selected == 0) return;
  PtrTable model{};
  for (int i = 0; i < n_images; ++i) model.p[i] = b.model.get() + size_t(i) * b.n_selected;
  AutoMaskKernel<<<ScatterGrid(b.n_selected), 256, 0, stream>>>(
      b.positions.get(), model, n_images, int(b.n_selected), int(b.width), mask);
  ++h.launches;
}

// SubMinorLoop::UpdateComponentList (subminor_loop.cc:230-246): one component per selected pixel
// whose model is non-zero in any image, in the order of the selection (row-major).
void SubMinorUpdateComponentList(rdl_handle& h, int n_images, size_t scale_index,
                                 cudaStream_t stream) {
  SubMinorBuffers& b = h.sub;
  if (b.n_selected == 0) return;
  std::vector<float> model(size_t(n_images) * b.n_selected);
  std::vector<uint32_t> positions(b.n_selected);
  RDL_CUDA(cudaMemcpyAsync(model.data(), b.model.get(), model.size() * sizeof(float),
                           cudaMemcpyDeviceToHost, stream));
  RDL_CUDA(cudaMemcpyAsync(positions.data(), b.positions.get(),
                           positions.size() * sizeof(uint32_t), cudaMemcpyDeviceToHost, stream));
  RDL_CUDA(cudaStreamSynchronize(stream));
  float values[kMaxImages];
  for (size_t px = 0; px != b.n_selected; ++px) {
    bool non_zero = false;
    for (int i = 0; i != n_images; ++i) {
      values[i] = model[size_t(i) * b.n_selected + px];
      non_zero = non_zero || values[i] != 0.0f;
    }
    if (non_zero)
      h.components.Add(positions[px] & 0xffffu, positions[px] >> 16, scale_index, values);
  }
}

void SubMinorCollectTrace(rdl_handle& h, cudaStream_t stream) {
  const uint64_t n = h.sub.n_trace;
  if (!n) return;
  const size_t old = h.trace_host.size();
  h.trace_host.resize(old + n);
  RDL_CUDA(cudaMemcpyAsync(h.trace_host.data() + old, h.trace_dev.get(), n * sizeof(TraceEntry),
                           cudaMemcpyDeviceToHost, stream));
  RDL_CUDA(cudaStreamSynchronize(stream));
  h.sub.n_trace = 0;
}

// GenericClean::ExecuteMajorIteration with use_sub_minor_optimization (generic_clean.cc:56-162,
// 208-252).
rdl_result RunGenericCleanSubMinor(rdl_handle& h, DeviceCube& cube, cudaStream_t stream) {
  const size_t width = cube.width, height = cube.height, n = cube.Pixels();
  const uint64_t iteration_at_start = h.iteration_number;
  h.reference_convolutions.clear();
  // generic_clean.cc:63-66: float 1.1f padding, rounded up to even (NOT the 2357-smooth size)
  const float padding = 1.1f;
  size_t conv_w = size_t(std::ceil(padding * float(width)));
  size_t conv_h = size_t(std::ceil(padding * float(height)));
  if (conv_w % 2 != 0) ++conv_w;
  if (conv_h % 2 != 0) ++conv_h;
  // the FFT plan is made (or refused) before anything is changed: a failing call leaves the
  // handle and the images untouched
  FftConvolver& conv = GetConvolver(h, conv_w, conv_h, width, height);

  const IntegSpec linear = MakeLinearSpec(cube);
  ConstPtrTable resid{};
  ConstPtrTable psf_per_image{};
  for (size_t i = 0; i != cube.n_images; ++i) {
    resid.p[i] = cube.residual.p[i];
    psf_per_image.p[i] = cube.psf.p[cube.PsfIndex(i)];
  }
  const size_t hb = size_t(std::round(width * h.settings.clean_border_ratio));
  const size_t vb = size_t(std::round(height * h.settings.clean_border_ratio));
  const PeakWindow window = MakePeakWindow(width, height, 0, height, hb, vb, h.mask.get(),
                                           h.rms.get(), h.settings.allow_negative_components);
  h.key.Reserve(1);
  auto find_peak = [&](const IntegSpec& spec, const ConstPtrTable& t, bool& has, float& value) {
    LaunchFindPeak(spec, t, int(width), int(height), window, h.key.get(), cube.vec, stream);
    ++h.launches;
    unsigned long long key = 0;
    RDL_CUDA(cudaMemcpyAsync(&key, h.key.get(), sizeof(key), cudaMemcpyDeviceToHost, stream));
    RDL_CUDA(cudaStreamSynchronize(stream));
    uint32_t index = 0;
    has = DecodePeakKey(key, value, index);
  };

  rdl_result r{};
  bool has = false;
  float max_value = 0.0f;
  find_peak(linear, resid, has, max_value);
  r.has_starting_peak = has;
  r.starting_peak_value = has ? max_value : 0.0f;
  r.final_peak_value = r.starting_peak_value;
  r.iteration_number = h.iteration_number;
  h.trace_host.clear();
  if (!has) return r;
  if (h.iteration_number >= h.settings.max_iterations) return r;

  const float initial_max = std::fabs(max_value);
  float first_threshold = h.settings.threshold;
  const float major_iter_threshold = std::max(
      h.settings.major_iteration_threshold, initial_max * (1.0f - h.settings.major_loop_gain));
  if (major_iter_threshold > first_threshold) first_threshold = major_iter_threshold;

  SubMinorSetup s;
  s.width = width;
  s.height = height;
  s.threshold = first_threshold;
  s.gain = h.settings.minor_loop_gain;
  s.h_border = hb;
  s.v_border = vb;
  s.current_iteration = h.iteration_number;
  s.max_iterations = h.settings.max_iterations;
  s.allow_negative = h.settings.allow_negative_components;
  s.stop_on_negative = h.settings.stop_on_negative_component;
  s.mask = h.mask.get();
  s.rms_factor = h.rms.get();
  s.divergence_limit = h.settings.divergence_limit;
  s.trace_scale = 0;
  const SubMinorOutcome o = RunSubMinorLoop(h, linear, resid, psf_per_image, int(cube.n_images),
                                            int(cube.n_pol), s, stream);
  SubMinorCollectTrace(h, stream);
  h.iteration_number = o.iteration;
  bool diverging = o.diverging;
  has = o.has_value;
  max_value = o.max_value;

  // CorrectResidualDirty per image (subminor_loop.cc:195-218) + model += components
  h.fft_scratch.Reserve(n);
  h.fft_spectrum.Reserve(conv.SpectrumSize());
  size_t spectrum_of_psf = size_t(-1);
  for (size_t i = 0; i != cube.n_images; ++i) {
    if (o.n_selected == 0) break;  // convolving an all-zero model changes nothing
    const size_t p = cube.PsfIndex(i);
    if (p != spectrum_of_psf) {
      conv.MakePsfKernelSpectrum(cube.psf.p[p], width, height, h.fft_spectrum.get(), stream);
      spectrum_of_psf = p;
    }
    SubMinorScatterModel(h, int(i), h.fft_scratch.get(), n, stream);
    ++h.reference_convolutions[{conv_w, conv_h}];
    conv.ConvolveWith(h.fft_scratch.get(), width, height, h.fft_spectrum.get(),
                      cube.residual.p[i], width, height, 1, stream);
    SubMinorAddModelTo(h, int(i), cube.model.p[i], stream);
  }
  h.launches += conv.TakeLaunchCount();
  // generic_clean.cc:155-161: without a selected pixel the reference searches `integrated`,
  // which holds the all-zero trimmed convolution by then: still no peak.

  r.iteration_number = h.iteration_number;
  if (diverging) {
    if (has) r.final_peak_value = max_value;
    r.another_iteration_required = 0;
    r.is_diverging = 1;
  } else if (has) {
    const bool final_threshold_reached =
        std::fabs(max_value) <= h.settings.threshold || max_value == 0.0f;
    const bool negative_reached = max_value < 0.0f && h.settings.stop_on_negative_component;
    const bool mgain_reached = std::fabs(max_value) <= major_iter_threshold;
    const bool did_work = (h.iteration_number - iteration_at_start) != 0;
    r.another_iteration_required =
        mgain_reached && did_work && !negative_reached && !final_threshold_reached;
    r.final_peak_value = max_value;
  } else {
    r.another_iteration_required = 0;
  }
  RDL_CUDA(cudaStreamSynchronize(stream));
  return r;
}

}  // namespace rdl
