// sm_100a kernels of the Hogbom / GenericClean path:
//   * FindPeakKernel      — ImageSet integration fused with the bordered, masked arg-max
//                           (image_set.cc:309-462 + peak_finder.cc:19-131)
//   * SubtractKernel      — SubtractImage of all channels in one coalesced launch
//                           (simple_clean.cc:61-96)
//   * HogbomStepKernel    — one full path-B iteration (generic_clean.cc:169-206) per launch:
//                           PSF subtraction of the previous component + re-integration +
//                           arg-max in ONE pass over the cube; the last block to finish
//                           runs the loop control (thresholds, divergence, spectral fit,
//                           model update) so the host never synchronises inside the loop.
// All of it is HBM-bound streaming work: 128-bit coalesced loads, grid = multiple of the
// SM count, 64-bit packed (value, ~index, sign) keys reduced by warp shuffles and one
// atomicMax per block.
#include "clean_kernels.cuh"

namespace rdl {

namespace {

constexpr int kThreads = 256;

template <int VEC>
struct Vec;
template <>
struct Vec<4> {
  __device__ static void Load(const float* p, float (&v)[4]) {
    const float4 t = *reinterpret_cast<const float4*>(p);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
  __device__ static void Store(float* p, const float (&v)[4]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  }
  __device__ static void LoadMask(const uint8_t* p, bool (&m)[4]) {
    const uchar4 t = *reinterpret_cast<const uchar4*>(p);
    m[0] = t.x; m[1] = t.y; m[2] = t.z; m[3] = t.w;
  }
};
template <>
struct Vec<1> {
  __device__ static void Load(const float* p, float (&v)[1]) { v[0] = *p; }
  __device__ static void Store(float* p, const float (&v)[1]) { *p = v[0]; }
  __device__ static void LoadMask(const uint8_t* p, bool (&m)[1]) { m[0] = *p; }
};

// Integration of VEC consecutive pixels starting at i0 (general, all modes).
template <int VEC>
__device__ __forceinline__ void IntegrateVec(const IntegSpec& s, const ConstPtrTable& t,
                                             size_t i0, float (&out)[VEC]) {
  if (s.mode == kIntegCopy) {
    Vec<VEC>::Load(t.p[s.img[0]] + i0, out);
  } else if (s.mode == kIntegLinear) {
    float v[VEC];
    Vec<VEC>::Load(t.p[s.img[0]] + i0, v);
#pragma unroll
    for (int l = 0; l < VEC; ++l) out[l] = __fmul_rn(v[l], s.w[0]);
#pragma unroll 4
    for (int k = 1; k < s.n_terms; ++k) {
      Vec<VEC>::Load(t.p[s.img[k]] + i0, v);
#pragma unroll
      for (int l = 0; l < VEC; ++l) out[l] = __fmaf_rn(v[l], s.w[k], out[l]);
    }
#pragma unroll
    for (int l = 0; l < VEC; ++l) out[l] = __fmul_rn(out[l], s.factor);
  } else {
#pragma unroll
    for (int l = 0; l < VEC; ++l) out[l] = IntegratePixel(s, t, i0 + l);
  }
}

// Key of VEC integrated values (applies RMS factor, mask and the search window).
template <int VEC>
__device__ __forceinline__ unsigned long long KeysOf(const float (&val)[VEC], size_t i0, int y,
                                                     int x0, const PeakWindow& w,
                                                     uint32_t key_offset = 0u) {
  unsigned long long best = 0ull;
  if (y < w.y0 || y >= w.y1) return best;
  if (x0 + VEC <= w.x0 || x0 >= w.x1) return best;
  float r[VEC];
  bool m[VEC];
  // mask / RMS image are full-frame: index them (and the key) in frame coordinates
  if (w.rms_factor) Vec<VEC>::Load(w.rms_factor + i0 + key_offset, r);
  if (w.mask) Vec<VEC>::LoadMask(w.mask + i0 + key_offset, m);
#pragma unroll
  for (int l = 0; l < VEC; ++l) {
    const int x = x0 + l;
    if (x < w.x0 || x >= w.x1) continue;
    if (w.mask && !m[l]) continue;
    const float v = w.rms_factor ? __fmul_rn(val[l], r[l]) : val[l];
    const unsigned long long k =
        MakePeakKey(v, w.allow_negative != 0, uint32_t(i0 + l) + key_offset);
    best = k > best ? k : best;
  }
  return best;
}

template <int VEC>
__global__ void __launch_bounds__(kThreads)
    FindPeakKernel(IntegSpec spec, ConstPtrTable images, int width, int height,
                   PeakWindow window, unsigned long long* key) {
  __shared__ unsigned long long red[32];
  const size_t n_vec = size_t(width) * height / VEC;
  unsigned long long best = 0ull;
  // Only rows inside the search window are read at all.
  const size_t v_begin = size_t(window.y0) * width / VEC, v_end = size_t(window.y1) * width / VEC;
  for (size_t v = v_begin + size_t(blockIdx.x) * kThreads + threadIdx.x; v < v_end && v < n_vec;
       v += size_t(gridDim.x) * kThreads) {
    const size_t i0 = v * VEC;
    const int y = int(i0 / width), x0 = int(i0 - size_t(y) * width);
    if (x0 + VEC <= window.x0 || x0 >= window.x1) continue;
    float val[VEC];
    IntegrateVec<VEC>(spec, images, i0, val);
    const unsigned long long k = KeysOf<VEC>(val, i0, y, x0, window);
    best = k > best ? k : best;
  }
  best = BlockMaxU64(best, red);
  if (threadIdx.x == 0 && best) atomicMax(key, best);
}

template <int VEC>
__global__ void __launch_bounds__(kThreads)
    IntegrateKernel(IntegSpec spec, ConstPtrTable images, size_t n, float* dest) {
  const size_t n_vec = n / VEC;
  for (size_t v = size_t(blockIdx.x) * kThreads + threadIdx.x; v < n_vec;
       v += size_t(gridDim.x) * kThreads) {
    float val[VEC];
    IntegrateVec<VEC>(spec, images, v * VEC, val);
    Vec<VEC>::Store(dest + v * VEC, val);
  }
}

struct SubtractWindow {
  int off_x, off_y, sx, ex, sy, ey;
};

__device__ __forceinline__ SubtractWindow MakeSubtractWindow(int px, int py, int width,
                                                             int height) {
  // simple_clean.cc:61-96 (PartialSubtractImage over all rows)
  SubtractWindow w;
  w.off_x = px - width / 2;
  w.off_y = py - height / 2;
  w.sx = w.off_x > 0 ? w.off_x : 0;
  w.sy = w.off_y > 0 ? w.off_y : 0;
  w.ex = min(px + width / 2, width);
  w.ey = min(py + height / 2, height);
  return w;
}

struct SubtractFactors {
  float f[kMaxImages];
};

// One thread per float4 of the clipped window, all images per thread.
__global__ void __launch_bounds__(kThreads)
    SubtractKernel(PtrTable images, ConstPtrTable psfs, SubtractFactors factors, int n_images,
                   int width, int height, int px, int py) {
  const SubtractWindow w = MakeSubtractWindow(px, py, width, height);
  const int span = w.ex - w.sx, rows = w.ey - w.sy;
  if (span <= 0 || rows <= 0) return;
  // Groups of 4 pixels aligned to the image's float4 grid; head/tail lanes are predicated.
  const int g0 = w.sx >> 2, g1 = (w.ex + 3) >> 2;
  const int groups = g1 - g0;
  const size_t total = size_t(groups) * rows;
  const bool vec_ok = (width & 3) == 0;
  for (size_t t = size_t(blockIdx.x) * kThreads + threadIdx.x; t < total;
       t += size_t(gridDim.x) * kThreads) {
    const int row = int(t / groups), g = int(t - size_t(row) * groups) + g0;
    const int y = w.sy + row, x0 = g << 2;
    const size_t i0 = size_t(y) * width + x0;
    const size_t p0 = size_t(y - w.off_y) * width + (x0 - w.off_x);
    const bool full = vec_ok && x0 >= w.sx && x0 + 4 <= w.ex;
    for (int k = 0; k < n_images; ++k) {
      float* img = images.p[k];
      const float* psf = psfs.p[k];
      const float f = factors.f[k];
      if (full) {
        float4 v = *reinterpret_cast<const float4*>(img + i0);
        v.x = __fmaf_rn(-__ldg(psf + p0), f, v.x);
        v.y = __fmaf_rn(-__ldg(psf + p0 + 1), f, v.y);
        v.z = __fmaf_rn(-__ldg(psf + p0 + 2), f, v.z);
        v.w = __fmaf_rn(-__ldg(psf + p0 + 3), f, v.w);
        *reinterpret_cast<float4*>(img + i0) = v;
      } else {
        for (int l = 0; l < 4; ++l) {
          const int x = x0 + l;
          if (x >= w.sx && x < w.ex) img[i0 + l] = __fmaf_rn(-__ldg(psf + p0 + l), f, img[i0 + l]);
        }
      }
    }
  }
}

// Loop control run by warp 0 of the last block (generic_clean.cc:99-112, 169-206).
// `key` indexes the width x global_height frame. values_ready: s_values already holds the
// per-image residual values at the peak (row-sharded path); otherwise they are read here.
template <bool FIRST>
__device__ void HogbomControl(const HogbomParams& p, CleanState* st, float* s_values,
                              unsigned long long key, bool values_ready) {
  const int lane = threadIdx.x;
  const int slot = st->slot;
  float val = 0.0f;
  uint32_t index = 0;
  const bool has = DecodePeakKey(key, val, index);
  unsigned long long iteration = st->iteration;
  bool diverging = false;
  int done = 0;
  if (FIRST) {
    if (!has) {
      done = 2;  // "No peak found."
    } else if (iteration >= p.max_iterations) {
      done = 3;  // peak-only probe (ParallelDeconvolution pass 1)
    }
  } else {
    if (has && p.divergence_limit != 0.0f)
      diverging = fabsf(val) > __fmul_rn(st->initial_max, p.divergence_limit);
    ++iteration;
  }
  float first_threshold = st->first_threshold;
  if (FIRST && !done) {
    const float initial_max = fabsf(val);
    const float mit = fmaxf(p.major_iteration_threshold,
                            __fmul_rn(initial_max, __fsub_rn(1.0f, p.mgain)));
    first_threshold = p.threshold;
    if (mit > first_threshold) first_threshold = mit;
    if (lane == 0) {
      st->initial_max = initial_max;
      st->major_iter_threshold = mit;
      st->first_threshold = first_threshold;
    }
  }
  if (!done) {
    const bool cont = has && fabsf(val) > first_threshold && iteration < p.max_iterations &&
                      !(val < 0.0f && p.stop_on_negative) && !diverging;
    if (!cont) done = 1;
  }
  const int x = has ? int(index % uint32_t(p.width)) : p.width;
  const int y = has ? int(index / uint32_t(p.width)) : p.global_height;
  // position inside this rank's row band (the whole frame on a single GPU)
  int y_local = 0;
  const bool owner = has && LocalRow(p, y, y_local);
  const size_t local_index = owner ? size_t(y_local) * p.width + x : 0;
  if (!done) {
    // peakValues[i] = dirty[i][peak]; PerformSpectralFit; *= gain; model[i][peak] += ...
    if (!values_ready)
      for (int i = lane; i < p.n_images; i += 32) s_values[i] = __ldcg(p.img.p[i] + local_index);
    __syncwarp();
    if (p.projection) {
      // values of polarisation pol sit at [ch * n_pol + pol]
      float fitted[2];
      int slot_i = 0;
      for (int i = lane; i < p.n_images; i += 32, ++slot_i) {
        const int ch = i / p.n_pol, pol = i - ch * p.n_pol;
        double s = 0.0;
        for (int j = 0; j < p.n_fit; ++j)
          s = fma(p.projection[ch * p.n_fit + j], double(s_values[j * p.n_pol + pol]), s);
        fitted[slot_i] = float(s);
      }
      __syncwarp();
      slot_i = 0;
      for (int i = lane; i < p.n_images; i += 32, ++slot_i) s_values[i] = fitted[slot_i];
      __syncwarp();
    }
    for (int i = lane; i < p.n_images; i += 32) {
      const float v = __fmul_rn(s_values[i], p.gain);
      if (owner) {
        float* m = p.model.p[i] + local_index;
        *m = __fadd_rn(*m, v);
      }
      st->factors[i] = v;
    }
    if (lane == 0 && st->n_trace < p.trace_capacity) {
      p.trace[st->n_trace] = TraceEntry{uint32_t(iteration), 0u, uint32_t(x), uint32_t(y), val};
      st->n_trace = st->n_trace + 1;
    }
  }
  if (lane == 0) {
    st->key[slot] = 0ull;  // ready for the launch after next
    st->slot = slot ^ 1;
    st->ticket = 0u;
    st->iteration = iteration;
    st->has_peak = has ? 1 : 0;
    if (has) st->max_value = val;
    if (FIRST) {
      st->has_start = has ? 1 : 0;
      st->start_value = has ? val : 0.0f;
    }
    st->diverging = diverging ? 1 : 0;
    st->px = x;
    st->py = y;
    st->done = done;
  }
}

template <int VEC, bool FIRST, bool SIMPLE>
__global__ void __launch_bounds__(kThreads)
    HogbomStepKernel(const __grid_constant__ HogbomParams p, CleanState* st) {
  if (st->done) return;
  __shared__ unsigned long long red[32];
  __shared__ float s_factors[kMaxImages];
  __shared__ int s_is_last;
  const int slot = st->slot;
  const int width = p.width, height = p.height;
  SubtractWindow sw{};
  if (!FIRST) {
    sw = MakeSubtractWindow(st->px, st->py, width, p.global_height);
    if (threadIdx.x < p.n_images) s_factors[threadIdx.x] = st->factors[threadIdx.x];
    __syncthreads();
  }
  const IntegSpec& spec = FIRST ? p.first_spec : p.step_spec;
  const ConstPtrTable& cimg = reinterpret_cast<const ConstPtrTable&>(p.img);
  const size_t n_vec = size_t(width) * height / VEC;
  unsigned long long best = 0ull;
  for (size_t v = size_t(blockIdx.x) * kThreads + threadIdx.x; v < n_vec;
       v += size_t(gridDim.x) * kThreads) {
    const size_t i0 = v * VEC;
    const int y_local = int(i0 / width), x0 = int(i0 - size_t(y_local) * width);
    const int y = FrameRow(p, y_local);  // frame row
    const uint32_t key_offset = uint32_t(y - y_local) * uint32_t(width);
    const bool in_window =
        !FIRST && y >= sw.sy && y < sw.ey && x0 + VEC > sw.sx && x0 < sw.ex;
    float val[VEC];
    if (SIMPLE) {
      // Every image is one linear term (or the single copied image): subtract and
      // accumulate in the same pass, one load and at most one store per image.
      const size_t p0 = in_window ? size_t(y - sw.off_y) * width + (x0 - sw.off_x) : 0;
#pragma unroll 4
      for (int k = 0; k < p.n_images; ++k) {
        float r[VEC];
        Vec<VEC>::Load(p.img.p[k] + i0, r);
        if (in_window) {
          const float* psf = p.psf.p[k] + p0;
          const float f = s_factors[k];
#pragma unroll
          for (int l = 0; l < VEC; ++l) {
            const int x = x0 + l;
            if (x >= sw.sx && x < sw.ex) r[l] = __fmaf_rn(-__ldg(psf + l), f, r[l]);
          }
          Vec<VEC>::Store(p.img.p[k] + i0, r);
        }
        if (spec.mode == kIntegCopy) {
#pragma unroll
          for (int l = 0; l < VEC; ++l) val[l] = r[l];
        } else if (k == 0) {
#pragma unroll
          for (int l = 0; l < VEC; ++l) val[l] = __fmul_rn(r[l], spec.w[0]);
        } else {
#pragma unroll
          for (int l = 0; l < VEC; ++l) val[l] = __fmaf_rn(r[l], spec.w[k], val[l]);
        }
      }
      if (spec.mode != kIntegCopy) {
#pragma unroll
        for (int l = 0; l < VEC; ++l) val[l] = __fmul_rn(val[l], spec.factor);
      }
    } else {
      if (in_window) {
        const size_t p0 = size_t(y - sw.off_y) * width + (x0 - sw.off_x);
        for (int k = 0; k < p.n_images; ++k) {
          float r[VEC];
          Vec<VEC>::Load(p.img.p[k] + i0, r);
          const float* psf = p.psf.p[k] + p0;
          const float f = s_factors[k];
#pragma unroll
          for (int l = 0; l < VEC; ++l) {
            const int x = x0 + l;
            if (x >= sw.sx && x < sw.ex) r[l] = __fmaf_rn(-__ldg(psf + l), f, r[l]);
          }
          Vec<VEC>::Store(p.img.p[k] + i0, r);
        }
      }
      IntegrateVec<VEC>(spec, cimg, i0, val);  // re-reads this thread's own stores
    }
    const unsigned long long k = KeysOf<VEC>(val, i0, y, x0, p.window, key_offset);
    best = k > best ? k : best;
  }
  __threadfence();  // publish this thread's image stores before the ticket
  best = BlockMaxU64(best, red);
  if (threadIdx.x == 0) {
    if (best) atomicMax(&st->key[slot], best);
    __threadfence();
    const unsigned int t = atomicAdd(&st->ticket, 1u);
    s_is_last = (t == gridDim.x - 1);
  }
  __syncthreads();
  if (s_is_last && threadIdx.x < 32) {
    __threadfence();
    const unsigned long long key = __ldcg(&st->key[slot]);
    if (!p.dist) {
      HogbomControl<FIRST>(p, st, s_factors, key, false);
    } else {
      // row-sharded: publish this rank's candidate; the joined control runs after the exchange
      float val;
      uint32_t index;
      const bool has = DecodePeakKey(key, val, index);
      int yl = 0;
      if (has) LocalRow(p, int(index / uint32_t(width)), yl);
      const size_t local_index = has ? size_t(yl) * width + index % uint32_t(width) : 0;
      for (int i = threadIdx.x; i < kMaxImages; i += 32)
        p.cand_local->values[i] =
            (has && i < p.n_images) ? __ldcg(p.img.p[i] + local_index) : 0.0f;
      if (threadIdx.x == 0) {
        p.cand_local->key = key;  // the joined control resets key / slot / ticket
      }
    }
  }
}

template <bool FIRST>
__global__ void __launch_bounds__(32)
    HogbomDistControlKernel(const __grid_constant__ HogbomParams p, CleanState* st,
                            const DistCandidate* all, int world) {
  if (st->done) return;
  __shared__ float s_values[kMaxImages];
  unsigned long long best = 0ull;
  int winner = 0;
  for (int r = 0; r < world; ++r) {
    const unsigned long long k = all[r].key;
    if (k > best) {
      best = k;
      winner = r;
    }
  }
  for (int i = threadIdx.x; i < p.n_images; i += 32) s_values[i] = all[winner].values[i];
  __syncwarp();
  HogbomControl<FIRST>(p, st, s_values, best, true);
}

int GridFor(size_t work_items, int blocks_per_sm) {
  const size_t want = (work_items + kThreads - 1) / kThreads;
  const size_t cap = size_t(kSmCount) * blocks_per_sm;
  return int(want < 1 ? 1 : (want > cap ? cap : want));
}

}  // namespace

void LaunchHogbomStep(const HogbomParams& p, CleanState* state, bool first, int vec,
                      cudaStream_t stream) {
  const size_t n = size_t(p.width) * p.height;
  const int grid = GridFor(n / vec, 8);
  const bool simple = first ? p.simple_first : p.simple_step;
#define RDL_LAUNCH(V, F, S) HogbomStepKernel<V, F, S><<<grid, kThreads, 0, stream>>>(p, state)
  if (vec == 4) {
    if (first) { if (simple) RDL_LAUNCH(4, true, true); else RDL_LAUNCH(4, true, false); }
    else       { if (simple) RDL_LAUNCH(4, false, true); else RDL_LAUNCH(4, false, false); }
  } else {
    if (first) { if (simple) RDL_LAUNCH(1, true, true); else RDL_LAUNCH(1, true, false); }
    else       { if (simple) RDL_LAUNCH(1, false, true); else RDL_LAUNCH(1, false, false); }
  }
#undef RDL_LAUNCH
}

void LaunchHogbomDistControl(const HogbomParams& p, CleanState* state, const DistCandidate* all,
                             int world, bool first, cudaStream_t stream) {
  if (first)
    HogbomDistControlKernel<true><<<1, 32, 0, stream>>>(p, state, all, world);
  else
    HogbomDistControlKernel<false><<<1, 32, 0, stream>>>(p, state, all, world);
}

void LaunchFindPeak(const IntegSpec& spec, const ConstPtrTable& images, int width, int height,
                    const PeakWindow& window, unsigned long long* key, int vec,
                    cudaStream_t stream) {
  RDL_CUDA(cudaMemsetAsync(key, 0, sizeof(unsigned long long), stream));
  const size_t rows = window.y1 > window.y0 ? size_t(window.y1 - window.y0) : 0;
  if (rows == 0 || window.x1 <= window.x0) return;
  const int grid = GridFor(rows * width / vec, 8);
  if (vec == 4)
    FindPeakKernel<4><<<grid, kThreads, 0, stream>>>(spec, images, width, height, window, key);
  else
    FindPeakKernel<1><<<grid, kThreads, 0, stream>>>(spec, images, width, height, window, key);
}

void LaunchIntegrate(const IntegSpec& spec, const ConstPtrTable& images, size_t n, float* dest,
                     int vec, cudaStream_t stream) {
  const int grid = GridFor(n / vec, 8);
  if (vec == 4)
    IntegrateKernel<4><<<grid, kThreads, 0, stream>>>(spec, images, n, dest);
  else
    IntegrateKernel<1><<<grid, kThreads, 0, stream>>>(spec, images, n, dest);
}

void LaunchSubtract(const PtrTable& images, const ConstPtrTable& psfs, int n_images, int width,
                    int height, int x, int y, const float* factors_host, cudaStream_t stream) {
  SubtractFactors f{};
  for (int i = 0; i < n_images; ++i) f.f[i] = factors_host[i];
  const int span = (width / 2) * 2, rows = (height / 2) * 2;
  const int grid = GridFor(size_t(span / 4 + 2) * rows, 8);
  SubtractKernel<<<grid, kThreads, 0, stream>>>(images, psfs, f, n_images, width, height, x, y);
}

}  // namespace rdl
