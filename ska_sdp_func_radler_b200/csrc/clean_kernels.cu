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
#include <cstdlib>
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
    const unsigned long long k_log = iteration - p.log_first_iteration;
    for (int i = lane; i < p.n_images; i += 32) {
      const float v = __fmul_rn(s_values[i], p.gain);
      if (p.log_values) {
        p.log_values[k_log * p.n_images + i] = v;
      } else if (owner) {
        float* m = p.model.p[i] + local_index;
        *m = __fadd_rn(*m, v);
      }
      st->factors[i] = v;
    }
    if (p.log_values && lane == 0) p.log_index[k_log] = index;
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

template <int VEC, bool FIRST, bool SIMPLE, int U>
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
      // U images at a time: all their loads (image vectors, then the shifted PSF values) are
      // issued before the first store. Written out by hand because the compiler cannot move a
      // load of image k + 1 above the store to image k (the planes may alias for all it knows),
      // which left one 16-byte load in flight per thread inside the window.
      for (int k0 = 0; k0 < p.n_images; k0 += U) {
        float r[U][VEC];
#pragma unroll
        for (int u = 0; u < U; ++u)
          if (k0 + u < p.n_images) Vec<VEC>::Load(p.img.p[k0 + u] + i0, r[u]);
        if (in_window) {
          float q[U][VEC];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            if (k0 + u < p.n_images) {
              const float* psf = p.psf.p[k0 + u] + p0;
#pragma unroll
              for (int l = 0; l < VEC; ++l) {
                const int x = x0 + l;
                q[u][l] = (x >= sw.sx && x < sw.ex) ? __ldg(psf + l) : 0.0f;
              }
            }
          }
#pragma unroll
          for (int u = 0; u < U; ++u) {
            if (k0 + u < p.n_images) {
              const float f = s_factors[k0 + u];
#pragma unroll
              for (int l = 0; l < VEC; ++l) {
                const int x = x0 + l;
                if (x >= sw.sx && x < sw.ex) r[u][l] = __fmaf_rn(-q[u][l], f, r[u][l]);
              }
              Vec<VEC>::Store(p.img.p[k0 + u] + i0, r[u]);
            }
          }
        }
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int k = k0 + u;
          if (k < p.n_images) {
            if (spec.mode == kIntegCopy) {
#pragma unroll
              for (int l = 0; l < VEC; ++l) val[l] = r[u][l];
            } else if (k == 0) {
#pragma unroll
              for (int l = 0; l < VEC; ++l) val[l] = __fmul_rn(r[u][l], spec.w[0]);
            } else {
#pragma unroll
              for (int l = 0; l < VEC; ++l) val[l] = __fmaf_rn(r[u][l], spec.w[k], val[l]);
            }
          }
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

// ---- persistent loop for frames that live in L2 ---------------------------------------------
// The whole path-B loop (generic_clean.cc:163-207) in ONE cooperative launch: at K1 size
// (1024^2 x 1 = 4 MB) a step is a few microseconds of L2 traffic, so a kernel launch plus the
// last-CTA control per iteration (13 us measured) is most of the time. Here every CTA keeps a
// fixed share of the pixels; per iteration it subtracts the previous component inside the
// window, re-integrates, joins the arg-max through one atomicMax and ONE grid barrier, and then
// runs the loop control redundantly (same inputs, same arithmetic => same decisions in every
// CTA). The per-image values at a CTA's best pixel travel with its candidate (two slots per CTA,
// alternating), because the owner may already be subtracting the next component from that pixel
// when a slower CTA gets to read it. Same arithmetic, operation order and tie-break as
// HogbomStepKernel + HogbomControl: the results are the same bits.
constexpr int kPersistThreads = 1024;
constexpr int kPersistCandStride = kMaxImages;

// Exchange between the CTAs without a counter barrier: every CTA publishes one 16-byte record
// per iteration {key, tag = iteration sequence number, payload}, written with a single 128-bit
// store, into its own slot (two sets of slots, alternating); every CTA then polls ALL slots of
// the set until each carries the current tag. A slot set is reused two iterations later, and a
// CTA can only get there after every other CTA has published for the iteration in between,
// i.e. after it has finished reading this one. The dependent chain per iteration is one store
// and one (polled) load through L2 instead of fence + atomic + flag + load.
struct TaggedRecord {
  unsigned long long key;
  uint32_t tag;
  uint32_t payload;
};

__device__ __forceinline__ void StoreRecord(void* slot, unsigned long long key, uint32_t tag,
                                            uint32_t payload) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(slot),
               "r"(uint32_t(key)), "r"(uint32_t(key >> 32)), "r"(tag), "r"(payload)
               : "memory");
}
__device__ __forceinline__ TaggedRecord LoadRecord(const void* slot) {
  uint32_t a, b, c, d;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
               : "=r"(a), "=r"(b), "=r"(c), "=r"(d)
               : "l"(slot)
               : "memory");
  TaggedRecord r;
  r.key = (unsigned long long)(a) | ((unsigned long long)(b) << 32);
  r.tag = c;
  r.payload = d;
  return r;
}

struct PersistSync {
  unsigned long long key[3];
  unsigned int bar_count, bar_gen;
};

__device__ __forceinline__ void PersistBarrier(PersistSync* sy) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    volatile unsigned int* gen = &sy->bar_gen;
    const unsigned int g = *gen;
    if (atomicAdd(&sy->bar_count, 1u) == gridDim.x - 1) {
      sy->bar_count = 0u;
      __threadfence();
      atomicAdd(&sy->bar_gen, 1u);
    } else {
      while (*gen == g) {
      }
    }
    __threadfence();
  }
  __syncthreads();
}

template <int VEC>
__global__ void __launch_bounds__(kPersistThreads, 1)
    HogbomPersistentKernel(const __grid_constant__ HogbomParams p, CleanState* st, PersistSync* sy,
                           float* cand) {
  __shared__ unsigned long long red[32];
  __shared__ float s_values[kMaxImages];  // peak values, then factors of the current component
  __shared__ unsigned long long s_key;
  const int width = p.width, height = p.height;
  const size_t n_vec = size_t(width) * height / VEC;
  const size_t total_threads = size_t(gridDim.x) * kPersistThreads;
  const size_t gtid = size_t(blockIdx.x) * kPersistThreads + threadIdx.x;
  const ConstPtrTable& cimg = reinterpret_cast<const ConstPtrTable&>(p.img);
  unsigned long long iteration = st->iteration, n_trace = 0;
  float initial_max = 0.0f, first_threshold = 0.0f, mit = 0.0f;
  float start_value = 0.0f, max_value = 0.0f;
  bool first = true, has_start = false, has_peak = false, diverging = false;
  int done = 0, slot = 0, parity = 0, px = 0, py = 0;
  SubtractWindow sw{};

  for (;;) {
    const IntegSpec& spec = first ? p.first_spec : p.step_spec;
    unsigned long long best = 0ull;
    for (size_t v = gtid; v < n_vec; v += total_threads) {
      const size_t i0 = v * VEC;
      const int y = int(i0 / width), x0 = int(i0 - size_t(y) * width);
      if (!first && y >= sw.sy && y < sw.ey && x0 + VEC > sw.sx && x0 < sw.ex) {
        const size_t p0 = size_t(y - sw.off_y) * width + (x0 - sw.off_x);
        for (int k = 0; k < p.n_images; ++k) {
          float r[VEC];
          Vec<VEC>::Load(p.img.p[k] + i0, r);
          const float* psf = p.psf.p[k] + p0;
          const float f = s_values[k];
#pragma unroll
          for (int l = 0; l < VEC; ++l) {
            const int x = x0 + l;
            if (x >= sw.sx && x < sw.ex) r[l] = __fmaf_rn(-__ldg(psf + l), f, r[l]);
          }
          Vec<VEC>::Store(p.img.p[k] + i0, r);
        }
      }
      float val[VEC];
      IntegrateVec<VEC>(spec, cimg, i0, val);  // re-reads this thread's own stores
      const unsigned long long k = KeysOf<VEC>(val, i0, y, x0, p.window);
      best = k > best ? k : best;
    }
    best = BlockMaxU64(best, red);  // contains a __syncthreads: the block's stores are visible
    if (threadIdx.x == 0) s_key = best;
    __syncthreads();
    best = s_key;
    if (best && threadIdx.x < 32) {
      float v;
      uint32_t index;
      DecodePeakKey(best, v, index);
      float* c = cand + (size_t(blockIdx.x) * 2 + parity) * kPersistCandStride;
      for (int i = threadIdx.x; i < p.n_images; i += 32) __stcg(c + i, p.img.p[i][index]);
      if (threadIdx.x == 0) atomicMax(&sy->key[slot], best);
    }
    PersistBarrier(sy);

    // ---- everyone reads the winner and runs the control (HogbomControl) --------------------
    const unsigned long long key = __ldcg(&sy->key[slot]);
    float val = 0.0f;
    uint32_t index = 0;
    const bool has = DecodePeakKey(key, val, index);
    if (has) {
      const size_t owner_thread = (size_t(index) / VEC) % total_threads;
      const float* c =
          cand + (size_t(owner_thread / kPersistThreads) * 2 + parity) * kPersistCandStride;
      if (threadIdx.x < p.n_images) s_values[threadIdx.x] = __ldcg(c + threadIdx.x);
    }
    if (first) {
      if (!has) {
        done = 2;
      } else if (iteration >= p.max_iterations) {
        done = 3;
      }
      has_start = has;
      start_value = has ? val : 0.0f;
    } else {
      if (has && p.divergence_limit != 0.0f)
        diverging = fabsf(val) > __fmul_rn(initial_max, p.divergence_limit);
      else
        diverging = false;
      ++iteration;
    }
    if (first && !done) {
      initial_max = fabsf(val);
      mit = fmaxf(p.major_iteration_threshold, __fmul_rn(initial_max, __fsub_rn(1.0f, p.mgain)));
      first_threshold = p.threshold;
      if (mit > first_threshold) first_threshold = mit;
    }
    if (!done) {
      const bool cont = has && fabsf(val) > first_threshold && iteration < p.max_iterations &&
                        !(val < 0.0f && p.stop_on_negative) && !diverging;
      if (!cont) done = 1;
    }
    has_peak = has;
    if (has) max_value = val;
    px = has ? int(index % uint32_t(width)) : width;
    py = has ? int(index / uint32_t(width)) : height;
    __syncthreads();  // s_values are in place
    if (done) break;

    if (threadIdx.x < 32) {
      const int lane = threadIdx.x;
      if (p.projection) {
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
        const float f = __fmul_rn(s_values[i], p.gain);
        s_values[i] = f;
        if (blockIdx.x == 0) {
          if (p.log_values) {
            p.log_values[(iteration - p.log_first_iteration) * p.n_images + i] = f;
          } else {
            float* m = p.model.p[i] + index;
            *m = __fadd_rn(*m, f);
          }
        }
      }
      if (blockIdx.x == 0 && lane == 0 && p.log_values)
        p.log_index[iteration - p.log_first_iteration] = index;
      if (blockIdx.x == 0 && lane == 0) {
        if (n_trace < p.trace_capacity) {
          p.trace[n_trace] = TraceEntry{uint32_t(iteration), 0u, uint32_t(px), uint32_t(py), val};
          ++n_trace;
        }
        sy->key[(slot + 2) % 3] = 0ull;  // the slot after next; a barrier separates its use
      }
    }
    __syncthreads();
    sw = MakeSubtractWindow(px, py, width, height);
    first = false;
    slot = (slot + 1) % 3;
    parity ^= 1;
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    st->iteration = iteration;
    st->has_peak = has_peak ? 1 : 0;
    st->max_value = max_value;
    st->has_start = has_start ? 1 : 0;
    st->start_value = start_value;
    st->initial_max = initial_max;
    st->first_threshold = first_threshold;
    st->major_iter_threshold = mit;
    st->diverging = diverging ? 1 : 0;
    st->px = px;
    st->py = py;
    st->done = done;
    st->n_trace = n_trace;
  }
}

// ---- the same loop with the cube in REGISTERS ------------------------------------------------
// When every thread's share of the cube is at most 4 float4 (K1: 2 float4 of one image) the images are loaded once, stay in registers for the whole run and are
// written back at the end: per iteration only the shifted PSF values come from L2, and the
// dependent chain is PSF load -> arg-max -> candidate store -> grid barrier -> one read of all
// candidates. Candidates (key + the C values at the CTA's best pixel) go to a slot per CTA, two
// sets alternating; after the barrier every CTA reads all of them at once and reduces locally,
// so there is no atomicMax and no second dependent read. Specs that are one term per image only
// (IsSimpleSpec); float4 layout only. Same arithmetic as HogbomStepKernel's SIMPLE path.
struct ResidentSlot {  // two tagged 16-byte halves
  TaggedRecord head;    // key, tag, value of image 0
  float v1, v2, v3;     // values of images 1..3
  uint32_t tag_b;
};

template <int VPT, int C>
__global__ void __launch_bounds__(kPersistThreads, 1)
    HogbomResidentKernel(const __grid_constant__ HogbomParams p, CleanState* st,
                         ResidentSlot* slots) {
  __shared__ unsigned long long red[32];
  __shared__ float s_values[4];
  __shared__ unsigned long long s_key;
  const int width = p.width, height = p.height;
  const size_t n_vec = size_t(width) * height / 4;
  const size_t total_threads = size_t(gridDim.x) * kPersistThreads;
  const size_t gtid = size_t(blockIdx.x) * kPersistThreads + threadIdx.x;
  float r[VPT][C][4];
#pragma unroll
  for (int j = 0; j < VPT; ++j) {
    const size_t v = gtid + size_t(j) * total_threads;
    if (v < n_vec) {
#pragma unroll
      for (int k = 0; k < C; ++k) Vec<4>::Load(p.img.p[k] + v * 4, r[j][k]);
    }
  }
  unsigned long long iteration = st->iteration, n_trace = 0;
  float initial_max = 0.0f, first_threshold = 0.0f, mit = 0.0f;
  float start_value = 0.0f, max_value = 0.0f;
  float factors[C];
  float* model_ptr[C];
  float model_old[C], model_add[C];
  bool model_pending = false;  // thread 0 of CTA 0 only
  bool first = true, has_start = false, has_peak = false, diverging = false;
  int done = 0, parity = 0, px = 0, py = 0;
  uint32_t seq = 0;
  SubtractWindow sw{};

  for (;;) {
    ++seq;
    const IntegSpec& spec = first ? p.first_spec : p.step_spec;
    unsigned long long best = 0ull;
    int best_j = 0;
#pragma unroll
    for (int j = 0; j < VPT; ++j) {
      const size_t v = gtid + size_t(j) * total_threads;
      if (v >= n_vec) continue;
      const uint32_t i0 = uint32_t(v) * 4u;  // frames are below 2^32 pixels
      const int y = int(i0 / uint32_t(width)), x0 = int(i0 - uint32_t(y) * uint32_t(width));
      if (!first && y >= sw.sy && y < sw.ey && x0 + 4 > sw.sx && x0 < sw.ex) {
        // PSF values of pixels x0..x0+3: columns cpos..cpos+3 of PSF row y - off_y. cpos is not a
        // multiple of 4 in general, but its remainder is the same for every thread (x0 is), so two
        // aligned 128-bit loads and a warp-uniform select replace four scalar loads.
        const int cpos = x0 - sw.off_x, m = cpos & 3, a0 = cpos - m;
        float q[C][4];
#pragma unroll
        for (int k = 0; k < C; ++k) {
          const float* row = p.psf.p[k] + size_t(y - sw.off_y) * width;
          float4 a = make_float4(0.0f, 0.0f, 0.0f, 0.0f), b = a;
          if (a0 >= 0) a = __ldg(reinterpret_cast<const float4*>(row + a0));
          if (m != 0 && a0 + 4 < width) b = __ldg(reinterpret_cast<const float4*>(row + a0 + 4));
          switch (m) {
            case 0: q[k][0] = a.x; q[k][1] = a.y; q[k][2] = a.z; q[k][3] = a.w; break;
            case 1: q[k][0] = a.y; q[k][1] = a.z; q[k][2] = a.w; q[k][3] = b.x; break;
            case 2: q[k][0] = a.z; q[k][1] = a.w; q[k][2] = b.x; q[k][3] = b.y; break;
            default: q[k][0] = a.w; q[k][1] = b.x; q[k][2] = b.y; q[k][3] = b.z; break;
          }
        }
#pragma unroll
        for (int k = 0; k < C; ++k) {
#pragma unroll
          for (int l = 0; l < 4; ++l) {
            const int x = x0 + l;
            if (x >= sw.sx && x < sw.ex) r[j][k][l] = __fmaf_rn(-q[k][l], factors[k], r[j][k][l]);
          }
        }
      }
      float val[4];
#pragma unroll
      for (int k = 0; k < C; ++k) {
        if (spec.mode == kIntegCopy) {
#pragma unroll
          for (int l = 0; l < 4; ++l) val[l] = r[j][k][l];
        } else if (k == 0) {
#pragma unroll
          for (int l = 0; l < 4; ++l) val[l] = __fmul_rn(r[j][0][l], spec.w[0]);
        } else {
#pragma unroll
          for (int l = 0; l < 4; ++l) val[l] = __fmaf_rn(r[j][k][l], spec.w[k], val[l]);
        }
      }
      if (spec.mode != kIntegCopy) {
#pragma unroll
        for (int l = 0; l < 4; ++l) val[l] = __fmul_rn(val[l], spec.factor);
      }
      const unsigned long long key = KeysOf<4>(val, i0, y, x0, p.window);
      if (key > best) {
        best = key;
        best_j = j;
      }
    }
    if (model_pending) {
#pragma unroll
      for (int k = 0; k < C; ++k) __stcg(model_ptr[k], __fadd_rn(model_old[k], model_add[k]));
      model_pending = false;
    }
    const unsigned long long mine = best;
    best = BlockMaxU64(best, red);
    if (threadIdx.x == 0) s_key = best;
    __syncthreads();
    best = s_key;
    ResidentSlot* my_slot = slots + size_t(parity) * gridDim.x + blockIdx.x;
    if (best == 0ull) {
      if (threadIdx.x == 0) {
        if (C > 1)
          asm volatile("st.volatile.global.v4.u32 [%0], {%1, %1, %1, %2};" ::"l"(&my_slot->v1), "r"(0u),
                       "r"(seq)
                       : "memory");
        StoreRecord(&my_slot->head, 0ull, seq, 0u);
      }
    } else if (mine == best) {  // exactly one thread: keys carry the pixel index
      float v;
      uint32_t index;
      DecodePeakKey(best, v, index);
      float vals[C];
#pragma unroll
      for (int j = 0; j < VPT; ++j) {
        if (j != best_j) continue;
        const int l_sel = int(index - uint32_t(gtid + size_t(j) * total_threads) * 4u);
#pragma unroll
        for (int k = 0; k < C; ++k) {
#pragma unroll
          for (int l = 0; l < 4; ++l)
            if (l == l_sel) vals[k] = r[j][k][l];
        }
      }
      float w4[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
      for (int k = 0; k < C; ++k) w4[k] = vals[k];
      if (C > 1)
        asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(&my_slot->v1),
                     "r"(__float_as_uint(w4[1])), "r"(__float_as_uint(w4[2])),
                     "r"(__float_as_uint(w4[3])), "r"(seq)
                     : "memory");
      StoreRecord(&my_slot->head, best, seq, __float_as_uint(w4[0]));
    }

    // ---- every CTA collects all candidates and runs the control (HogbomControl) --------------
    unsigned long long key = 0ull;
    float cv[4] = {0.0f, 0.0f, 0.0f, 0.0f};
    if (threadIdx.x < gridDim.x) {
      const ResidentSlot* sl = slots + size_t(parity) * gridDim.x + threadIdx.x;
      TaggedRecord rec = LoadRecord(&sl->head);
      while (rec.tag != seq) rec = LoadRecord(&sl->head);  // (a __nanosleep back-off measured slower)
      key = rec.key;
      cv[0] = __uint_as_float(rec.payload);
      if (C > 1) {
        uint32_t a, b, c, d;
        do {
          asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];"
                       : "=r"(a), "=r"(b), "=r"(c), "=r"(d)
                       : "l"(&sl->v1)
                       : "memory");
        } while (d != seq);
        cv[1] = __uint_as_float(a);
        cv[2] = __uint_as_float(b);
        cv[3] = __uint_as_float(c);
      }
    }
    const unsigned long long my_key = key;
    key = BlockMaxU64(key, red);
    if (threadIdx.x == 0) s_key = key;
    __syncthreads();
    key = s_key;
    if (key != 0ull && my_key == key) {
#pragma unroll
      for (int k = 0; k < C; ++k) s_values[k] = cv[k];
    }
    __syncthreads();
    float val = 0.0f;
    uint32_t index = 0;
    const bool has = DecodePeakKey(key, val, index);
    if (first) {
      if (!has) {
        done = 2;
      } else if (iteration >= p.max_iterations) {
        done = 3;
      }
      has_start = has;
      start_value = has ? val : 0.0f;
    } else {
      diverging = has && p.divergence_limit != 0.0f &&
                  fabsf(val) > __fmul_rn(initial_max, p.divergence_limit);
      ++iteration;
    }
    if (first && !done) {
      initial_max = fabsf(val);
      mit = fmaxf(p.major_iteration_threshold, __fmul_rn(initial_max, __fsub_rn(1.0f, p.mgain)));
      first_threshold = p.threshold;
      if (mit > first_threshold) first_threshold = mit;
    }
    if (!done) {
      const bool cont = has && fabsf(val) > first_threshold && iteration < p.max_iterations &&
                        !(val < 0.0f && p.stop_on_negative) && !diverging;
      if (!cont) done = 1;
    }
    has_peak = has;
    if (has) max_value = val;
    px = has ? int(index % uint32_t(width)) : width;
    py = has ? int(index / uint32_t(width)) : height;
    if (done) break;

    // peak values -> spectral fit -> * gain (every thread, redundantly: C <= 4)
    float pv[C];
#pragma unroll
    for (int k = 0; k < C; ++k) pv[k] = s_values[k];
    if (p.projection) {
      float fitted[C];
#pragma unroll
      for (int i = 0; i < C; ++i) {
        const int ch = i / p.n_pol, pol = i - ch * p.n_pol;
        double s = 0.0;
        for (int j = 0; j < p.n_fit; ++j) {
          float pvj = 0.0f;
#pragma unroll
          for (int k = 0; k < C; ++k)
            if (k == j * p.n_pol + pol) pvj = pv[k];
          s = fma(p.projection[ch * p.n_fit + j], double(pvj), s);
        }
        fitted[i] = float(s);
      }
#pragma unroll
      for (int k = 0; k < C; ++k) pv[k] = fitted[k];
    }
#pragma unroll
    for (int k = 0; k < C; ++k) factors[k] = __fmul_rn(pv[k], p.gain);
    if (blockIdx.x == 0 && threadIdx.x == 0) {
      // model += factor, software-pipelined: the load is issued here, the add and the store wait
      // until this thread has done its share of the next iteration (every CTA is waiting for this
      // one's next candidate, so the L2 round trip must not sit in front of it)
      if (p.log_values) {
        const unsigned long long k_log = iteration - p.log_first_iteration;
#pragma unroll
        for (int k = 0; k < C; ++k) p.log_values[k_log * C + k] = factors[k];
        p.log_index[k_log] = index;
      } else {
#pragma unroll
        for (int k = 0; k < C; ++k) {
          model_ptr[k] = p.model.p[k] + index;
          model_old[k] = __ldcg(model_ptr[k]);
          model_add[k] = factors[k];
        }
        model_pending = true;
      }
      if (n_trace < p.trace_capacity) {
        p.trace[n_trace] = TraceEntry{uint32_t(iteration), 0u, uint32_t(px), uint32_t(py), val};
        ++n_trace;
      }
    }
    sw = MakeSubtractWindow(px, py, width, height);
    first = false;
    parity ^= 1;
  }
  if (model_pending) {
#pragma unroll
    for (int k = 0; k < C; ++k) __stcg(model_ptr[k], __fadd_rn(model_old[k], model_add[k]));
  }
  // the cube goes back to memory
#pragma unroll
  for (int j = 0; j < VPT; ++j) {
    const size_t v = gtid + size_t(j) * total_threads;
    if (v >= n_vec) continue;
#pragma unroll
    for (int k = 0; k < C; ++k) Vec<4>::Store(p.img.p[k] + v * 4, r[j][k]);
  }
  if (blockIdx.x == 0 && threadIdx.x == 0) {
    st->iteration = iteration;
    st->has_peak = has_peak ? 1 : 0;
    st->max_value = max_value;
    st->has_start = has_start ? 1 : 0;
    st->start_value = start_value;
    st->initial_max = initial_max;
    st->first_threshold = first_threshold;
    st->major_iter_threshold = mit;
    st->diverging = diverging ? 1 : 0;
    st->px = px;
    st->py = py;
    st->done = done;
    st->n_trace = n_trace;
  }
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
  // knobs for experiments; the defaults (2 images per batch, 8 CTAs per SM = two full waves at 64
  // registers) measured fastest and most repeatable on B200: profiles/r02_hogbom_batch_sweep.jsonl
  static const int batch = [] { const char* e = getenv("RDL_HOGBOM_BATCH"); return e ? atoi(e) : 2; }();
  static const int bps = [] { const char* e = getenv("RDL_HOGBOM_BPS"); return e ? atoi(e) : 8; }();
  const int grid = GridFor(n / vec, bps);
  const bool simple = first ? p.simple_first : p.simple_step;
#define RDL_LAUNCH(V, F, S, UU) HogbomStepKernel<V, F, S, UU><<<grid, kThreads, 0, stream>>>(p, state)
  if (vec == 4) {
    if (first) { if (simple) RDL_LAUNCH(4, true, true, 4); else RDL_LAUNCH(4, true, false, 4); }
    else if (!simple) RDL_LAUNCH(4, false, false, 4);
    else if (batch == 8) RDL_LAUNCH(4, false, true, 8);
    else if (batch == 2) RDL_LAUNCH(4, false, true, 2);
    else RDL_LAUNCH(4, false, true, 4);
  } else {
    if (first) { if (simple) RDL_LAUNCH(1, true, true, 4); else RDL_LAUNCH(1, true, false, 4); }
    else       { if (simple) RDL_LAUNCH(1, false, true, 4); else RDL_LAUNCH(1, false, false, 4); }
  }
#undef RDL_LAUNCH
}

size_t HogbomPersistentScratchBytes() {
  const size_t cand = size_t(kSmCount) * 2 * kPersistCandStride * sizeof(float);
  const size_t slots = size_t(kSmCount) * 2 * sizeof(ResidentSlot);
  return std::max(sizeof(PersistSync) + cand, slots);
}

namespace {
template <int VPT, int C>
void LaunchResident(const HogbomParams& p, CleanState* state, void* scratch, int grid,
                    cudaStream_t stream) {
  HogbomParams params = p;
  ResidentSlot* slots = static_cast<ResidentSlot*>(scratch);
  void* args[] = {&params, &state, &slots};
  RDL_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(HogbomResidentKernel<VPT, C>),
                                       dim3(grid), dim3(kPersistThreads), args, 0, stream));
}
}  // namespace

bool LaunchHogbomPersistent(const HogbomParams& p, CleanState* state, void* scratch, int vec,
                            cudaStream_t stream) {
  static const int max_blocks = [] {
    int a = 0, b = 0, c = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&a, HogbomPersistentKernel<4>, kPersistThreads,
                                                      0) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, HogbomPersistentKernel<1>, kPersistThreads,
                                                      0) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&c, HogbomResidentKernel<4, 1>, kPersistThreads,
                                                      0) != cudaSuccess)
      return 0;
    int sms = 0, dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    return std::min(a, std::min(b, c)) > 0 ? std::min(sms, kSmCount) : 0;
  }();
  if (max_blocks == 0) return false;
  const size_t n_vec = size_t(p.width) * p.height / vec;
  const int grid = int(std::min<size_t>(size_t(max_blocks), (n_vec + kPersistThreads - 1) / kPersistThreads));
  // tags start at 1: a zeroed slot never looks published
  RDL_CUDA(cudaMemsetAsync(scratch, 0, HogbomPersistentScratchBytes(), stream));
  // registers when the share of every thread is at most 4 float4 over all images
  static const bool no_resident = [] { const char* e = getenv("RDL_HOGBOM_RESIDENT"); return e && atoi(e) == 0; }();
  const size_t threads = size_t(grid) * kPersistThreads;
  const size_t vpt = (n_vec + threads - 1) / threads;
  const int c = p.n_images;
  // (single images only: with 2 and 4 images the generic kernel measured as fast or faster)
  if (!no_resident && vec == 4 && p.simple_first && p.simple_step && c == 1 && vpt <= 4) {
    const int v = vpt <= 1 ? 1 : vpt <= 2 ? 2 : 4;
#define RDL_RES(V, CC)                                      \
  if (v == V && c == CC) {                                  \
    LaunchResident<V, CC>(p, state, scratch, grid, stream); \
    return true;                                            \
  }
    RDL_RES(1, 1) RDL_RES(2, 1) RDL_RES(4, 1)
#undef RDL_RES
  }
  PersistSync* sy = static_cast<PersistSync*>(scratch);
  float* cand = reinterpret_cast<float*>(sy + 1);
  HogbomParams params = p;
  void* args[] = {&params, &state, &sy, &cand};
  if (vec == 4)
    RDL_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(HogbomPersistentKernel<4>), dim3(grid),
                                         dim3(kPersistThreads), args, 0, stream));
  else
    RDL_CUDA(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(HogbomPersistentKernel<1>), dim3(grid),
                                         dim3(kPersistThreads), args, 0, stream));
  return true;
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
