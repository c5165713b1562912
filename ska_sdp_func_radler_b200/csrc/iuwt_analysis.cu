// IUWT structure selection and plane measurements on the device (iuwt.cuh for the reference
// lines): the significant-structure test, the seeded flood fills over (scale, y, x), the median
// absolute deviation of a coefficient plane and its largest coefficient.
//
// SelectStructures of the reference starts a depth-first fill at every voxel that passes the
// test and is not yet marked. Every such voxel is a seed itself, so the union of the fills is
// exactly the set {exceeds threshold, not marked on entry, inside the prior mask, inside the
// border}: one elementwise pass. Only the fills from ONE seed need connectivity; they run as
// tile-local label propagation on a one-byte state volume (bit 0: may be filled, bit 1: filled),
// repeated until a round changes nothing. Results (masks, areas, positions) are order-free, so
// they are bit-identical to the reference's walk.
#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <limits>
#include <stdexcept>

#include "common.cuh"
#include "iuwt.cuh"

namespace rdl {
namespace {

constexpr int kMaxFillScales = 24;
constexpr int kFillTX = 64, kFillTY = 16;  // tile of the propagation kernel
constexpr int kFillThreads = 256;
constexpr uint8_t kMayFill = 1, kFilled = 2;

struct FillGeometry {
  size_t width, height;
  int n_scales;              // planes in the volume
  int min_scale, end_scale;  // moves along the scale axis stay inside [min_scale, end_scale)
  size_t min_x, max_x, min_y, max_y;
  int use_abs;               // FloodFill2D's fabs(v) > t flavour
  float thresholds[kMaxFillScales];
};

FillGeometry MakeGeometry(size_t width, size_t height, int n_scales, const float* thresholds,
                          int min_scale, int end_scale, float clean_border) {
  if (width == 0 || height == 0 || n_scales < 1 || n_scales > kMaxFillScales)
    throw std::invalid_argument("IUWT structure selection: bad volume size");
  if (min_scale < 0 || end_scale > n_scales || min_scale >= end_scale)
    throw std::invalid_argument("IUWT structure selection: bad scale range");
  if (!(clean_border >= 0.0f) || !(clean_border < 0.5f))
    throw std::invalid_argument("IUWT structure selection: clean border outside [0, 0.5)");
  if (width * height >= (size_t(1) << 32))
    throw std::invalid_argument("IUWT structure selection: plane too large");
  FillGeometry g{};
  g.width = width;
  g.height = height;
  g.n_scales = n_scales;
  g.min_scale = min_scale;
  g.end_scale = end_scale;
  // image_analysis.cc:91-94: float product truncated to size_t
  const size_t bx = clean_border * width, by = clean_border * height;
  g.min_x = bx;
  g.max_x = width - bx;
  g.min_y = by;
  g.max_y = height - by;
  for (int s = 0; s != n_scales; ++s) g.thresholds[s] = thresholds[s];
  return g;
}

// image_analysis.cc:9-19
__device__ __forceinline__ bool Exceeds(float v, float threshold, bool use_abs) {
  if (use_abs) return fabsf(v) > threshold;
  if (threshold >= 0.0f) return v > threshold;
  return v < threshold || v > -threshold;
}

__device__ __forceinline__ void CountInto(unsigned long long* counter, unsigned n) {
  for (int o = 16; o; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
  if ((threadIdx.x & 31) == 0 && n) atomicAdd(counter, static_cast<unsigned long long>(n));
}

// One thread per VEC consecutive pixels of one plane (VEC = 4 needs width % 4 == 0, so the four
// share a row). MODE 0: SelectStructures, masks |= passing set inside the border, count.
// MODE 1: write the fill state (kMayFill where a fill may enter) for a seeded fill.
template <int VEC, int MODE>
__global__ void __launch_bounds__(256)
StructureTestKernel(FillGeometry g, const float* __restrict__ coefficients, uint8_t* masks,
                    const uint8_t* __restrict__ prior, uint8_t* state, int first_scale,
                    unsigned long long* counter) {
  const size_t area = g.width * g.height;
  const int scale = first_scale + blockIdx.y;
  const size_t px = (size_t(blockIdx.x) * blockDim.x + threadIdx.x) * VEC;
  unsigned added = 0;
  if (px < area) {
    const size_t vox = size_t(scale) * area + px;
    const float t = g.thresholds[scale];
    float v[VEC];
    uint8_t m[VEC], p[VEC];
    if (VEC == 4) {
      const float4 v4 = *reinterpret_cast<const float4*>(coefficients + vox);
      v[0] = v4.x, v[1] = v4.y, v[2] = v4.z, v[3] = v4.w;
      const uchar4 m4 = *reinterpret_cast<const uchar4*>(masks + vox);
      m[0] = m4.x, m[1] = m4.y, m[2] = m4.z, m[3] = m4.w;
      uchar4 p4 = make_uchar4(1, 1, 1, 1);
      if (prior) p4 = *reinterpret_cast<const uchar4*>(prior + px);
      p[0] = p4.x, p[1] = p4.y, p[2] = p4.z, p[3] = p4.w;
    } else {
      v[0] = coefficients[vox];
      m[0] = masks[vox];
      p[0] = prior ? prior[px] : 1;
    }
    const size_t y = px / g.width, x0 = px - y * g.width;
    const bool row_in = y >= g.min_y && y < g.max_y;
    uint8_t out[VEC];
    bool any = false;
#pragma unroll
    for (int k = 0; k != VEC; ++k) {
      bool pass = Exceeds(v[k], t, g.use_abs != 0) && !m[k] && p[k];
      if (MODE == 0) {
        pass = pass && row_in && (x0 + k) >= g.min_x && (x0 + k) < g.max_x;
        out[k] = (m[k] || pass) ? 1 : 0;
        added += pass;
        any |= pass;
      } else {
        out[k] = pass ? kMayFill : 0;
      }
    }
    if (MODE == 0) {
      if (any) {
        if (VEC == 4)
          *reinterpret_cast<uchar4*>(masks + vox) = make_uchar4(out[0], out[1], out[2], out[3]);
        else
          masks[vox] = out[0];
      }
    } else {
      if (VEC == 4)
        *reinterpret_cast<uchar4*>(state + vox) = make_uchar4(out[0], out[1], out[2], out[3]);
      else
        state[vox] = out[0];
    }
  }
  if (MODE == 0) CountInto(counter, added);
}

__global__ void SeedKernel(uint8_t* state, size_t voxel) { state[voxel] = kMayFill | kFilled; }

// One CTA owns a kFillTX x kFillTY tile of every plane in [s_lo, s_hi]. It loads the state of
// the tile plus a one-pixel halo, spreads the filled bit until the tile is stable and writes the
// new voxels back. The moves are the reference's (image_analysis.cc:105-146): they are bounded
// by the position of the voxel that is LEFT, which is why the tests below look at the neighbour.
// A tile only has work in round r if it, or one of its four neighbours, filled a voxel in round
// r - 1 (or holds the seed): due[tile] is the latest round the tile was asked to run in.
__global__ void __launch_bounds__(kFillThreads)
FillPropagateKernel(FillGeometry g, uint8_t* state, int s_lo, int s_hi, unsigned* changed,
                    int* due, int round) {
  extern __shared__ uint8_t tile[];
  constexpr int PW = kFillTX + 2, PH = kFillTY + 2;
  const int tile_id = blockIdx.y * gridDim.x + blockIdx.x;
  if (due[tile_id] < round) return;
  const int n_s = s_hi - s_lo + 1;
  const size_t area = g.width * g.height;
  const long x0 = long(blockIdx.x) * kFillTX - 1, y0 = long(blockIdx.y) * kFillTY - 1;
  int seen = 0;
  for (int i = threadIdx.x; i < n_s * PW * PH; i += kFillThreads) {
    const int s = i / (PW * PH), r = i - s * (PW * PH), ty = r / PW, tx = r - ty * PW;
    const long gx = x0 + tx, gy = y0 + ty;
    uint8_t v = 0;
    if (gx >= 0 && gy >= 0 && gx < long(g.width) && gy < long(g.height))
      v = state[size_t(s_lo + s) * area + size_t(gy) * g.width + size_t(gx)];
    tile[i] = v;
    seen |= v & kFilled;
  }
  if (!__syncthreads_or(seen)) return;  // nothing filled in or next to this tile

  bool wrote = false;
  for (;;) {
    int grew = 0;
    for (int i = threadIdx.x; i < n_s * kFillTX * kFillTY; i += kFillThreads) {
      const int s = i / (kFillTX * kFillTY), r = i - s * (kFillTX * kFillTY);
      const int ty = r / kFillTX + 1, tx = r % kFillTX + 1;
      const int at = (s * PH + ty) * PW + tx;
      if (tile[at] != kMayFill) continue;
      const size_t gx = size_t(x0 + tx), gy = size_t(y0 + ty);
      const int scale = s_lo + s;
      bool reach = false;
      // from the right neighbour moving left: c.x > min_x
      reach |= (tile[at + 1] & kFilled) && gx + 1 > g.min_x;
      // from the left neighbour moving right: c.x < max_x - 1
      reach |= (tile[at - 1] & kFilled) && gx < g.max_x;
      reach |= (tile[at + PW] & kFilled) && gy + 1 > g.min_y;
      reach |= (tile[at - PW] & kFilled) && gy < g.max_y;
      // from the plane above moving down: c.scale > min_scale
      if (s + 1 < n_s) reach |= (tile[at + PW * PH] & kFilled) && scale + 1 > g.min_scale;
      // from the plane below moving up: c.scale < end_scale - 1
      if (s > 0) reach |= (tile[at - PW * PH] & kFilled) && scale - 1 < g.end_scale - 1;
      if (reach) {
        tile[at] = kMayFill | kFilled;
        state[size_t(scale) * area + gy * g.width + gx] = kMayFill | kFilled;
        grew = 1;
      }
    }
    if (!__syncthreads_or(grew)) break;
    wrote = true;
  }
  if (wrote && threadIdx.x == 0) {
    atomicAdd(changed, 1u);
    atomicMax(&due[tile_id], round + 1);
    if (blockIdx.x > 0) atomicMax(&due[tile_id - 1], round + 1);
    if (blockIdx.x + 1 < gridDim.x) atomicMax(&due[tile_id + 1], round + 1);
    if (blockIdx.y > 0) atomicMax(&due[tile_id - int(gridDim.x)], round + 1);
    if (blockIdx.y + 1 < gridDim.y) atomicMax(&due[tile_id + int(gridDim.x)], round + 1);
  }
}

// the seed's tile and its neighbours run in round 0
__global__ void SeedTilesKernel(int* due, int tiles_x, int tiles_y, int tx, int ty) {
  const int dx = int(threadIdx.x) - 1, dy = int(threadIdx.y) - 1;
  const int x = tx + dx, y = ty + dy;
  if (x >= 0 && y >= 0 && x < tiles_x && y < tiles_y) due[y * tiles_x + x] = 0;
}

// masks |= filled; count the filled voxels (the seed included, image_analysis.cc:101-104)
template <int VEC>
__global__ void __launch_bounds__(256)
FillCommitKernel(const uint8_t* __restrict__ state, uint8_t* masks, size_t n,
                 unsigned long long* counter) {
  const size_t i = (size_t(blockIdx.x) * blockDim.x + threadIdx.x) * VEC;
  unsigned filled = 0;
  if (i < n) {
    if (VEC == 4) {
      const uchar4 s4 = *reinterpret_cast<const uchar4*>(state + i);
      const unsigned f = (s4.x & kFilled) | (s4.y & kFilled) << 8 | (s4.z & kFilled) << 16 |
                         (s4.w & kFilled) << 24;
      if (f) {
        uchar4 m4 = *reinterpret_cast<uchar4*>(masks + i);
        if (s4.x & kFilled) m4.x = 1;
        if (s4.y & kFilled) m4.y = 1;
        if (s4.z & kFilled) m4.z = 1;
        if (s4.w & kFilled) m4.w = 1;
        *reinterpret_cast<uchar4*>(masks + i) = m4;
        filled = __popc(f);
      }
    } else if (state[i] & kFilled) {
      masks[i] = 1;
      filled = 1;
    }
  }
  CountInto(counter, filled);
}

bool Aligned(const void* p, size_t a) { return reinterpret_cast<uintptr_t>(p) % a == 0; }

uint64_t ReadCounter(rdl_handle& h, cudaStream_t stream) {
  unsigned long long v = 0;
  RDL_CUDA(cudaMemcpyAsync(&v, h.iuwt_scalars.get(), sizeof(v), cudaMemcpyDeviceToHost, stream));
  RDL_CUDA(cudaStreamSynchronize(stream));
  return v;
}

void ZeroScalars(rdl_handle& h, cudaStream_t stream) {
  h.iuwt_scalars.Reserve(kIuwtScalarWords);
  RDL_CUDA(cudaMemsetAsync(h.iuwt_scalars.get(), 0, kIuwtScalarWords * sizeof(unsigned long long),
                           stream));
}

}  // namespace

// ---- significant structures ----------------------------------------------------------------

uint64_t IuwtSelectStructures(rdl_handle& h, const float* coefficients, uint8_t* masks,
                              size_t width, size_t height, int n_scales, const float* thresholds,
                              int min_scale, int end_scale, float clean_border,
                              const uint8_t* prior_mask, cudaStream_t stream) {
  const FillGeometry g =
      MakeGeometry(width, height, n_scales, thresholds, min_scale, end_scale, clean_border);
  ZeroScalars(h, stream);
  const size_t area = width * height;
  const bool vec = width % 4 == 0 && Aligned(coefficients, 16) && Aligned(masks, 4) &&
                   (!prior_mask || Aligned(prior_mask, 4));
  const dim3 grid(unsigned((area / (vec ? 4 : 1) + 255) / 256), unsigned(end_scale - min_scale));
  if (vec)
    StructureTestKernel<4, 0><<<grid, 256, 0, stream>>>(g, coefficients, masks, prior_mask,
                                                        nullptr, min_scale, h.iuwt_scalars.get());
  else
    StructureTestKernel<1, 0><<<grid, 256, 0, stream>>>(g, coefficients, masks, prior_mask,
                                                        nullptr, min_scale, h.iuwt_scalars.get());
  RDL_CUDA(cudaGetLastError());
  return ReadCounter(h, stream);
}

namespace {
// Shared by the 3-D and the 2-D fill: state volume, seed, propagation rounds, commit.
uint64_t RunFill(rdl_handle& h, const FillGeometry& g, const float* coefficients, uint8_t* masks,
                 const uint8_t* prior_mask, size_t seed_x, size_t seed_y, int seed_scale,
                 cudaStream_t stream) {
  if (seed_x >= g.width || seed_y >= g.height || seed_scale < 0 || seed_scale >= g.n_scales)
    throw std::invalid_argument("IUWT flood fill: seed outside the volume");
  const size_t area = g.width * g.height;
  // planes the walk can visit: down while scale > min_scale, up while scale < end_scale - 1
  const int s_lo = std::min(seed_scale, g.min_scale);
  const int s_hi = std::max(seed_scale, g.end_scale - 1);
  const int n_s = s_hi - s_lo + 1;
  ZeroScalars(h, stream);
  h.iuwt_state.Reserve(size_t(g.n_scales) * area);
  uint8_t* state = h.iuwt_state.get();
  const bool vec = g.width % 4 == 0 && Aligned(coefficients, 16) && Aligned(masks, 4) &&
                   (!prior_mask || Aligned(prior_mask, 4));
  const unsigned blocks = unsigned((area / (vec ? 4 : 1) + 255) / 256);
  const dim3 grid(blocks, unsigned(n_s));
  if (vec)
    StructureTestKernel<4, 1><<<grid, 256, 0, stream>>>(g, coefficients, masks, prior_mask, state,
                                                        s_lo, nullptr);
  else
    StructureTestKernel<1, 1><<<grid, 256, 0, stream>>>(g, coefficients, masks, prior_mask, state,
                                                        s_lo, nullptr);
  SeedKernel<<<1, 1, 0, stream>>>(state, size_t(seed_scale) * area + seed_y * g.width + seed_x);
  RDL_CUDA(cudaGetLastError());

  const size_t smem = size_t(n_s) * (kFillTX + 2) * (kFillTY + 2);
  if (smem > 48 * 1024)
    RDL_CUDA(cudaFuncSetAttribute(FillPropagateKernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                  int(smem)));
  const dim3 tiles(unsigned((g.width + kFillTX - 1) / kFillTX),
                   unsigned((g.height + kFillTY - 1) / kFillTY));
  // rounds are launched in batches; each round has its own change counter, and the first
  // round that changed nothing ends the fill
  unsigned* changed = reinterpret_cast<unsigned*>(h.iuwt_scalars.get() + kIuwtChangedWord);
  h.iuwt_due.Reserve(size_t(tiles.x) * tiles.y);
  int* due = h.iuwt_due.get();
  RDL_CUDA(cudaMemsetAsync(due, 0xff, size_t(tiles.x) * tiles.y * sizeof(int), stream));  // -1
  SeedTilesKernel<<<1, dim3(3, 3), 0, stream>>>(due, int(tiles.x), int(tiles.y),
                                                 int(seed_x / kFillTX), int(seed_y / kFillTY));
  for (int round = 0;;) {
    unsigned flags[kIuwtFillBatch];
    RDL_CUDA(cudaMemsetAsync(changed, 0, sizeof(flags), stream));
    for (int r = 0; r != kIuwtFillBatch; ++r, ++round)
      FillPropagateKernel<<<tiles, kFillThreads, smem, stream>>>(g, state, s_lo, s_hi, changed + r,
                                                                 due, round);
    RDL_CUDA(cudaGetLastError());
    RDL_CUDA(cudaMemcpyAsync(flags, changed, sizeof(flags), cudaMemcpyDeviceToHost, stream));
    RDL_CUDA(cudaStreamSynchronize(stream));
    if (std::find(flags, flags + kIuwtFillBatch, 0u) != flags + kIuwtFillBatch) break;
  }

  // only the planes the walk could visit hold state
  const size_t n = size_t(n_s) * area, offset = size_t(s_lo) * area;
  const bool vec_commit = n % 4 == 0 && Aligned(state + offset, 4) && Aligned(masks + offset, 4);
  if (vec_commit)
    FillCommitKernel<4><<<unsigned((n / 4 + 255) / 256), 256, 0, stream>>>(
        state + offset, masks + offset, n, h.iuwt_scalars.get());
  else
    FillCommitKernel<1><<<unsigned((n + 255) / 256), 256, 0, stream>>>(
        state + offset, masks + offset, n, h.iuwt_scalars.get());
  RDL_CUDA(cudaGetLastError());
  return ReadCounter(h, stream);
}
}  // namespace

uint64_t IuwtFloodfill(rdl_handle& h, const float* coefficients, uint8_t* masks, size_t width,
                       size_t height, int n_scales, const float* thresholds, int min_scale,
                       int end_scale, size_t seed_x, size_t seed_y, int seed_scale,
                       float clean_border, const uint8_t* prior_mask, cudaStream_t stream) {
  // image_analysis.cc:97: the fill never leaves the decomposition
  const FillGeometry g = MakeGeometry(width, height, n_scales, thresholds, min_scale,
                                      std::min(end_scale, n_scales), clean_border);
  return RunFill(h, g, coefficients, masks, prior_mask, seed_x, seed_y, seed_scale, stream);
}

uint64_t IuwtFloodfill2D(rdl_handle& h, const float* image, uint8_t* mask, size_t width,
                         size_t height, float threshold, bool use_abs, size_t seed_x,
                         size_t seed_y, cudaStream_t stream) {
  FillGeometry g = MakeGeometry(width, height, 1, &threshold, 0, 1, 0.0f);
  g.use_abs = use_abs ? 1 : 0;
  return RunFill(h, g, image, mask, nullptr, seed_x, seed_y, 0, stream);
}

// ---- median absolute deviation: exact radix select on the bit patterns of |v| ---------------
// Non-negative floats order like their bit patterns, so element n/2 of sorted |v| is found by
// three histogram passes over the 31 value bits (11 + 10 + 10), each narrowing the prefix.

namespace {

struct SelectPass {
  int shift, bits, prefix_shift;  // bin = (key >> shift) & ((1 << bits) - 1); prefix = key >> prefix_shift
};

__global__ void __launch_bounds__(512)
MadHistogramKernel(const float* __restrict__ image, size_t n, SelectPass pass,
                   const unsigned long long* __restrict__ select, unsigned* histogram) {
  __shared__ unsigned bins[2048];
  for (int i = threadIdx.x; i < 2048; i += blockDim.x) bins[i] = 0;
  __syncthreads();
  const unsigned prefix = unsigned(select[0]);
  const unsigned mask = (1u << pass.bits) - 1u;
  const size_t stride = size_t(gridDim.x) * blockDim.x;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const unsigned key = __float_as_uint(image[i]) & 0x7fffffffu;
    if (pass.prefix_shift < 31 && (key >> pass.prefix_shift) != prefix) continue;
    atomicAdd(&bins[(key >> pass.shift) & mask], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < 2048; i += blockDim.x)
    if (bins[i]) atomicAdd(&histogram[i], bins[i]);
}

// select[0] = prefix found so far, select[1] = rank still to be skipped inside that prefix
__global__ void MadNarrowKernel(unsigned long long* select, unsigned* histogram, int bits) {
  __shared__ unsigned long long chunk_sum[32];
  // 2048 bins over 32 lanes: each lane sums its 64-bin chunk, lane 0 walks the chunks
  const int lane = threadIdx.x;
  const int n_bins = 1 << bits, per = (n_bins + 31) / 32;
  unsigned long long sum = 0;
  for (int b = lane * per; b < min(n_bins, (lane + 1) * per); ++b) sum += histogram[b];
  chunk_sum[lane] = sum;
  __syncwarp();
  if (lane == 0) {
    unsigned long long rank = select[1];
    int c = 0;
    while (c < 31 && rank >= chunk_sum[c]) rank -= chunk_sum[c++];
    int b = c * per;
    while (b < n_bins - 1 && rank >= histogram[b]) rank -= histogram[b++];
    select[0] = (select[0] << bits) | unsigned(b);
    select[1] = rank;
  }
  __syncwarp();
  for (int b = lane; b < 2048; b += 32) histogram[b] = 0;
}

}  // namespace

float IuwtMad(rdl_handle& h, const float* image, size_t n, cudaStream_t stream) {
  if (n == 0 || n >= (size_t(1) << 32)) throw std::invalid_argument("IUWT Mad: bad plane size");
  ZeroScalars(h, stream);
  unsigned long long* select = h.iuwt_scalars.get() + 2;
  unsigned* histogram = reinterpret_cast<unsigned*>(h.iuwt_scalars.get() + kIuwtHistWord);
  const unsigned long long init[2] = {0ull, static_cast<unsigned long long>(n / 2)};
  RDL_CUDA(cudaMemcpyAsync(select, init, sizeof(init), cudaMemcpyHostToDevice, stream));
  const SelectPass passes[3] = {{20, 11, 31}, {10, 10, 20}, {0, 10, 10}};
  const unsigned grid = unsigned(std::min<size_t>((n + 511) / 512, size_t(kSmCount) * 4));
  for (const SelectPass& p : passes) {
    MadHistogramKernel<<<grid, 512, 0, stream>>>(image, n, p, select, histogram);
    MadNarrowKernel<<<1, 32, 0, stream>>>(select, histogram, p.bits);
  }
  RDL_CUDA(cudaGetLastError());
  unsigned long long found = 0;
  RDL_CUDA(cudaMemcpyAsync(&found, select, sizeof(found), cudaMemcpyDeviceToHost, stream));
  RDL_CUDA(cudaStreamSynchronize(stream));
  const unsigned bits = unsigned(found);
  float median;
  static_assert(sizeof(median) == sizeof(bits), "float is 32 bits");
  std::memcpy(&median, &bits, sizeof(median));
  return median / 0.674559;  // double division, as in the reference
}

// ---- largest coefficient of a plane ---------------------------------------------------------

namespace {

// larger float <=> larger unsigned
__device__ __forceinline__ unsigned OrderedBits(float v) {
  const unsigned u = __float_as_uint(v);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

__global__ void __launch_bounds__(256)
MaxAbsKernel(const float* __restrict__ image, const uint8_t* __restrict__ mask, size_t width,
             size_t min_x, size_t max_x, size_t min_y, size_t max_y, int allow_negative,
             unsigned long long* best) {
  const size_t span = max_x - min_x, n = span * (max_y - min_y);
  const size_t stride = size_t(gridDim.x) * blockDim.x;
  unsigned long long key = 0;
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n; i += stride) {
    const size_t y = min_y + i / span, x = min_x + i % span, px = y * width + x;
    if (mask && !mask[px]) continue;
    float v = image[px];
    if (allow_negative) v = fabsf(v);
    if (!(v > -FLT_MAX)) continue;  // strict '>' against lowest(); NaN never wins
    v += 0.0f;                      // -0 and +0 compare equal: the first of them wins
    const unsigned long long k =
        (static_cast<unsigned long long>(OrderedBits(v)) << 32) | (~unsigned(px));
    key = k > key ? k : key;
  }
  for (int o = 16; o; o >>= 1) {
    const unsigned long long other = __shfl_xor_sync(0xffffffffu, key, o);
    key = other > key ? other : key;
  }
  if ((threadIdx.x & 31) == 0 && key) atomicMax(best, key);
}

}  // namespace

float IuwtMaxAbs(rdl_handle& h, const float* image, size_t width, size_t height,
                 float clean_border, bool allow_negative, const uint8_t* mask, size_t& x,
                 size_t& y, cudaStream_t stream) {
  const float no_threshold = 0.0f;
  const FillGeometry g = MakeGeometry(width, height, 1, &no_threshold, 0, 1, clean_border);
  ZeroScalars(h, stream);
  unsigned long long* best = h.iuwt_scalars.get() + 4;
  const size_t n = (g.max_x - g.min_x) * (g.max_y - g.min_y);
  const unsigned grid = unsigned(std::min<size_t>((n + 255) / 256, size_t(kSmCount) * 8));
  MaxAbsKernel<<<grid, 256, 0, stream>>>(image, mask, width, g.min_x, g.max_x, g.min_y, g.max_y,
                                         allow_negative ? 1 : 0, best);
  RDL_CUDA(cudaGetLastError());
  unsigned long long key = 0;
  RDL_CUDA(cudaMemcpyAsync(&key, best, sizeof(key), cudaMemcpyDeviceToHost, stream));
  RDL_CUDA(cudaStreamSynchronize(stream));
  x = width;
  y = height;
  if (key == 0) return std::numeric_limits<float>::lowest();
  const size_t px = ~unsigned(key & 0xffffffffu);
  x = px % width;
  y = px / width;
  float v;
  RDL_CUDA(cudaMemcpyAsync(&v, image + px, sizeof(v), cudaMemcpyDeviceToHost, stream));
  RDL_CUDA(cudaStreamSynchronize(stream));
  return allow_negative ? std::fabs(v) : v;
}

}  // namespace rdl
