// rdl_iuwt_decompose_static: IuwtDecomposition::DecomposeMt for callers that may not allocate
// device memory (CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
// No counterpart in the reference.
//
// rdl_iuwt_decompose keeps its two ping-pong planes in DeviceBuffers of the handle (cudaMalloc on
// first use / growth). Here they are the two halves of ONE file-scope device array of
// 2 * kStaticMaxPixels floats (8 MiB), which every GPU that loads the library holds until exit.
// The entry point calls no cudaMalloc / cudaFree and no allocating library routine; images with
// more than kStaticMaxPixels pixels are rejected.
//
// Every call on a device shares that array: it keeps the previous call's contents (nothing in
// this algorithm needs it zeroed -- each plane is fully overwritten by a smoothing pass before it
// is read), and two calls that overlap on the GPU (different streams, different handles, other
// host threads) race on it. Serialise them.
#include <algorithm>
#include <mutex>
#include <stdexcept>
#include <string>

#include "../../include/radler_b200.h"
#include "common.cuh"
#include "handle.h"
#include "iuwt.cuh"

namespace rdl {
namespace {

constexpr size_t kStaticMaxPixels = size_t(1) << 20;  // 1024 x 1024, or any shape with as many pixels
constexpr int kStaticMaxDevices = 64;

// In host code this name is only a stand-in for the symbol, not a device address: the address a
// kernel can use comes from cudaGetSymbolAddress, per device (ScratchOfCurrentDevice).
static __device__ __align__(256) float g_static_scratch[2 * kStaticMaxPixels];

std::mutex g_scratch_mutex;
float* g_scratch_address[kStaticMaxDevices] = {};

// The array's address on `device`, which must be the current device. Looked up once per device.
float* ScratchOfCurrentDevice(int device) {
  if (device < 0 || device >= kStaticMaxDevices)
    throw std::invalid_argument("device index beyond the static scratch table");
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (!g_scratch_address[device]) {
    float* p = nullptr;
    RDL_CUDA(cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_static_scratch));
    if (!p) throw std::runtime_error("cudaGetSymbolAddress returned no address");
    g_scratch_address[device] = p;
  }
  return g_scratch_address[device];
}

// w = i0 - w  (differenceMT with the doubly smoothed image already in the coefficient plane)
__global__ void DifferenceInPlaceKernel(float* __restrict__ w, const float* __restrict__ i0, size_t n) {
  for (size_t i = size_t(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
       i += size_t(gridDim.x) * blockDim.x)
    w[i] = __fsub_rn(i0[i], w[i]);
}

}  // namespace
}  // namespace rdl

extern "C" rdl_status rdl_iuwt_decompose_static(rdl_handle* h, const float* device_input,
                                                uint64_t width, uint64_t height, int32_t n_scales,
                                                int32_t include_largest, float* device_coefficients,
                                                void* stream_v) {
  using namespace rdl;
  if (!h || !device_input || !device_coefficients) return RDL_ERR_INVALID_ARGUMENT;
  try {
    if (n_scales < 1) throw std::invalid_argument("IUWT needs at least one scale");
    if (width == 0 || height == 0 || width > kStaticMaxPixels || height > kStaticMaxPixels ||
        width * height > kStaticMaxPixels)
      throw std::invalid_argument("rdl_iuwt_decompose_static: " + std::to_string(width) + " x " +
                                  std::to_string(height) + " exceeds the static scratch of " +
                                  std::to_string(kStaticMaxPixels) + " pixels");
    RDL_CUDA(cudaSetDevice(h->device));
    cudaStream_t stream = stream_v ? static_cast<cudaStream_t>(stream_v) : h->own_stream;
    float* scratch = ScratchOfCurrentDevice(h->device);
    const size_t a = size_t(width) * height;
    float* planes[2] = {scratch, scratch + kStaticMaxPixels};
    const int grid = int(std::min<size_t>((a + 255) / 256, size_t(kSmCount) * 16));
    const float* i0 = device_input;
    int next = 0;
    for (int s = 0; s != n_scales; ++s) {
      const bool last = s + 1 == n_scales;
      float* i1 = (last && include_largest) ? device_coefficients + size_t(n_scales) * a : planes[next];
      float* w = device_coefficients + size_t(s) * a;
      IuwtSmooth(*h, i0, i1, width, height, s + 1, false, stream);  // i1 = B(i0)
      IuwtSmooth(*h, i1, w, width, height, s + 1, false, stream);   // w = B(i1)
      DifferenceInPlaceKernel<<<grid, 256, 0, stream>>>(w, i0, a);  // w = i0 - B(i1)
      RDL_CUDA(cudaGetLastError());
      ++h->launches;
      i0 = i1;
      next ^= 1;
    }
    return RDL_OK;
  } catch (const CudaError& e) {
    h->error = e.what();
    cudaGetLastError();
    return RDL_ERR_CUDA;
  } catch (const std::invalid_argument& e) {
    h->error = e.what();
    return RDL_ERR_INVALID_ARGUMENT;
  } catch (const std::exception& e) {
    h->error = e.what();
    return RDL_ERR_INVALID_ARGUMENT;
  }
}
