// Size-keyed cache of device allocations behind DeviceBuffer (handle.h).
// A multiscale major iteration takes ~150 buffers of a few hundred MB (convolved PSFs, spectra,
// per-image scratch) and gives them back at the end of the call; cudaMalloc / cudaFree of that
// much memory cost a variable 0.1-0.4 s of host time per call. Freed blocks are kept per
// (device, size) instead and handed out again; the sizes in play are few (frame planes,
// spectra). A block is given back only after the device is idle (cudaFree synchronises too), so
// reuse from another handle's stream is safe.
#include <cstdlib>
#include <map>
#include <mutex>
#include <string>
#include <vector>

#include "common.cuh"

namespace rdl {
namespace {

struct Pool {
  std::mutex lock;
  std::map<std::pair<int, size_t>, std::vector<void*>> blocks;
  size_t cached_bytes = 0;
  // Three quarters of the device's memory unless RDL_POOL_MAX_GB says otherwise (135 GB on a
  // B200). A K4 multiscale call holds ~110 GB of planes and spectra; with the earlier fixed cap of
  // 96 GB the surplus went back to the driver at the end of every call and was allocated again in
  // the next, at 0.4 s for some calls and nothing for others (which blocks stayed was a matter
  // of release order): major iterations of 390 and 800-900 ms alternated in one process.
  size_t max_cached_bytes = size_t(96) << 30;
  Pool() {
    size_t free_bytes = 0, total_bytes = 0;
    if (cudaMemGetInfo(&free_bytes, &total_bytes) == cudaSuccess && total_bytes > 0)
      max_cached_bytes = total_bytes / 4 * 3;
    else
      cudaGetLastError();
    if (const char* gb = std::getenv("RDL_POOL_MAX_GB")) max_cached_bytes = size_t(std::atof(gb) * double(1 << 30));
  }
};

Pool& ThePool() {
  static Pool* pool = new Pool;  // never destroyed: the driver may be gone at exit
  return *pool;
}

constexpr size_t kPooledFrom = size_t(1) << 20;  // smaller blocks go straight to the driver

int CurrentDevice() {
  int device = 0;
  cudaGetDevice(&device);
  return device;
}

}  // namespace

void DeviceTrim() {
  Pool& pool = ThePool();
  std::lock_guard<std::mutex> guard(pool.lock);
  for (auto& entry : pool.blocks)
    for (void* p : entry.second) cudaFree(p);
  pool.blocks.clear();
  pool.cached_bytes = 0;
}

size_t DeviceAvailableBytes() {
  size_t free_bytes = 0, total_bytes = 0;
  if (cudaMemGetInfo(&free_bytes, &total_bytes) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  Pool& pool = ThePool();
  std::lock_guard<std::mutex> guard(pool.lock);
  const int device = CurrentDevice();
  for (const auto& entry : pool.blocks)
    if (entry.first.first == device) free_bytes += entry.first.second * entry.second.size();
  return free_bytes;
}

void* DeviceAlloc(size_t bytes) {
  Pool& pool = ThePool();
  if (bytes >= kPooledFrom) {
    std::lock_guard<std::mutex> guard(pool.lock);
    auto it = pool.blocks.find({CurrentDevice(), bytes});
    if (it != pool.blocks.end() && !it->second.empty()) {
      void* p = it->second.back();
      it->second.pop_back();
      pool.cached_bytes -= bytes;
      return p;
    }
  }
  void* p = nullptr;
  cudaError_t e = cudaMalloc(&p, bytes);
  if (e != cudaSuccess) {  // give the cache back to the driver and try once more
    cudaGetLastError();
    DeviceTrim();
    e = cudaMalloc(&p, bytes);
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    throw CudaError(std::string("cudaMalloc(") + std::to_string(bytes) + " bytes): " +
                    cudaGetErrorString(e));
  }
  return p;
}

void DeviceFree(void* p, size_t bytes) {
  if (!p) return;
  Pool& pool = ThePool();
  if (bytes >= kPooledFrom) {
    // the block's own device, which need not be the calling thread's current one
    cudaPointerAttributes attr{};
    const int current = CurrentDevice();
    int device = current;
    if (cudaPointerGetAttributes(&attr, p) == cudaSuccess) device = attr.device;
    else cudaGetLastError();
    if (device != current) cudaSetDevice(device);
    cudaDeviceSynchronize();  // what cudaFree would have done: nothing in flight uses the block
    if (device != current) cudaSetDevice(current);
    std::lock_guard<std::mutex> guard(pool.lock);
    if (pool.cached_bytes + bytes <= pool.max_cached_bytes) {
      pool.blocks[{device, bytes}].push_back(p);
      pool.cached_bytes += bytes;
      return;
    }
  }
  cudaFree(p);
}

}  // namespace rdl
