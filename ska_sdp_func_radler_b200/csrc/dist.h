// Multi-GPU plumbing: NCCL (over NVLink / NVSwitch) is loaded at run time, so single-GPU users
// carry no NCCL dependency. One process per GPU.
//   row-sharded Hogbom loop: the only exchange is the joined arg-max, an all-gather of one
//     DistCandidate (264 bytes) per rank per minor iteration;
//   facet-sharded ParallelDeconvolution: an all-gather of the facet results after each of the two
//     passes (24 bytes per facet) and one in-place sum of the residual / model cubes at the end,
//     in which every pixel has exactly one non-zero contributor;
//   channel-sharded MultiScaleAlgorithm: every rank runs the FFT convolutions of its own channels
//     and broadcasts their results; everything joined (integration over channels, scale selection,
//     the sub-minor loop on the compact sets) is computed identically on every rank.
#ifndef RDL_DIST_H_
#define RDL_DIST_H_

#include <cstddef>
#include <cstdint>

#include <cuda_runtime.h>

namespace rdl {

constexpr size_t kUniqueIdBytes = 128;  // sizeof(ncclUniqueId)

struct DistContext {
  void* comm = nullptr;  // ncclComm_t
  int rank = 0, world = 1;
  bool band_set = false;
  size_t y_offset = 0, global_height = 0;
  size_t block_rows = 0;  // > 0: cyclic row blocks (rank owns blocks b with b % world == rank)
};

// Throws CudaError when libnccl cannot be loaded or a call fails.
void NcclGetUniqueId(uint8_t out[kUniqueIdBytes]);
void NcclInit(DistContext& ctx, const uint8_t id[kUniqueIdBytes], int rank, int world);
void NcclDestroy(DistContext& ctx);
void NcclAllGatherBytes(DistContext& ctx, const void* send, void* recv, size_t bytes_per_rank,
                        cudaStream_t stream);

// planes[i] (count floats, the same address range on every rank's own copy) is sent from rank
// roots[i] to everybody, all planes in one NCCL group.
void NcclBroadcastPlanes(DistContext& ctx, float* const* planes, const int* roots, size_t n_planes,
                         size_t count, cudaStream_t stream);
void NcclSendFloat(DistContext& ctx, const float* data, size_t count, int peer, cudaStream_t stream);
void NcclRecvFloat(DistContext& ctx, float* data, size_t count, int peer, cudaStream_t stream);

// In-place sum over the ranks.
void NcclAllReduceSumFloat(DistContext& ctx, float* data, size_t count, cudaStream_t stream);

}  // namespace rdl
#endif
