// NCCL bindings resolved with dlopen (see dist.h).
#include "dist.h"

#include <dlfcn.h>
#include <nccl.h>

#include <mutex>
#include <string>

#include "common.cuh"

namespace rdl {
namespace {

struct NcclApi {
  ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
  ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t,
                            cudaStream_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t,
                            cudaStream_t) = nullptr;
  ncclResult_t (*Broadcast)(const void*, void*, size_t, ncclDataType_t, int, ncclComm_t,
                            cudaStream_t) = nullptr;
  ncclResult_t (*Send)(const void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*Recv)(void*, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t) = nullptr;
  ncclResult_t (*GroupStart)() = nullptr;
  ncclResult_t (*GroupEnd)() = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
};

const NcclApi& Api() {
  static NcclApi api;
  static std::once_flag once;
  static std::string error;
  std::call_once(once, [] {
    // a process that imported torch already has its bundled libnccl.so.2 loaded: same SONAME
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) {
      error = std::string("cannot load libnccl: ") + dlerror();
      return;
    }
    auto sym = [&](const char* name) {
      void* p = dlsym(lib, name);
      if (!p && error.empty()) error = std::string("libnccl lacks ") + name;
      return p;
    };
    api.GetUniqueId = reinterpret_cast<decltype(api.GetUniqueId)>(sym("ncclGetUniqueId"));
    api.CommInitRank = reinterpret_cast<decltype(api.CommInitRank)>(sym("ncclCommInitRank"));
    api.CommDestroy = reinterpret_cast<decltype(api.CommDestroy)>(sym("ncclCommDestroy"));
    api.AllGather = reinterpret_cast<decltype(api.AllGather)>(sym("ncclAllGather"));
    api.AllReduce = reinterpret_cast<decltype(api.AllReduce)>(sym("ncclAllReduce"));
    api.Broadcast = reinterpret_cast<decltype(api.Broadcast)>(sym("ncclBroadcast"));
    api.Send = reinterpret_cast<decltype(api.Send)>(sym("ncclSend"));
    api.Recv = reinterpret_cast<decltype(api.Recv)>(sym("ncclRecv"));
    api.GroupStart = reinterpret_cast<decltype(api.GroupStart)>(sym("ncclGroupStart"));
    api.GroupEnd = reinterpret_cast<decltype(api.GroupEnd)>(sym("ncclGroupEnd"));
    api.GetErrorString = reinterpret_cast<decltype(api.GetErrorString)>(sym("ncclGetErrorString"));
  });
  if (!error.empty()) throw CudaError(error);
  return api;
}

void Check(ncclResult_t r, const char* what) {
  if (r != ncclSuccess) throw CudaError(std::string(what) + ": " + Api().GetErrorString(r));
}

static_assert(sizeof(ncclUniqueId) == kUniqueIdBytes, "ncclUniqueId size");

}  // namespace

void NcclGetUniqueId(uint8_t out[kUniqueIdBytes]) {
  ncclUniqueId id;
  Check(Api().GetUniqueId(&id), "ncclGetUniqueId");
  memcpy(out, &id, kUniqueIdBytes);
}

void NcclInit(DistContext& ctx, const uint8_t id_bytes[kUniqueIdBytes], int rank, int world) {
  if (world < 1 || rank < 0 || rank >= world) throw std::invalid_argument("bad rank / world size");
  NcclDestroy(ctx);
  ctx.rank = rank;
  ctx.world = world;
  if (world == 1) return;
  ncclUniqueId id;
  memcpy(&id, id_bytes, kUniqueIdBytes);
  ncclComm_t comm = nullptr;
  Check(Api().CommInitRank(&comm, world, id, rank), "ncclCommInitRank");
  ctx.comm = comm;
}

void NcclDestroy(DistContext& ctx) {
  if (ctx.comm) Api().CommDestroy(static_cast<ncclComm_t>(ctx.comm));
  ctx.comm = nullptr;
}

void NcclAllGatherBytes(DistContext& ctx, const void* send, void* recv, size_t bytes_per_rank,
                        cudaStream_t stream) {
  if (!ctx.comm) throw std::invalid_argument("rdl_dist_init was not called on this handle");
  Check(Api().AllGather(send, recv, bytes_per_rank, ncclChar, static_cast<ncclComm_t>(ctx.comm),
                        stream),
        "ncclAllGather");
}

void NcclBroadcastPlanes(DistContext& ctx, float* const* planes, const int* roots, size_t n_planes,
                         size_t count, cudaStream_t stream) {
  if (!ctx.comm) throw std::invalid_argument("rdl_dist_init was not called on this handle");
  if (n_planes == 0) return;
  ncclComm_t comm = static_cast<ncclComm_t>(ctx.comm);
  Check(Api().GroupStart(), "ncclGroupStart");
  for (size_t i = 0; i != n_planes; ++i)
    Check(Api().Broadcast(planes[i], planes[i], count, ncclFloat, roots[i], comm, stream),
          "ncclBroadcast");
  Check(Api().GroupEnd(), "ncclGroupEnd");
}

void NcclSendFloat(DistContext& ctx, const float* data, size_t count, int peer, cudaStream_t stream) {
  if (!ctx.comm) throw std::invalid_argument("rdl_dist_init was not called on this handle");
  Check(Api().Send(data, count, ncclFloat, peer, static_cast<ncclComm_t>(ctx.comm), stream), "ncclSend");
}

void NcclRecvFloat(DistContext& ctx, float* data, size_t count, int peer, cudaStream_t stream) {
  if (!ctx.comm) throw std::invalid_argument("rdl_dist_init was not called on this handle");
  Check(Api().Recv(data, count, ncclFloat, peer, static_cast<ncclComm_t>(ctx.comm), stream), "ncclRecv");
}

void NcclAllReduceSumFloat(DistContext& ctx, float* data, size_t count, cudaStream_t stream) {
  if (!ctx.comm) throw std::invalid_argument("rdl_dist_init was not called on this handle");
  Check(Api().AllReduce(data, data, count, ncclFloat, ncclSum, static_cast<ncclComm_t>(ctx.comm),
                        stream),
        "ncclAllReduce");
}

}  // namespace rdl
