"""How fast one 8192^2 float plane travels between ranks: ncclBroadcast (what the channel-sharded
multiscale uses) vs send/recv, torch.distributed over NCCL. torchrun --nproc-per-node N ..."""
import json
import os

import torch
import torch.distributed as dist

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
local = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n = 8192 * 8192
planes = [torch.randn(n, device="cuda") for _ in range(8)]


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def bcast():
    for i, p in enumerate(planes):
        dist.broadcast(p, src=i % world)


def p2p():
    ops = []
    for i, p in enumerate(planes):
        src = i % world
        if rank == src:
            ops += [dist.P2POp(dist.isend, p, peer) for peer in range(world) if peer != src]
        else:
            ops.append(dist.P2POp(dist.irecv, p, src))
    for w in dist.batch_isend_irecv(ops):
        w.wait()


out = {"world": world, "plane_MB": n * 4 / 1e6}
for name, fn in (("broadcast", bcast), ("send_recv", p2p)):
    ms = timed(fn)
    # bytes RECEIVED per rank per round: planes it does not own
    recv = sum(1 for i in range(len(planes)) if i % world != rank) * n * 4
    out[name] = {"ms_per_8_planes": ms, "receive_GBps_per_rank": recv / ms / 1e6}
if rank == 0:
    print(json.dumps(out))
dist.destroy_process_group()
