"""K5: ParallelDeconvolution 16384^2 with 8x8 facets, Hogbom per facet, wall clock + the host
share (Dijkstra dividers). python tools/facet_bench.py [SIZE] [GRID] [MAX_ITERATIONS]
Under torchrun (WORLD_SIZE > 1) facet i runs on rank i % WORLD_SIZE and the cubes are merged over
NCCL; every rank holds the whole frame, the time printed is the maximum over the ranks."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import ska_sdp_func_radler_b200 as rdl
from bench import make_cube

n = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
g = int(sys.argv[2]) if len(sys.argv) > 2 else 8
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 2000
rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
if world > 1:
    import torch.distributed as dist
    from ska_sdp_func_radler_b200 import distributed as rd
    dist.init_process_group("nccl", device_id=dev)
res0, psfs, weights, _ = make_cube(n, 1, dev)


def make_algorithm():
    a = rdl.GenericClean(use_sub_minor_optimization=False, device=local)
    a.apply(max_iterations=iters, clean_border_ratio=0.0, minor_loop_gain=0.1, major_loop_gain=0.8,
            threshold=0.0)
    if world > 1:
        a.dist_init(rd.share_unique_id(rdl.dist_unique_id, rank, world, device=dev), rank, world)
    return a


alg = make_algorithm()
cube = rdl.Cube(n, n, 1, 1, None)
for rep in range(2):
    res, model = res0.clone(), torch.zeros_like(res0)
    alg.set_iteration_number(0)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    l0 = alg.kernel_launch_count
    t0 = time.perf_counter()
    r = alg.parallel_execute_major_iteration(g, g, cube, res, model, psfs)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([wall], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        wall = float(t)
    facets = alg.facets()
    geom = alg.parallel_timing()[0] * 1e-3  # this rank's divider search inside the call
    if rank == 0:
      print(json.dumps({"n": n, "grid": g, "n_gpus": world, "rep": rep, "wall_ms": wall * 1e3,
                      "residual_checksum": float(res.double().abs().sum()),
                      "model_checksum": float(model.double().sum()),
                      "host_geometry_ms": geom * 1e3, "device_and_copies_ms": (wall - geom) * 1e3,
                      "iterations_facet0": r.iteration_number,
                      "iterations_all_facets": int(sum(f["iteration_number"] for f in facets)),
                      "another_iteration_required": r.another_iteration_required,
                      "launches": alg.kernel_launch_count - l0,
                      "facet_boxes": [(f["width"], f["height"]) for f in facets[:4]]}), flush=True)
    if world > 1:
        alg.dist_finalize()
    alg = make_algorithm()
if world > 1:
    dist.destroy_process_group()
