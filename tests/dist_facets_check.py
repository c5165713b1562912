"""torchrun script: facet-sharded ParallelDeconvolution (facet i on rank i % WORLD_SIZE, results
merged over NCCL) vs the same facet run on ONE GPU inside the same process: per-call result,
facet records, residual and model must be BIT-identical on every rank (GenericClean facets do not
depend on the order in which neighbours finish). Two major iterations per case; device and host
buffers. Prints "FACETS-OK <world>" on rank 0.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/dist_facets_check.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
import torch
import torch.distributed as dist

import ska_sdp_func_radler_b200 as rdl
from helpers import hogbom_case
from ska_sdp_func_radler_b200 import distributed as rd

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
ok = True
SETTINGS = dict(max_iterations=400, minor_loop_gain=0.1, major_loop_gain=0.7, clean_border_ratio=0.0)


def run(alg, n, c, grid, res, psfs, weights, host):
    cube = rdl.Cube(n, n, c, 1, weights)
    if host:
        r, m, p = res.copy(), np.zeros_like(res), psfs
    else:
        r, m, p = (torch.from_numpy(res.copy()).to(dev), torch.zeros((c, n, n), device=dev),
                   torch.from_numpy(psfs).to(dev))
    results = [alg.parallel_execute_major_iteration(grid[0], grid[1], cube, r, m, p) for _ in range(2)]
    if not host:
        r, m = r.cpu().numpy(), m.cpu().numpy()
    return results, alg.facets(), r, m


for (n, c, grid, sub_minor, host) in [(128, 1, (2, 2), False, False), (192, 2, (3, 2), False, False),
                                      (160, 1, (3, 3), False, True), (128, 1, (2, 2), True, False),
                                      (256, 1, (1, 1), False, False)]:
    res, psfs, weights, _ = hogbom_case(n, c)
    single = rdl.GenericClean(use_sub_minor_optimization=sub_minor, device=local)
    single.apply(**SETTINGS)
    want = run(single, n, c, grid, res, psfs, weights, host)
    alg = rdl.GenericClean(use_sub_minor_optimization=sub_minor, device=local)
    alg.apply(**SETTINGS)
    uid = rd.share_unique_id(rdl.dist_unique_id, rank, world, device=dev)
    alg.dist_init(uid, rank, world)
    got = run(alg, n, c, grid, res, psfs, weights, host)
    same_results = all(
        (g.starting_peak_value, g.final_peak_value, g.another_iteration_required, g.is_diverging,
         g.iteration_number) ==
        (w.starting_peak_value, w.final_peak_value, w.another_iteration_required, w.is_diverging,
         w.iteration_number) for g, w in zip(got[0], want[0]))
    good = (same_results and got[1] == want[1] and np.array_equal(got[2], want[2]) and
            np.array_equal(got[3], want[3]) and
            sum(f["iteration_number"] for f in want[1]) > 0)
    if not good:
        print(f"rank {rank}: MISMATCH n={n} c={c} grid={grid} sub_minor={sub_minor} host={host}: "
              f"{got[0]} vs {want[0]}; residual diff "
              f"{np.abs(got[2] - want[2]).max()} model diff {np.abs(got[3] - want[3]).max()}", flush=True)
    ok = ok and good
    alg.dist_finalize()
flag = torch.tensor([1 if ok else 0], device=dev)
if world > 1:
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
if rank == 0:
    print(f"FACETS-OK {world}" if int(flag) else "FACETS-FAIL", flush=True)
sys.exit(0 if int(flag) else 1)
