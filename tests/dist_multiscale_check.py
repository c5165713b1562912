"""torchrun script: channel-sharded MultiScaleAlgorithm (csrc/multiscale.cu, "Several GPUs") on
WORLD_SIZE GPUs vs the same call on one GPU without a distributed context: iteration count, peak
trace, per-scale counters, residual and model cubes must be BIT-identical on every rank (the
ranks run the same kernels on the same data; only WHO convolves a channel changes). Also checks the
single-GPU result against the CPU oracle's peak trace. Prints "MS-DIST-OK <world>" on rank 0.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/dist_multiscale_check.py
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
from helpers import multiscale_case
from ska_sdp_func_radler_b200 import distributed as rd

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
ok = True


def run(alg, cube, res, psfs, calls, automask):
    d_res = torch.from_numpy(res.copy()).to(dev)
    d_model = torch.zeros_like(d_res)
    d_psf = torch.from_numpy(psfs).to(dev)
    out = []
    for k in range(calls):
        if automask:
            alg.set_auto_mask_mode(k == 0, k > 0)
            alg.set_threshold(0.3 if k == 0 else 0.1)
        g = alg.execute_major_iteration(cube, d_res, d_model, d_psf)
        out.append((g.iteration_number, g.final_peak_value, g.another_iteration_required,
                    alg.trace().copy(), [tuple(sorted(s.items())) for s in alg.scale_infos()]))
    torch.cuda.synchronize()
    return out, d_res.cpu().numpy(), d_model.cpu().numpy()


# (size, channels, max iterations, fit, calls, automask, track components)
for (n, c, iters, fit, calls, automask, track) in [(256, 4, 800, True, 1, False, False),
                                                   (192, 3, 500, False, 2, False, True),
                                                   (256, 5, 600, True, 2, True, False),
                                                   (128, 2, 300, False, 1, False, False)]:
    res, psfs, weights, freqs = multiscale_case(n, c)
    cube = rdl.Cube(n, n, c, 1, weights)

    def make():
        a = rdl.MultiScaleAlgorithm(6.0, 1.0, max_scales=4, device=local, track_components=track)
        a.apply(max_iterations=iters, clean_border_ratio=0.0, major_loop_gain=0.8,
                trace_capacity=1 << 16)
        if fit:
            a.set_spectral_fitter(rdl.FIT_POLYNOMIAL, 2, freqs, weights, 1)
        return a

    single = make()
    want, w_res, w_model = run(single, cube, res, psfs, calls, automask)
    sharded = make()
    uid = rd.share_unique_id(rdl.dist_unique_id, rank, world, device=dev)
    sharded.dist_init(uid, rank, world)
    got, g_res, g_model = run(sharded, cube, res, psfs, calls, automask)
    good = np.array_equal(g_res, w_res) and np.array_equal(g_model, w_model) and want[0][0] > 10
    for (wi, wp, wa, wt, ws), (gi, gp, ga, gt, gs) in zip(want, got):
        good = good and wi == gi and wp == gp and wa == ga and np.array_equal(wt, gt) and ws == gs
    if track:
        a, b = single.component_list(c), sharded.component_list(c)
        good = good and len(a) == len(b) and all(
            np.array_equal(x0, x1) and np.array_equal(y0, y1) and np.array_equal(v0, v1)
            for (x0, y0, v0), (x1, y1, v1) in zip(a, b))
    if not good:
        print(f"rank {rank}: MISMATCH n={n} c={c} iterations {want[0][0]} vs {got[0][0]}", flush=True)
    ok = ok and good
    sharded.dist_finalize()

flag = torch.tensor([1 if ok else 0], device=dev)
if world > 1:
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0 and int(flag[0]) == 1:
    print(f"MS-DIST-OK {world}")
if world > 1:
    dist.destroy_process_group()
