"""torchrun script: row-sharded GenericClean (NCCL joined arg-max) on WORLD_SIZE GPUs vs the CPU
oracle on the same cube: peak sequence, iteration count and every rank's band of the residual /
model must be BIT-identical to the single-process result. Prints "DIST-OK <world>" on rank 0.
    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 tests/dist_hogbom_check.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)
import torch
import torch.distributed as dist

import oracle
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
# (size, channels, iterations, spectral fit, cyclic block rows or 0 = contiguous bands)
for (n, c, iters, fit, block) in [(256, 1, 150, False, 0), (192, 3, 120, False, 0),
                                  (256, 4, 100, True, 0), (256, 2, 150, False, 16),
                                  (192, 3, 100, True, 4)]:
    res, psfs, weights, freqs = hogbom_case(n, c)
    st = oracle.Settings(max_iterations=iters, clean_border_ratio=0.02, record_trace=True)
    if fit:
        st.spectral_mode, st.n_terms, st.frequencies, st.fit_weights = 1, 2, freqs, weights
    o_res, o_model = res.copy(), np.zeros_like(res)
    o = oracle.GenericClean(False).execute(oracle.Cube(n, n, c, 1, weights), st, o_res, o_model, psfs)
    alg = rdl.GenericClean(use_sub_minor_optimization=False, device=local)
    alg.apply(max_iterations=iters, clean_border_ratio=0.02, trace_capacity=iters)
    if fit:
        alg.set_spectral_fitter(rdl.FIT_POLYNOMIAL, 2, freqs, weights, 1)
    uid = rd.share_unique_id(rdl.dist_unique_id, rank, world, device=dev)
    alg.dist_init(uid, rank, world)
    if block:
        rows = rd.cyclic_rows(n, rank, world, block)
        alg.dist_set_row_blocks(block, n)
    else:
        y0, y1 = rd.row_band(n, rank, world)
        rows = np.arange(y0, y1)
        alg.dist_set_row_band(y0, n)
    d_res = torch.from_numpy(np.ascontiguousarray(res[:, rows])).to(dev)
    d_model = torch.zeros_like(d_res)
    g = alg.execute_major_iteration(rdl.Cube(n, len(rows), c, 1, weights), d_res, d_model,
                                    torch.from_numpy(psfs).to(dev))
    tr = alg.trace()
    good = (g.iteration_number == o.iteration_number and
            np.array_equal(tr["x"], o.trace["x"]) and np.array_equal(tr["y"], o.trace["y"]) and
            np.array_equal(tr["value"], o.trace["value"]) and
            np.array_equal(d_res.cpu().numpy(), o_res[:, rows]) and
            np.array_equal(d_model.cpu().numpy(), o_model[:, rows]) and
            g.final_peak_value == o.final_peak_value)
    if not good:
        print(f"rank {rank}: MISMATCH case n={n} c={c} fit={fit} block={block}: iterations "
              f"{g.iteration_number} vs {o.iteration_number}", flush=True)
    ok = ok and good
    # the same run through HOST buffers: every rank passes its rows and the full PSFs; each PSF
    # plane is uploaded by one rank and broadcast (abi.cu). Two calls, the second with the PSFs
    # kept on the device.
    h_res = np.ascontiguousarray(res[:, rows])
    h_model = np.zeros_like(h_res)
    alg.set_iteration_number(0)
    alg.set_psfs_constant(True)
    band_cube = rdl.Cube(n, len(rows), c, 1, weights)
    alg.apply(max_iterations=iters // 2, clean_border_ratio=0.02, trace_capacity=iters)
    alg.execute_major_iteration(band_cube, h_res, h_model, psfs)
    alg.apply(max_iterations=iters, clean_border_ratio=0.02, trace_capacity=iters)
    g2 = alg.execute_major_iteration(band_cube, h_res, h_model, psfs)
    good = (g2.iteration_number == o.iteration_number and np.array_equal(h_res, o_res[:, rows]) and
            np.array_equal(h_model, o_model[:, rows]))
    if not good:
        print(f"rank {rank}: HOST-BUFFER MISMATCH case n={n} c={c} fit={fit} block={block}", flush=True)
    ok = ok and good
    alg.dist_finalize()
flag = torch.tensor([1 if ok else 0], device=dev)
if world > 1:
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    dist.destroy_process_group()
if rank == 0:
    print(f"DIST-OK {world}" if int(flag) else "DIST-FAIL", flush=True)
sys.exit(0 if int(flag) else 1)
