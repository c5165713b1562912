"""CPU figures from the reference's own sources (oracle/_ref, all host cores) beside the device for
BASELINE configs 3 and 5, on bounded versions of the configs, wall clock:
  K3  IuwtDeconvolution::ExecuteMajorIteration, 2048^2 x 1
  K5  ParallelDeconvolution 8 x 8 facets, Hogbom per facet (divider search included), 4096^2 x 1
python tests/tools/k3_k5_cpu_bench.py [k3_size] [k5_size]"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch

import oracle
import ska_sdp_func_radler_b200 as rdl
from bench import make_cube, make_ms_cube

oracle.build()
cores = os.cpu_count() or 1
k3_n = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
k5_n = int(sys.argv[2]) if len(sys.argv) > 2 else 4096


def rel(a, b):
    return float(np.abs(a.astype(np.float64) - b).max() / max(np.abs(b).max(), 1e-30))


# ---- K3 -------------------------------------------------------------------------------------
res, psfs, _, _ = make_ms_cube(k3_n, 1, "cpu")
res, psfs = res.numpy(), psfs.numpy()
kw = dict(max_iterations=20, minor_loop_gain=0.1, major_loop_gain=0.8, clean_border_ratio=0.0)
ref = oracle.Reference("iuwt", k3_n, k3_n, 1, **kw)
r_res, r_model = res.copy(), np.zeros_like(res)
t0 = time.perf_counter()
w = ref.execute(r_res, r_model, psfs)
cpu_s = time.perf_counter() - t0
alg = rdl.DeconvolutionAlgorithm(rdl.IUWT)
alg.apply(**kw)
cube = rdl.Cube(k3_n, k3_n, 1, 1, None)
d_psf = torch.from_numpy(psfs).cuda()
times = []
for rep in range(2):
    d_res, d_model = torch.from_numpy(res.copy()).cuda(), torch.zeros((1, k3_n, k3_n), device="cuda")
    alg.set_iteration_number(0)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    g = alg.execute_major_iteration(cube, d_res, d_model, d_psf)
    torch.cuda.synchronize()
    times.append(time.perf_counter() - t0)
print(json.dumps({"config": f"K3 IuwtDeconvolution {k3_n}^2 x 1, max_iterations 20, mgain 0.8",
                  "cpu_reference_sources_s": cpu_s, "cores": cores, "gpu_s": min(times),
                  "gpu_first_call_s": times[0], "speedup": cpu_s / min(times),
                  "iterations": [int(ref.iteration_number), int(g.iteration_number)],
                  "residual_rel_diff": rel(d_res.cpu().numpy(), r_res),
                  "model_rel_diff": rel(d_model.cpu().numpy(), r_model)}), flush=True)
del d_res, d_model, d_psf

# ---- K5 -------------------------------------------------------------------------------------
res0, psfs, weights, _ = make_cube(k5_n, 1, "cpu")
res, psfs = res0.numpy(), psfs.numpy()
kw = dict(max_iterations=2000, minor_loop_gain=0.1, major_loop_gain=0.8, clean_border_ratio=0.0)
ref = oracle.Reference("generic", k5_n, k5_n, 1, use_sub_minor_optimization=False, grid=(8, 8),
                       threads=cores, **kw)
r_res, r_model = res.copy(), np.zeros_like(res)
t0 = time.perf_counter()
w = ref.execute(r_res, r_model, psfs)
cpu_s = time.perf_counter() - t0
alg = rdl.GenericClean(use_sub_minor_optimization=False)
alg.apply(**kw)
cube = rdl.Cube(k5_n, k5_n, 1, 1, None)
d_psf = torch.from_numpy(psfs).cuda()
times, geom = [], []
for rep in range(2):
    d_res, d_model = torch.from_numpy(res.copy()).cuda(), torch.zeros((1, k5_n, k5_n), device="cuda")
    alg.set_iteration_number(0)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    g = alg.parallel_execute_major_iteration(8, 8, cube, d_res, d_model, d_psf)
    torch.cuda.synchronize()
    times.append(time.perf_counter() - t0)
    geom.append(alg.parallel_timing()[0] * 1e-3)
print(json.dumps({"config": f"K5 ParallelDeconvolution {k5_n}^2 x 1, 8 x 8 facets, Hogbom path B, "
                            "max_iterations 2000 per facet, mgain 0.8 (divider search included on both sides)",
                  "cpu_reference_sources_s": cpu_s, "cores": cores, "gpu_s": min(times),
                  "gpu_host_divider_search_s": geom[int(np.argmin(times))],
                  "speedup": cpu_s / min(times),
                  "identical": bool(np.array_equal(d_res.cpu().numpy(), r_res) and
                                    np.array_equal(d_model.cpu().numpy(), r_model)),
                  "note": "the reference's threaded facet run is order dependent only for multiscale; "
                          "GenericClean facets are bit-identical"}), flush=True)
