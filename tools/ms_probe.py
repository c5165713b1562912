"""Times one MultiScaleAlgorithm::ExecuteMajorIteration on the device at a given size
(CUDA events, cube resident in HBM): python tools/ms_probe.py SIZE CHANNELS MAX_ITERATIONS"""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import ska_sdp_func_radler_b200 as rdl
from bench import make_ms_cube

n, c, iters = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
dev = torch.device("cuda", 0)
res0, psfs, weights, freqs = make_ms_cube(n, c, dev)
alg = rdl.MultiScaleAlgorithm(6.0, 1.0, max_scales=5)
alg.apply(max_iterations=iters, clean_border_ratio=0.0, minor_loop_gain=0.1, major_loop_gain=0.8,
          threshold=0.0, trace_capacity=0)
if c > 1:
    alg.set_spectral_fitter(rdl.FIT_POLYNOMIAL, 2, freqs, weights, 1)
cube = rdl.Cube(n, n, c, 1, weights)
stream = torch.cuda.current_stream().cuda_stream
for rep in range(int(sys.argv[4]) if len(sys.argv) > 4 else 2):
    res = res0.clone()
    model = torch.zeros_like(res)
    alg.set_iteration_number(0)
    torch.cuda.synchronize()
    l0 = alg.kernel_launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    e0.record()
    r = alg.execute_major_iteration(cube, res, model, psfs, stream=stream)
    e1.record()
    torch.cuda.synchronize()
    print(json.dumps({"n": n, "channels": c, "rep": rep, "ms": e0.elapsed_time(e1),
                      "wall_ms": (time.perf_counter() - t0) * 1e3,
                      "iterations": r.iteration_number,
                      "minor_iterations": alg.minor_iteration_count,
                      "launches": alg.kernel_launch_count - l0,
                      "scales": [s["scale"] for s in alg.scale_infos()],
                      "mem_GB": torch.cuda.mem_get_info()[0] / 1e9}))
