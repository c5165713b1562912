"""K5: ParallelDeconvolution 16384^2 with 8x8 facets, Hogbom per facet (one B200), CUDA events +
the host share (Dijkstra dividers). python tools/facet_bench.py [SIZE] [GRID] [MAX_ITERATIONS]"""
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
dev = torch.device("cuda", 0)
res0, psfs, weights, _ = make_cube(n, 1, dev)
alg = rdl.GenericClean(use_sub_minor_optimization=False)
alg.apply(max_iterations=iters, clean_border_ratio=0.0, minor_loop_gain=0.1, major_loop_gain=0.8,
          threshold=0.0)
cube = rdl.Cube(n, n, 1, 1, None)
for rep in range(2):
    res, model = res0.clone(), torch.zeros_like(res0)
    alg.set_iteration_number(0)
    torch.cuda.synchronize()
    l0 = alg.kernel_launch_count
    t0 = time.perf_counter()
    r = alg.parallel_execute_major_iteration(g, g, cube, res, model, psfs)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    facets = alg.facets()
    host = res0[0].cpu().numpy()
    t1 = time.perf_counter()
    rdl.facet_grid(host, g, g)
    geom = time.perf_counter() - t1
    print(json.dumps({"n": n, "grid": g, "rep": rep, "wall_ms": wall * 1e3,
                      "host_geometry_ms": geom * 1e3, "device_and_copies_ms": (wall - geom) * 1e3,
                      "iterations_facet0": r.iteration_number,
                      "iterations_all_facets": int(sum(f["iteration_number"] for f in facets)),
                      "another_iteration_required": r.another_iteration_required,
                      "launches": alg.kernel_launch_count - l0,
                      "facet_boxes": [(f["width"], f["height"]) for f in facets[:4]]}))
    alg = rdl.GenericClean(use_sub_minor_optimization=False)
    alg.apply(max_iterations=iters, clean_border_ratio=0.0, minor_loop_gain=0.1,
              major_loop_gain=0.8, threshold=0.0)
