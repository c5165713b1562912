"""Device-resident Högbom loop (GenericClean path B) at the bench size, one JSON line: iterations/s
and the roofline fraction, for trying kernel variants (RDL_HOGBOM_BATCH, RDL_HOGBOM_BPS).
python tools/hogbom_probe.py [size] [channels] [iterations] [steps]"""
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402
import ska_sdp_func_radler_b200 as rdl  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
c = int(sys.argv[2]) if len(sys.argv) > 2 else 16
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 256
steps = int(sys.argv[4]) if len(sys.argv) > 4 else 2
dev = torch.device("cuda", 0)
res0, psfs, weights, freqs = bench.make_cube(n, c, dev, seed=1)
cube = rdl.Cube(n, n, c, 1, weights)
alg = rdl.GenericClean(use_sub_minor_optimization=False)
alg.apply(max_iterations=iters, clean_border_ratio=0.0, minor_loop_gain=0.1, threshold=0.0,
          trace_capacity=iters)
res, model = res0.clone(), torch.zeros_like(res0)
# PROBE_PAD_KB: planes at a stride of n * n * 4 + pad bytes instead of back to back (power-of-two
# plane strides put pixel (x, y) of every image at the same offset modulo any power of two)
pad = int(os.environ.get("PROBE_PAD_KB", "0")) * 256  # floats
if pad:
    def spread(t):
        big = torch.empty((c, n * n + pad), device=dev)
        planes = [big[i, :n * n].view(n, n) for i in range(c)]
        for i in range(c):
            planes[i].copy_(t[i])
        return planes, big
    res, _keep1 = spread(res)
    model, _keep2 = spread(model)
    psfs, _keep3 = spread(psfs)
stream = torch.cuda.current_stream().cuda_stream


def step():
    alg.set_iteration_number(0)
    return alg.execute_major_iteration(cube, res, model, psfs, stream=stream)


step()
torch.cuda.synchronize()


def clocks():
    try:
        import pynvml
        pynvml.nvmlInit()
        hnd = pynvml.nvmlDeviceGetHandleByIndex(0)
        return {"sm_mhz": pynvml.nvmlDeviceGetClockInfo(hnd, pynvml.NVML_CLOCK_SM),
                "mem_mhz": pynvml.nvmlDeviceGetClockInfo(hnd, pynvml.NVML_CLOCK_MEM),
                "power_w": pynvml.nvmlDeviceGetPowerUsage(hnd) / 1000.0,
                "temp_c": pynvml.nvmlDeviceGetTemperature(hnd, 0)}
    except Exception as e:  # noqa: BLE001
        return {"error": str(e)}


total, alg_bytes, ms, per_step = 0, 0.0, 0.0, []
for _ in range(steps):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    r = step()
    e1.record()
    torch.cuda.synchronize()
    per_step.append(round(e0.elapsed_time(e1), 3))
    ms += e0.elapsed_time(e1)
    total += r.iteration_number
    tr = alg.trace()
    a_sub = bench.window_area(n, tr["x"].astype(np.int64), tr["y"].astype(np.int64), None)
    alg_bytes += float((4.0 * c * n * n + 8.0 * c * a_sub).sum())
clk = clocks()
peak, _ = bench.measured_peaks()
print(json.dumps({"size": n, "channels": c, "batch": os.environ.get("RDL_HOGBOM_BATCH", "default"),
                  "blocks_per_sm": os.environ.get("RDL_HOGBOM_BPS", "default"),
                  "iterations_per_s": total / (ms * 1e-3), "us_per_iteration": ms * 1e3 / total,
                  "achieved_gbs": alg_bytes / (ms * 1e-3) / 1e9,
                  "frac": alg_bytes / (ms * 1e-3) / 1e9 / peak, "ms_per_step": per_step,
                  "persistent_mb": os.environ.get("RDL_HOGBOM_PERSISTENT_MB", "default"),
                  "resident": os.environ.get("RDL_HOGBOM_RESIDENT", "default"), "plane_pad_kb": os.environ.get("PROBE_PAD_KB", "0"), "clocks": clk}))
