"""IUWT structure selection and plane measurements at 2048^2 x 6 scales (BASELINE config 3):
the device calls (host-visible result, so wall clock around a synchronising call, L2 flushed
between repetitions) beside the CPU oracle restatement on the same data (pass --cpu)."""
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import numpy as np
import torch

import ska_sdp_func_radler_b200 as rdl

n = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1].isdigit() else 2048
s = 6
with_cpu = "--cpu" in sys.argv
alg = rdl.DeconvolutionAlgorithm(rdl.IUWT)
g = torch.Generator(device="cuda").manual_seed(5)
c = torch.randn((s, n, n), device="cuda", generator=g)
yy, xx = torch.meshgrid(torch.arange(n, device="cuda"), torch.arange(n, device="cuda"), indexing="ij")
for k in range(40):  # extended sources: structures a few hundred pixels across
    cx, cy, r = (k * 397) % n, (k * 911) % n, 20 + 5 * (k % 9)
    c += 8.0 * torch.exp(-((xx - cx) ** 2 + (yy - cy) ** 2) / (2.0 * r * r))[None]
thr = [3.0] * s
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
masks = torch.zeros((s, n, n), dtype=torch.uint8, device="cuda")
seed = (397 % n, 911 % n, 2)


def timed(fn, reset=None, reps=7):
    ts = []
    for _ in range(reps):
        if reset:
            reset()
        flush.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        r = fn()
        ts.append((time.perf_counter() - t0) * 1e3)
    return sorted(ts)[len(ts) // 2], r


rows = [
    ("select_structures", lambda: alg.iuwt_select_structures(c, masks, thr, 0, s, 0.0),
     masks.zero_, 5 * s * n * n),
    ("floodfill", lambda: alg.iuwt_floodfill(c, masks, thr, 0, s, seed, 0.0), masks.zero_,
     7 * s * n * n),
    ("mad", lambda: alg.iuwt_mad(c[1]), None, 4 * n * n),
    ("max_abs", lambda: alg.iuwt_max_abs(c[1], 0.0, True), None, 4 * n * n),
]
cpu = {}
if with_cpu:
    import oracle
    sel, fill, _, mad, max_abs = oracle.iuwt_analysis()
    hc = c.cpu().numpy()
    hm = np.zeros((s, n, n), np.uint8)
    ht = np.array(thr, np.float32)
    for name, fn in (("select_structures", lambda: sel(hc, hm, ht, 0, s, 0.0)[1]),
                     ("floodfill", lambda: fill(hc, hm, ht, 0, s, seed, 0.0)[1]),
                     ("mad", lambda: mad(hc[1])), ("max_abs", lambda: max_abs(hc[1], 0.0, True))):
        t0 = time.perf_counter()
        r = fn()
        cpu[name] = ((time.perf_counter() - t0) * 1e3, r)
for name, fn, reset, bytes_ in rows:
    fn()
    ms, r = timed(fn, reset)
    line = {"op": name, "n": n, "scales": s, "ms": ms, "result": r, "algorithmic_bytes": bytes_,
            "GBps": bytes_ / ms / 1e6}
    if name in cpu:
        line["cpu_oracle_ms"] = cpu[name][0]
        line["cpu_result_equal"] = bool(cpu[name][1] == r)
    print(json.dumps(line))
