"""U8: IUWT decomposition / recomposition at 2048^2 x 6 scales (BASELINE config 3), CUDA events."""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import ska_sdp_func_radler_b200 as rdl

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
s = int(sys.argv[2]) if len(sys.argv) > 2 else 6
alg = rdl.DeconvolutionAlgorithm(rdl.IUWT)
side = torch.cuda.Stream()
torch.cuda.set_stream(side)
x = torch.randn((n, n), device="cuda")
coef = torch.empty((s + 1, n, n), device="cuda")
out = torch.empty((n, n), device="cuda")
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for name, fn, bytes_ in (
        ("decompose", lambda: alg.iuwt_decompose(x, coef, s, True, stream=side.cuda_stream), s * 20 * n * n),
        ("recompose", lambda: alg.iuwt_recompose(coef, out, s, True, stream=side.cuda_stream), (8 + s * 12) * n * n)):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(10):
        flush.zero_()  # L2 flush between timed iterations
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(side)
        fn()
        e1.record(side)
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = sorted(ts)[len(ts) // 2]
    print(json.dumps({"op": name, "n": n, "scales": s, "ms": ms, "algorithmic_bytes": bytes_,
                      "GBps": bytes_ / ms / 1e6, "reference_bytes": (s * 44 if name == "decompose" else s * 28) * n * n}))
