"""Times the FFT convolution engine alone (CUDA events, inputs larger than L2)."""
import json
import sys
import os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import ska_sdp_func_radler_b200 as rdl

sizes = [int(a) for a in sys.argv[1:]] or [4096, 4608, 8192, 9216, 9408]
alg = rdl.GenericClean()
side = torch.cuda.Stream()  # a NULL stream would mean "the handle's own stream"
torch.cuda.set_stream(side)
stream = side.cuda_stream
for n in sizes:
    n_img = 4
    imgs = torch.randn((n_img, n, n), device="cuda")
    ker = torch.randn((n, n), device="cuda")
    alg.convolve(imgs, ker, stream=stream)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 3
    e0.record()
    for _ in range(reps):
        alg.convolve(imgs, ker, stream=stream)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    # per call: 1 kernel spectrum (16 B/px) + n_img convolutions (28 B/px)
    bytes_ = (16 + 28 * n_img) * n * n
    print(json.dumps({"n": n, "ms_per_call": ms, "ms_per_convolution": ms / (n_img + 16 / 28),
                      "GBps": bytes_ / ms / 1e6}))
    del imgs, ker
