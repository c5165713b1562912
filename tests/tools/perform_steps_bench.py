"""Times csrc/image_statistics.cu (Radler::Perform's steps around the major iteration) on the
device against the reference's host code for the same step (oracle/_ref: image_set.cc compiled
against the aocommon stand-in), one JSON line per step.
    python tests/tools/perform_steps_bench.py [size] [n_original] [n_deconvolution]"""
import json
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import oracle  # noqa: E402  (the CPU arm only)
import ska_sdp_func_radler_b200 as rdl  # noqa: E402


def device_ms(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
    n_orig = int(sys.argv[2]) if len(sys.argv) > 2 else 16
    n_dec = int(sys.argv[3]) if len(sys.argv) > 3 else 4
    rng = np.random.default_rng(1)
    alg = rdl.GenericClean()
    plane = (rng.standard_normal((n, n)) * 1e-3).astype(np.float32)
    d = torch.from_numpy(plane).cuda()
    stats = alg.median_and_stddev_from_mad(d)
    ms = device_ms(lambda: alg.median_and_stddev_from_mad(d))
    line = {"step": "MedianAndStdDevFromMAD", "frame": n, "device_ms": round(ms, 3),
            "passes_over_the_plane": 6,
            "GB/s": round(6 * plane.nbytes / ms / 1e6, 1)}
    if oracle.ref_lib() is not None:
        t = time.perf_counter()
        rm, rs = oracle.ref_median_and_stddev_from_mad(plane)
        line["host_ms"] = round((time.perf_counter() - t) * 1e3, 1)
        line["identical"] = bool(np.float32(stats[0]) == rm and np.float32(stats[1]) == rs)
    print(json.dumps(line), flush=True)

    planes = (rng.standard_normal((n_orig, n, n))).astype(np.float32)
    weights = rng.uniform(0.5, 2.0, n_orig)
    index = [ch * n_dec // n_orig for ch in range(n_orig)]
    d_in = [torch.from_numpy(p).cuda() for p in planes]
    d_out = [torch.zeros((n, n), device="cuda") for _ in range(n_dec)]
    ms = device_ms(lambda: alg.load_and_average(d_in, weights, index, d_out))
    moved = (n_orig + n_dec) * plane.nbytes
    line = {"step": "LoadAndAverage", "frame": n, "original": n_orig, "deconvolution": n_dec,
            "device_ms": round(ms, 3), "GB/s": round(moved / ms / 1e6, 1)}
    if oracle.ref_lib() is not None:
        t = time.perf_counter()
        want = oracle.ref_load_and_average(planes, weights, 1, n_dec)
        line["host_ms"] = round((time.perf_counter() - t) * 1e3, 1)
        got = np.stack([t_.cpu().numpy() for t_ in d_out])
        line["identical"] = bool(np.array_equal(got.view(np.uint32), want.view(np.uint32)))
    print(json.dumps(line), flush=True)

    models = [torch.zeros((n, n), device="cuda") for _ in range(n_dec)]
    mask = torch.zeros((n, n), dtype=torch.uint8, device="cuda")
    ms = device_ms(lambda: alg.auto_mask_from_model(models, mask))
    print(json.dumps({"step": "auto-mask from model", "frame": n, "planes": n_dec,
                      "device_ms": round(ms, 3),
                      "GB/s": round((n_dec * plane.nbytes + n * n) / ms / 1e6, 1)}), flush=True)

    # InterpolateAndStoreModel: n_dec model planes (a few thousand components) -> n_orig planes
    orig = 1.2e8 + 2.0e6 * np.arange(n_orig)
    fit = orig.reshape(n_dec, -1).mean(axis=1)
    host_models = np.zeros((n_dec, n, n), np.float32)
    ys, xs = rng.integers(0, n, 5000), rng.integers(0, n, 5000)
    for d_ in range(n_dec):
        host_models[d_, ys, xs] = (rng.standard_normal(5000) + 3.0).astype(np.float32)
    alg.set_spectral_fitter(rdl.FIT_POLYNOMIAL, 2, fit, None)
    for m, hm in zip(models, host_models):
        m.copy_(torch.from_numpy(hm))
    outs = [torch.zeros((n, n), device="cuda") for _ in range(n_orig)]
    ms = device_ms(lambda: alg.interpolate_model(models, orig, outs))
    line = {"step": "InterpolateAndStoreModel", "frame": n, "deconvolution": n_dec,
            "original": n_orig, "terms": 2, "device_ms": round(ms, 3),
            "GB/s": round((n_dec + n_orig) * plane.nbytes / ms / 1e6, 1)}
    if oracle.ref_lib() is not None:
        t = time.perf_counter()
        want = oracle.ref_interpolate_model(host_models, 1, fit, np.ones(n_dec, np.float32), 2, orig)
        line["host_ms"] = round((time.perf_counter() - t) * 1e3, 1)
        got = np.stack([t_.cpu().numpy() for t_ in outs])
        line["identical"] = bool(np.array_equal(got, want))
    print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
