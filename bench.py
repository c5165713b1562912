#!/usr/bin/env python
"""bench.py — the Radler minor-loop hot path on B200, next to the reference's CPU path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
                    [--workload hogbom|multiscale] [--size 8192] [--channels 16]

A "step" is one ExecuteMajorIteration (deconvolution_algorithm.h:69-71) on one synthetic
cube. Workloads (BASELINE.json metric "CLEAN iters/sec & multiscale major-iter ms at
8192^2 x 16ch"):
  hogbom      GenericClean path B (FindPeak -> SubtractImage loop) on a joined 8192^2 x 16
              channel cube, `--iters` minor iterations per step; metric = CLEAN iterations/s.
Prints ONE JSON line (see DESIGN.md "Measurement").
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


# --------------------------------------------------------------------------------------
# synthetic cube: point sources (*) Gaussian PSF + white noise, power-law
# spectra, per-channel PSF widths, channel weights. Built with torch FFTs (plumbing, untimed).
# --------------------------------------------------------------------------------------
def make_cube(n, n_channels, device, seed=1):
    import torch
    g = torch.Generator(device="cpu").manual_seed(seed)
    n_src = max(8, n * n // 4096)
    xs = torch.randint(16, n - 16, (n_src,), generator=g)
    ys = torch.randint(16, n - 16, (n_src,), generator=g)
    flux = torch.exp(torch.empty(n_src).uniform_(np.log(0.01), np.log(10.0), generator=g))
    alpha = torch.empty(n_src).uniform_(-1.2, -0.2, generator=g)
    weights = torch.empty(n_channels).uniform_(0.5, 1.5, generator=g).numpy().astype(np.float32)
    freqs = np.linspace(100e6, 250e6, n_channels) if n_channels > 1 else np.array([150e6])
    yy, xx = torch.meshgrid(torch.arange(n, device=device, dtype=torch.float32),
                            torch.arange(n, device=device, dtype=torch.float32), indexing="ij")
    r = torch.hypot(xx - n // 2, yy - n // 2)
    del xx, yy
    res = torch.empty((n_channels, n, n), dtype=torch.float32, device=device)
    psfs = torch.empty((n_channels, n, n), dtype=torch.float32, device=device)
    for c in range(n_channels):
        fwhm = 4.0 * (freqs[0] / freqs[c]) ** 0.25
        sigma = fwhm / 2.3548200450309493
        p = torch.exp(-0.5 * (r / sigma) ** 2) + 0.01 * torch.cos(r / 8.0) * torch.exp(-r / 200.0) * (r > 0)
        p[p.abs() < 1e-7] = 0.0
        p[n // 2, n // 2] = 1.0
        psfs[c] = p
        sky = torch.zeros((n, n), dtype=torch.float32, device=device)
        f_c = flux * (freqs[c] / freqs[0]) ** alpha
        sky.index_put_((ys.to(device), xs.to(device)), f_c.to(device), accumulate=True)
        # circular FFT convolution with the PSF centre moved to the origin (compact PSF)
        k = torch.roll(p, shifts=(-(n // 2), -(n // 2)), dims=(0, 1))
        d = torch.fft.irfft2(torch.fft.rfft2(sky) * torch.fft.rfft2(k), s=(n, n))
        gen = torch.Generator(device=device).manual_seed(seed * 1000 + c)
        d += 1e-3 * torch.randn((n, n), generator=gen, device=device, dtype=torch.float32)
        res[c] = d
        del sky, k, d, p
    return res, psfs, weights, freqs


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                 "--format=csv,noheader,nounits", "-lms", "100"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([v.strip() for v in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm, mx, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
            except (ValueError, IndexError):
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                                "sw_power_cap"], r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def window_area(n, x, y):
    """Clipped PSF window of SubtractImage (simple_clean.cc:61-96), per component."""
    sx = np.maximum(x - n // 2, 0)
    ex = np.minimum(x + n // 2, n)
    sy = np.maximum(y - n // 2, 0)
    ey = np.minimum(y + n // 2, n)
    return (ex - sx).astype(np.float64) * (ey - sy)


# --------------------------------------------------------------------------------------
# CPU arm: the oracle (a port of the reference's CPU path; the reference itself cannot be
# built offline — see DESIGN.md) on all host cores, on a bounded sample of the workload.
# --------------------------------------------------------------------------------------
def cpu_hogbom(res, psfs, weights, iters):
    import oracle
    oracle.build()
    cores = os.cpu_count() or 1
    oracle.set_threads(cores)
    c, n, _ = res.shape
    model = np.zeros_like(res)
    alg = oracle.GenericClean(use_sub_minor_optimization=False)
    s = oracle.Settings(max_iterations=iters, clean_border_ratio=0.0, minor_loop_gain=0.1,
                        threshold=0.0)
    t0 = time.perf_counter()
    r = alg.execute(oracle.Cube(n, n, c, 1, weights), s, res, model, psfs)
    dt = time.perf_counter() - t0
    return r.iteration_number / dt, dt, cores, r.iteration_number


def run_reference(args):
    """--impl reference: the CPU path on the host cores, same metric/config, bounded sample."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import torch
    n, c = args.size, args.channels
    res, psfs, weights, _ = make_cube(n, c, "cpu")
    res, psfs = res.numpy(), psfs.numpy()
    sample_iters = args.cpu_iters
    vals, times = [], []
    for step in range(args.warmup + args.steps):
        work = res.copy()
        v, dt, cores, it = cpu_hogbom(work, psfs, weights, sample_iters)
        if step >= args.warmup:
            vals.append(v)
            times.append(dt)
    value = float(np.mean(vals))
    line = base_line(args, value, float(np.mean(times)) * 1e3)
    line.update({
        "impl": "reference", "n_gpus": args.gpus,
        "cpu_baseline": {"value": value, "unit": "iterations/s", "cores": cores, "kind": "port",
                         "sample": f"{sample_iters} minor iterations per step of the same "
                                   f"{n}^2 x {c} cube (oracle GenericClean path B, all host cores)"},
        "e2e": {"value": value, "unit": "iterations/s", "h2d_bytes_per_step": 0,
                "d2h_bytes_per_step": 0},
        "gpu_launches": 0, "dtype": "f32",
    })
    print(json.dumps(line))


def base_line(args, value, ms_per_step):
    return {
        "metric": "CLEAN minor iterations/s (GenericClean FindPeak+SubtractImage loop, joined channels)",
        "value": value, "unit": "iterations/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"hogbom {args.size}^2 x {args.channels} channels joined, "
                               f"{args.iters} minor iterations per major iteration (step), "
                               "GenericClean path B",
                   "size": args.size, "channels": args.channels, "iters_per_step": args.iters,
                   "l2": "inputs larger than L2 (cube >> 126 MB)" if
                   args.size * args.size * args.channels * 4 > 4 * 126e6 else "L2-resident cube",
                   "parallelism": "1 cube per GPU (replicas)" if args.gpus > 1 else "single GPU"},
    }


def run_ours(args):
    import torch
    import torch.distributed as dist

    import ska_sdp_func_radler_b200 as rdl
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)
    n, c, iters = args.size, args.channels, args.iters
    res0, psfs, weights, freqs = make_cube(n, c, dev, seed=1 + rank)
    cube = rdl.Cube(n, n, c, 1, weights)
    alg = rdl.GenericClean(use_sub_minor_optimization=False, device=local)
    alg.apply(max_iterations=iters, clean_border_ratio=0.0, minor_loop_gain=0.1, threshold=0.0,
              trace_capacity=iters)
    stream = torch.cuda.current_stream().cuda_stream
    res = res0.clone()
    model = torch.zeros_like(res)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device():
        alg.set_iteration_number(0)
        return alg.execute_major_iteration(cube, res, model, psfs, stream=stream)

    # ---- device-resident: cube already in HBM when the timed region starts ----------------
    for _ in range(args.warmup):
        step_device()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = alg.kernel_launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    total_iters, alg_bytes, u3_bytes = 0, 0.0, 0.0
    e0.record()
    for _ in range(args.steps):
        r = step_device()
        total_iters += r.iteration_number
        tr = alg.trace()
        a_sub = window_area(n, tr["x"].astype(np.int64), tr["y"].astype(np.int64))
        # fused kernel minimum: every image read once (4*C*A) + PSF read and residual
        # write-back inside the window (8*C*A_sub). SURVEY U3 = 12*C*A_sub + 4*C*A_pk.
        alg_bytes += float((4.0 * c * n * n + 8.0 * c * a_sub).sum())
        u3_bytes += float((4.0 * c * n * n + 12.0 * c * a_sub).sum())
    e1.record()
    barrier()
    dev_ms = e0.elapsed_time(e1)
    launches = alg.kernel_launch_count - launches0
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([dev_ms, float(total_iters)], dtype=torch.float64, device=dev)
    if world > 1:
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone()
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        dev_ms_max, iters_all = float(tmax[0]), float(tsum[1])
    else:
        dev_ms_max, iters_all = dev_ms, float(total_iters)
    value = iters_all / (dev_ms_max * 1e-3)

    # ---- end to end through the C ABI with HOST buffers (H2D + D2H inside the call) ---------
    h_res = torch.empty((c, n, n), dtype=torch.float32).pin_memory()
    h_model = torch.zeros((c, n, n), dtype=torch.float32).pin_memory()
    h_psf = torch.empty((c, n, n), dtype=torch.float32).pin_memory()
    h_res.copy_(res0)
    h_psf.copy_(psfs)
    del res, model
    torch.cuda.synchronize()
    np_res, np_model, np_psf = h_res.numpy(), h_model.numpy(), h_psf.numpy()
    e2e_steps = max(1, min(args.steps, 3))
    alg.set_iteration_number(0)
    alg.execute_major_iteration(cube, np_res, np_model, np_psf)  # warm-up (staging buffers)
    barrier()
    t0 = time.perf_counter()
    e2e_iters = 0
    for _ in range(e2e_steps):
        alg.set_iteration_number(0)
        e2e_iters += alg.execute_major_iteration(cube, np_res, np_model, np_psf).iteration_number
    barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s, float(e2e_iters)], dtype=torch.float64, device=dev)
    if world > 1:
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone()
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        e2e_s, e2e_iters = float(tmax[0]), float(tsum[1])
    plane = n * n * 4
    e2e = {"value": e2e_iters / e2e_s, "unit": "iterations/s",
           "h2d_bytes_per_step": 3 * c * plane, "d2h_bytes_per_step": 2 * c * plane,
           "ms_per_step": e2e_s / e2e_steps * 1e3,
           "note": "rdl_execute_major_iteration with RDL_MEM_HOST pinned buffers: residual, model "
                   "and PSFs copied to HBM, residual and model copied back, every step"}

    if rank == 0:
        peak, peak_src = measured_peaks()
        avg_launch_s = dev_ms * 1e-3 / max(launches, 1)
        achieved = alg_bytes / max(launches, 1) / avg_launch_s / 1e9
        line = base_line(args, value, dev_ms_max / args.steps)
        line.update({
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": "HogbomStepKernel (fused SubtractImage + "
                         "channel integration + arg-max)", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                         "peak_source": peak_src,
                         "bytes_per_launch": alg_bytes / max(launches, 1),
                         "avg_launch_us": avg_launch_s * 1e6,
                         "achieved_survey_u3": u3_bytes / max(launches, 1) / avg_launch_s / 1e9,
                         "note": "achieved uses the fused-kernel minimum bytes (each image read "
                                 "once); achieved_survey_u3 uses SURVEY 8(d) U3 = U1 + U2"},
        })
        if not args.no_cpu:
            h = res0.cpu().numpy()
            v, dt, cores, it = cpu_hogbom(h, np_psf, weights, args.cpu_iters)
            line["cpu_baseline"] = {
                "value": v, "unit": "iterations/s", "cores": cores, "kind": "port",
                "sample": f"{it} minor iterations of the same {n}^2 x {c} cube "
                          f"({dt:.1f} s, oracle GenericClean path B, all host cores)"}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="hogbom", choices=["hogbom"])
    ap.add_argument("--size", type=int, default=8192)
    ap.add_argument("--channels", type=int, default=16)
    ap.add_argument("--iters", type=int, default=256, help="minor iterations per step")
    ap.add_argument("--cpu-iters", type=int, default=6, help="CPU sample: minor iterations")
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3 if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
