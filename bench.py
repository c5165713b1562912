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
def make_cube(n, n_channels, device, seed=1, fwhm0=4.0, extended=False):
    import torch
    g = torch.Generator(device="cpu").manual_seed(seed)
    n_src = max(8, n * n // 4096)
    xs = torch.randint(16, n - 16, (n_src,), generator=g)
    ys = torch.randint(16, n - 16, (n_src,), generator=g)
    flux = torch.exp(torch.empty(n_src).uniform_(np.log(0.01), np.log(10.0), generator=g))
    alpha = torch.empty(n_src).uniform_(-1.2, -0.2, generator=g)
    weights = torch.empty(n_channels).uniform_(0.5, 1.5, generator=g).numpy().astype(np.float32)
    # K2/K4: A/262144 extended Gaussian blobs, sigma in {6,12,24,48} px, peak 0.05-0.5 Jy/px
    n_ext = max(2, n * n // 262144) if extended else 0
    ex = torch.randint(n // 8, n - n // 8, (n_ext,), generator=g) if n_ext else None
    ey = torch.randint(n // 8, n - n // 8, (n_ext,), generator=g) if n_ext else None
    es = torch.tensor([6.0, 12.0, 24.0, 48.0])[torch.randint(0, 4, (n_ext,), generator=g)] if n_ext else None
    ep = torch.empty(n_ext).uniform_(0.05, 0.5, generator=g) if n_ext else None
    ea = torch.empty(n_ext).uniform_(-1.2, -0.2, generator=g) if n_ext else None
    freqs = np.linspace(100e6, 250e6, n_channels) if n_channels > 1 else np.array([150e6])
    yy, xx = torch.meshgrid(torch.arange(n, device=device, dtype=torch.float32),
                            torch.arange(n, device=device, dtype=torch.float32), indexing="ij")
    r = torch.hypot(xx - n // 2, yy - n // 2)
    del xx, yy
    res = torch.empty((n_channels, n, n), dtype=torch.float32, device=device)
    psfs = torch.empty((n_channels, n, n), dtype=torch.float32, device=device)
    for c in range(n_channels):
        fwhm = fwhm0 * (freqs[0] / freqs[c]) ** 0.25
        sigma = fwhm / 2.3548200450309493
        p = torch.exp(-0.5 * (r / sigma) ** 2) + 0.01 * torch.cos(r / 8.0) * torch.exp(-r / 200.0) * (r > 0)
        p[p.abs() < 1e-7] = 0.0
        p[n // 2, n // 2] = 1.0
        psfs[c] = p
        sky = torch.zeros((n, n), dtype=torch.float32, device=device)
        f_c = flux * (freqs[c] / freqs[0]) ** alpha
        sky.index_put_((ys.to(device), xs.to(device)), f_c.to(device), accumulate=True)
        for b in range(n_ext):
            hw = int(5 * float(es[b]))
            x0, y0 = int(ex[b]), int(ey[b])
            xa, xb, ya, yb = max(x0 - hw, 0), min(x0 + hw + 1, n), max(y0 - hw, 0), min(y0 + hw + 1, n)
            gx = torch.arange(xa, xb, device=device, dtype=torch.float32) - x0
            gy = torch.arange(ya, yb, device=device, dtype=torch.float32) - y0
            amp = float(ep[b]) * (freqs[c] / freqs[0]) ** float(ea[b])
            sky[ya:yb, xa:xb] += amp * torch.exp(-0.5 * (gy[:, None] ** 2 + gx[None, :] ** 2) / float(es[b]) ** 2)
        # circular FFT convolution with the PSF centre moved to the origin (compact PSF)
        k = torch.roll(p, shifts=(-(n // 2), -(n // 2)), dims=(0, 1))
        d = torch.fft.irfft2(torch.fft.rfft2(sky) * torch.fft.rfft2(k), s=(n, n))
        gen = torch.Generator(device=device).manual_seed(seed * 1000 + c)
        d += 1e-3 * torch.randn((n, n), generator=gen, device=device, dtype=torch.float32)
        res[c] = d
        del sky, k, d, p
    return res, psfs, weights, freqs


def make_ms_cube(n, n_channels, device, seed=1):
    """K2/K4 cube of SURVEY 8(d): 6 px beam, point sources + extended blobs."""
    return make_cube(n, n_channels, device, seed=seed, fwhm0=6.0, extended=True)


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap,power.limit,clocks.mem")

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
        power, limit, mem = [], None, []
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = float(r[1])
            except (ValueError, IndexError):
                continue
            try:  # what a power-capped box looks like: draw at the limit, SM clock below max
                power.append(float(r[2]))
                limit = float(r[7])
                mem.append(float(r[8]))
            except (ValueError, IndexError):
                pass
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                                "sw_power_cap"], r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": mx,
                "reasons": sorted(reasons), "samples": len(sm),
                "sm_min_mhz": float(min(sm)) if sm else None,
                "power_w": float(np.median(power)) if power else None, "power_limit_w": limit,
                "mem_mhz": float(np.median(mem)) if mem else None}


def profile_traffic(name, kernels=None):
    """DRAM bytes (read + write) of the named kernels from an `ncu --set full` summary committed
    under profiles/ (tools/ncu_full_summary.py), or None."""
    path = os.path.join(ROOT, "profiles", name)
    if not os.path.exists(path):
        return None
    total, seen = 0.0, set()
    for k in json.load(open(path))["kernels"]:
        key = k["kernel"].split("(")[0]
        if kernels is not None and not any(key.startswith(w) for w in kernels):
            continue
        if key in seen:  # one launch per kernel
            continue
        seen.add(key)
        total += (k.get("dram_read_MB", 0.0) + k.get("dram_write_MB", 0.0)) * 1e6
    return total or None


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def window_area(n, x, y, owned_rows=None):
    """Clipped PSF window of SubtractImage (simple_clean.cc:61-96), per component, restricted
    to the frame rows a rank owns (all rows when None)."""
    sx = np.maximum(x - n // 2, 0)
    ex = np.minimum(x + n // 2, n)
    sy = np.maximum(y - n // 2, 0)
    ey = np.minimum(y + n // 2, n)
    if owned_rows is None:
        rows = ey - sy
    else:
        own = np.zeros(n + 1, np.int64)
        own[np.asarray(owned_rows) + 1] = 1
        cum = np.cumsum(own)
        rows = cum[ey] - cum[sy]
    return (ex - sx).astype(np.float64) * rows


# --------------------------------------------------------------------------------------
# CPU arm: the oracle (a port of the reference's CPU path; the reference itself cannot be
# built offline — see DESIGN.md) on all host cores, on a bounded sample of the workload.
# --------------------------------------------------------------------------------------
def cpu_kind():
    import oracle
    oracle.build()
    return "reference" if oracle.ref_lib() is not None and oracle._refa() is not None else "port"


def cpu_multiscale(res, psfs, weights, freqs, iters):
    """One MultiScaleAlgorithm::ExecuteMajorIteration on the host cores (K2/K4 settings)."""
    import oracle
    oracle.build()
    cores = os.cpu_count() or 1
    oracle.set_threads(cores)
    c, n, _ = res.shape
    model = np.zeros_like(res)
    kind = cpu_kind()
    if kind == "reference":
        alg = oracle.Reference("multiscale", n, n, c, weights=weights, frequencies=freqs,
                               max_iterations=iters, clean_border_ratio=0.0, minor_loop_gain=0.1,
                               major_loop_gain=0.8, threshold=0.0, max_scales=5, beam_size=6.0,
                               pixel_scale=1.0, fit_terms=2 if c > 1 else 0, threads=cores)
        t0 = time.perf_counter()
        r = alg.execute(res, model, psfs)
    else:
        alg = oracle.MultiScale(beam_size=6.0, pixel_scale=1.0, max_scales=5)
        st = oracle.Settings(max_iterations=iters, clean_border_ratio=0.0, minor_loop_gain=0.1,
                             major_loop_gain=0.8, threshold=0.0)
        if c > 1:
            st.spectral_mode, st.n_terms, st.frequencies, st.fit_weights = 1, 2, freqs, weights
        t0 = time.perf_counter()
        r = alg.execute(oracle.Cube(n, n, c, 1, weights), st, res, model, psfs)
    dt = time.perf_counter() - t0
    return dt, cores, r.iteration_number, kind


def cpu_hogbom_slope(res0, psfs, weights, iters_a, iters_b):
    """Reference CPU arm of the Hogbom loop as a SLOPE: two runs of iters_a and iters_b minor
    iterations from the same cube; (iters_b - iters_a) / (t_b - t_a) leaves out what a call pays
    once (building the ImageSets, copying 21 GB of planes in and out, the first integrate +
    FindPeak). Returns the rate, both timings and the images of the longer run."""
    work = res0.copy()
    _, t_a, cores, it_a, _ = cpu_hogbom(work, psfs, weights, iters_a, return_model=True)
    work = res0.copy()
    _, t_b, cores, it_b, model = cpu_hogbom(work, psfs, weights, iters_b, return_model=True)
    rate = (it_b - it_a) / max(t_b - t_a, 1e-9)
    return {"rate": rate, "t_a": t_a, "t_b": t_b, "it_a": it_a, "it_b": it_b, "cores": cores,
            "lump_rate": it_b / t_b, "residual": work, "model": model}


def cpu_convolution_seconds(size, cores, reps=2):
    """One schaapcommon::math::Convolve of a size^2 frame on the host cores: (a) through the code
    the reference arm really runs (the stand-in of schaapcommon forwards to oracle/fft.cc), (b) the
    same three transforms with MKL (torch.fft on the CPU). BASELINE.md 2.3: the FASTER one is the
    denominator, so the speed-up is not flattered by our own CPU FFT."""
    import torch

    import oracle
    oracle.build()
    oracle.set_threads(cores)
    rng = np.random.default_rng(size)
    img = rng.standard_normal((size, size), dtype=np.float32)
    ker = rng.standard_normal((size, size), dtype=np.float32)
    t_own = []
    for _ in range(reps):
        t0 = time.perf_counter()
        oracle.convolve(img, ker)
        t_own.append(time.perf_counter() - t0)
    torch.set_num_threads(cores)
    a, k = torch.from_numpy(img), torch.from_numpy(ker)
    t_mkl = []
    for _ in range(reps + 1):
        t0 = time.perf_counter()
        torch.fft.irfft2(torch.fft.rfft2(a) * torch.fft.rfft2(k), s=(size, size))
        t_mkl.append(time.perf_counter() - t0)
    return min(t_own), min(t_mkl[1:])


def price_reference_convolutions(counts, cores):
    """CPU time of the convolutions the reference makes for a major iteration: its own count by
    frame size (rdl_reference_convolutions, checked against the reference's sources in
    tests/test_gpu_parity_surface.py) x the measured per-convolution time on this host."""
    total_own = total_mkl = total_best = 0.0
    per_size = {}
    for (w, h), count in sorted(counts.items()):
        own, mkl = cpu_convolution_seconds(w, cores)
        per_size[f"{w}x{h}"] = {"count": count, "reference_fft_s": own, "mkl_fft_s": mkl}
        total_own += count * own
        total_mkl += count * mkl
        total_best += count * min(own, mkl)
    return total_own, total_mkl, total_best, per_size


def cpu_hogbom(res, psfs, weights, iters, return_model=False):
    if return_model:
        v, dt, cores, it = cpu_hogbom(res, psfs, weights, iters)  # res is updated in place
        return v, dt, cores, it, _LAST_MODEL.pop()
    import oracle
    oracle.build()
    cores = os.cpu_count() or 1
    oracle.set_threads(cores)
    c, n, _ = res.shape
    model = np.zeros_like(res)
    _LAST_MODEL.clear()
    if cpu_kind() == "reference":
        alg = oracle.Reference("generic", n, n, c, weights=weights,
                               use_sub_minor_optimization=False, max_iterations=iters,
                               clean_border_ratio=0.0, minor_loop_gain=0.1, threshold=0.0,
                               threads=cores)
        t0 = time.perf_counter()
        r = alg.execute(res, model, psfs)
        dt = time.perf_counter() - t0
        _LAST_MODEL.append(model)
        return r.iteration_number / dt, dt, cores, r.iteration_number
    alg = oracle.GenericClean(use_sub_minor_optimization=False)
    s = oracle.Settings(max_iterations=iters, clean_border_ratio=0.0, minor_loop_gain=0.1,
                        threshold=0.0)
    t0 = time.perf_counter()
    r = alg.execute(oracle.Cube(n, n, c, 1, weights), s, res, model, psfs)
    dt = time.perf_counter() - t0
    _LAST_MODEL.append(model)
    return r.iteration_number / dt, dt, cores, r.iteration_number


_LAST_MODEL = []


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
    vals, times, lumps = [], [], []
    # the CPU arm needs no warm-up beyond one step (no clocks to ramp, no JIT)
    for step in range(min(args.warmup, 1) + args.steps):
        sl = cpu_hogbom_slope(res, psfs, weights, 2, 2 + sample_iters)
        cores = sl["cores"]
        if step >= min(args.warmup, 1):
            vals.append(sl["rate"])
            lumps.append(sl["lump_rate"])
            times.append((sl["t_b"] - sl["t_a"]) / max(sl["it_b"] - sl["it_a"], 1) * args.iters)
    value = float(np.mean(vals))
    line = base_line(args, value, float(np.mean(times)) * 1e3)
    line.update({
        "impl": "reference", "n_gpus": args.gpus,
        "cpu_baseline": {"value": value, "unit": "iterations/s", "cores": cores, "kind": cpu_kind(),
                         "sample": f"slope between runs of 2 and {2 + sample_iters} minor iterations "
                                   f"of the same {n}^2 x {c} cube (GenericClean path B, reference "
                                   "sources, all host cores); ms_per_step = that rate applied to "
                                   f"the {args.iters} iterations of a step",
                         "lump_value": float(np.mean(lumps)),
                         "lump_note": "iterations / wall time of the longer run, ImageSet set-up "
                                      "and plane copies included (what round 1 reported)"},
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
        "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"hogbom {args.size}^2 x {args.channels} channels joined, "
                               f"{args.iters} minor iterations per major iteration (step), "
                               "GenericClean path B",
                   "size": args.size, "channels": args.channels, "iters_per_step": args.iters,
                   "l2": "inputs larger than L2 (cube >> 126 MB)" if
                   args.size * args.size * args.channels * 4 > 4 * 126e6 else "L2-resident cube",
                   "parallelism": f"rows of one cube sharded over {args.gpus} GPUs (cyclic 64-row blocks), NCCL joined "
                                  "arg-max (all-gather of one 264-byte candidate per rank per "
                                  "minor iteration)" if args.gpus > 1 else "single GPU"},
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
    from ska_sdp_func_radler_b200 import distributed as rd
    n, c, iters = args.size, args.channels, args.iters
    # N > 1: ONE cube, its rows sharded over the ranks (strong scaling); every rank generates the
    # same synthetic frame and keeps its band. The joined arg-max is the only exchange.
    res0, psfs, weights, freqs = make_cube(n, c, dev, seed=1)
    block_rows = 64  # cyclic row blocks: every rank sees the same share of each PSF window
    owned = None
    if world > 1:
        owned = rd.cyclic_rows(n, rank, world, block_rows)
        res0 = res0.index_select(1, torch.from_numpy(owned).to(dev)).contiguous()
    rows = n // world
    cube = rdl.Cube(n, rows, c, 1, weights)
    alg = rdl.GenericClean(use_sub_minor_optimization=False, device=local)
    alg.apply(max_iterations=iters, clean_border_ratio=0.0, minor_loop_gain=0.1, threshold=0.0,
              trace_capacity=iters)
    if world > 1:
        alg.dist_init(rd.share_unique_id(rdl.dist_unique_id, rank, world, device=dev), rank, world)
        alg.dist_set_row_blocks(block_rows, n)
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
    traces = []
    e0.record()
    for _ in range(args.steps):
        r = step_device()
        total_iters += r.iteration_number
        traces.append(alg.trace().copy())  # the byte accounting below stays out of the timed region
    e1.record()
    barrier()
    for tr in traces:
        a_sub = window_area(n, tr["x"].astype(np.int64), tr["y"].astype(np.int64), owned)
        # fused kernel minimum: every image read once (4*C*A) + PSF read and residual
        # write-back inside the window (8*C*A_sub). SURVEY U3 = 12*C*A_sub + 4*C*A_pk.
        # (A = this rank's row band, A_sub = the window clipped to it.)
        alg_bytes += float((4.0 * c * n * rows + 8.0 * c * a_sub).sum())
        u3_bytes += float((4.0 * c * n * rows + 12.0 * c * a_sub).sum())
    dev_ms = e0.elapsed_time(e1)
    launches = alg.kernel_launch_count - launches0
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([dev_ms, float(total_iters)], dtype=torch.float64, device=dev)
    if world > 1:
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        tsum = t.clone()
        dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
        # one sharded cube: every rank counts the same (joined) minor iterations
        dev_ms_max, iters_all = float(tmax[0]), float(total_iters)
    else:
        dev_ms_max, iters_all = dev_ms, float(total_iters)
    value = iters_all / (dev_ms_max * 1e-3)

    # ---- end to end through the C ABI with HOST buffers (H2D + D2H inside the call) ---------
    h_res = torch.empty((c, rows, n), dtype=torch.float32).pin_memory()
    h_model = torch.zeros((c, rows, n), dtype=torch.float32).pin_memory()
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
        e2e_s = float(tmax[0])
    plane = n * n * 4
    band = n * rows * 4
    # one GPU: the model planes stay on the host (the call returns a component log and applies it
    # there); row-sharded runs still move them
    model_planes = 0 if world == 1 else 1
    log_bytes = (c + 1) * 4 * iters if world == 1 else 0
    e2e = {"value": e2e_iters / e2e_s, "unit": "iterations/s",
           # the PSFs cross PCIe once per job: row-sharded ranks upload a share each and broadcast
           "h2d_bytes_per_step": (1 + model_planes) * c * band * world + c * plane,
           "d2h_bytes_per_step": (1 + model_planes) * c * band * world + log_bytes,
           "ms_per_step": e2e_s / e2e_steps * 1e3,
           "note": "rdl_execute_major_iteration with RDL_MEM_HOST pinned buffers, every step: "
                   "residual and PSFs copied to HBM, residual copied back; the model planes stay on "
                   "the host (one GPU: the call logs its components on the device and applies them "
                   "to the caller's planes, same fp32 adds; row-sharded: model planes copied both ways)"}

    def time_e2e(r_, m_, p_, steps):
        alg.set_iteration_number(0)
        alg.execute_major_iteration(cube, r_, m_, p_)  # warm-up: staging, registration, PSF upload
        barrier()
        t0_ = time.perf_counter()
        its = 0
        for _ in range(steps):
            alg.set_iteration_number(0)
            its += alg.execute_major_iteration(cube, r_, m_, p_).iteration_number
        barrier()
        tt = torch.tensor([time.perf_counter() - t0_], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        return its / float(tt[0]), float(tt[0]) / steps * 1e3

    # the PSFs of a run do not change between major iterations: rdl_set_psfs_constant keeps the
    # staged copies (the adapter of INTEGRATION.md switches it on)
    alg.set_psfs_constant(True)
    v_cached, ms_cached = time_e2e(np_res, np_model, np_psf, e2e_steps)
    e2e["psfs_cached"] = {"value": v_cached, "ms_per_step": ms_cached,
                          "h2d_bytes_per_step": (1 + model_planes) * c * band * world,
                          "note": "same call with rdl_set_psfs_constant(1): the PSF planes cross "
                                  "PCIe on the first call only"}
    # what the C++ adapter really passes: pageable aocommon::Image buffers
    del h_res, h_model, np_res, np_model
    pg_res, pg_model = res0.cpu().numpy().copy(), np.zeros((c, rows, n), np.float32)
    pg_psf = np_psf.copy()
    del h_psf, np_psf
    v_page, ms_page = time_e2e(pg_res, pg_model, pg_psf, 1)
    alg.set_pin_host_buffers(True)
    v_reg, ms_reg = time_e2e(pg_res, pg_model, pg_psf, e2e_steps)
    alg.set_pin_host_buffers(False)
    alg.set_psfs_constant(False)
    e2e["pageable"] = {"value": v_page, "ms_per_step": ms_page,
                       "note": "pageable numpy buffers (what an aocommon::Image is), PSFs cached"}
    e2e["pageable_registered"] = {"value": v_reg, "ms_per_step": ms_reg,
                                  "note": "the same pageable buffers with rdl_set_pin_host_buffers(1)"
                                          ": cudaHostRegister on first use, PSFs cached"}
    np_psf = pg_psf

    if rank == 0:
        peak, peak_src = measured_peaks()
        # sharded runs launch the step kernel + the joined control kernel per minor iteration; the
        # roofline is the step kernel's (its time here includes the exchange it waits for)
        step_launches = launches // 2 if world > 1 else launches
        avg_launch_s = dev_ms * 1e-3 / max(step_launches, 1)
        achieved = alg_bytes / max(step_launches, 1) / avg_launch_s / 1e9
        line = base_line(args, value, dev_ms_max / args.steps)
        line.update({
            "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
            "roofline": {"bound": "hbm", "kernel": "HogbomStepKernel (fused SubtractImage + "
                         "channel integration + arg-max)", "achieved": achieved, "peak": peak,
                         "unit": "GB/s", "frac": achieved / peak,
                         "traffic": profile_traffic("r02_ncu_full_hogbom_8192x16.json", ["HogbomStepKernel"]),
                         "traffic_source": "dram__bytes_read.sum + dram__bytes_write.sum of one launch, "
                                           "profiles/r02_ncu_full_hogbom_8192x16.json (single GPU, 8192^2 x 16)",
                         "peak_source": peak_src,
                         "bytes_per_launch": alg_bytes / max(step_launches, 1),
                         "avg_launch_us": avg_launch_s * 1e6,
                         "achieved_survey_u3": u3_bytes / max(step_launches, 1) / avg_launch_s / 1e9,
                         "note": "achieved uses the fused-kernel minimum bytes (each image read "
                                 "once); achieved_survey_u3 uses SURVEY 8(d) U3 = U1 + U2"},
        })
        if not args.no_cpu and world == 1:
            h = res0.cpu().numpy()
            sl = cpu_hogbom_slope(h, np_psf, weights, 2, 2 + args.cpu_iters)
            line["cpu_baseline"] = {
                "value": sl["rate"], "unit": "iterations/s", "cores": sl["cores"], "kind": cpu_kind(),
                "sample": f"slope between runs of {sl['it_a']} and {sl['it_b']} minor iterations of the "
                          f"same {n}^2 x {c} cube ({sl['t_a']:.1f} s and {sl['t_b']:.1f} s; GenericClean "
                          "path B, the reference's sources, all host cores)",
                "lump_value": sl["lump_rate"],
                "lump_note": "iterations / wall time of the longer run, set-up and plane copies "
                             "included (the figure round 1 reported)"}
            # parity AT THE TIMED SIZE: the same number of minor iterations on the device, images
            # compared bit for bit with what the reference's sources just produced
            p_res, p_model = res0.clone(), torch.zeros_like(res0)
            alg.apply(max_iterations=sl["it_b"], trace_capacity=sl["it_b"])
            alg.set_iteration_number(0)
            pr = alg.execute_major_iteration(cube, p_res, p_model, psfs, stream=stream)
            torch.cuda.synchronize()
            same_res = bool(np.array_equal(p_res.cpu().numpy(), sl["residual"]))
            same_model = bool(np.array_equal(p_model.cpu().numpy(), sl["model"]))
            line["parity"] = {"checked": int(pr.iteration_number),
                              "identical": same_res and same_model and pr.iteration_number == sl["it_b"],
                              "what": f"residual and model cubes ({n}^2 x {c}) after {sl['it_b']} minor "
                                      "iterations, device vs the reference's own sources on the host: "
                                      "np.array_equal"}
            alg.apply(max_iterations=iters, trace_capacity=iters)
            del p_res, p_model, sl, h
        if world == 1 and not args.no_ms:
            # second half of BASELINE.json's metric: multiscale major-iteration ms on the K4 cube,
            # with its own roofline (the FFT convolution is 90+ % of it), end-to-end figure and CPU arm
            del res0
            torch.cuda.empty_cache()
            ms_res0, ms_psfs, ms_w, ms_f = make_ms_cube(n, c, dev)
            ms_cube = rdl.Cube(n, n, c, 1, ms_w)
            ms_alg = ms_algorithm(rdl, local, c, ms_f, ms_w, args.ms_iters)
            ms_total, ms_r, ms_minor, ms_launches = run_multiscale_device(
                rdl, ms_alg, ms_cube, ms_res0, ms_psfs, 2, 1, barrier, stream)
            conv_counts = ms_alg.reference_convolutions()
            n_pad = int(rdl.convolution_size(ms_alg.scale_infos()[-1]["scale"], n, 1.1))
            ms = {
                "ms_per_major_iteration": ms_total / 2, "max_iterations": args.ms_iters,
                "components_cleaned": int(ms_r.iteration_number),
                "minor_iterations": ms_minor / 2, "gpu_launches": int(ms_launches),
                "config": f"multiscale {n}^2 x {c} channels joined, 5 scales, polynomial fit 2 "
                          f"terms, mgain 0.8, at most {args.ms_iters} components per major iteration; "
                          "device-resident, CUDA events",
                "roofline": fft_roofline(rdl, n_pad, dev)}
            # the natural major iteration: no component cap, stops on mgain
            un_alg = ms_algorithm(rdl, local, c, ms_f, ms_w, 10 ** 9)
            un_total, un_r, un_minor, _ = run_multiscale_device(rdl, un_alg, ms_cube, ms_res0, ms_psfs,
                                                                1, 1, barrier, stream)
            ms["uncapped"] = {"ms_per_major_iteration": un_total,
                              "components_cleaned": int(un_r.iteration_number),
                              "minor_iterations": un_minor}
            del un_alg
            # end to end through host buffers (pinned; PSFs uploaded on every call)
            ms["e2e"] = multiscale_e2e(ms_alg, ms_cube, ms_res0, ms_psfs, barrier, 1)
            # major iterations 2..n of a run: the PSFs are the ones of the first (radler.cc:175-268),
            # so with rdl_set_psfs_constant(1) what the call derives from them alone is kept
            ms_alg.set_psfs_constant(True)
            k_total, k_r, _, _ = run_multiscale_device(rdl, ms_alg, ms_cube, ms_res0, ms_psfs, 2, 1,
                                                       barrier, stream)
            k_e2e = multiscale_e2e(ms_alg, ms_cube, ms_res0, ms_psfs, barrier, 1)
            ms_alg.set_psfs_constant(False)
            ms["psfs_constant"] = {
                "ms_per_major_iteration": k_total / 2, "e2e_ms_per_step": k_e2e["ms_per_step"],
                "components_cleaned": int(k_r.iteration_number),
                "e2e_h2d_bytes_per_step": 2 * c * n * n * 4,
                "note": "the same major iteration from the second call of a run on, with "
                        "rdl_set_psfs_constant(1): scale-convolved / twice-convolved PSFs, kernel "
                        "spectra and PSF peaks kept from the previous call, PSF planes not uploaded "
                        "again; results bit-identical (tests/test_gpu_multiscale.py)"}
            if not args.no_cpu:
                cores = os.cpu_count() or 1
                own, mkl, best, per_size = price_reference_convolutions(conv_counts, cores)
                ms["cpu_baseline"] = {
                    "value": best * 1e3, "unit": "ms per major iteration", "cores": cores,
                    "kind": "port",
                    "sample": "the convolutions the reference makes for this major iteration ("
                              "rdl_reference_convolutions: count by frame size, equal to the "
                              "reference's own count in tests/test_gpu_parity_surface.py) x the "
                              "per-convolution time measured here (2 convolutions per size). A LOWER "
                              "bound on the CPU major iteration: everything that is not an FFT "
                              "convolution (7-8 % of the reference's run) is left out",
                    "with_reference_fft_ms": own * 1e3, "with_mkl_fft_ms": mkl * 1e3,
                    "fft": "faster of the two per size (BASELINE.md 2.3)", "per_size": per_size,
                    "speedup_device": best * 1e3 / (ms_total / 2),
                    "speedup_e2e": best * 1e3 / ms["e2e"]["ms_per_step"]}
            line["multiscale"] = ms
            try:
                line["perform_steps"] = perform_steps(rdl, ms_alg, ms_cube, ms_res0, ms_f, ms_w)
            except Exception as e:  # informational: never costs the line
                line["perform_steps"] = {"error": str(e)}
            del ms_res0, ms_psfs, ms_alg
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def perform_steps(rdl, alg, cube, planes, freqs, weights, n_dec=4):
    """What Radler::Perform does either side of ExecuteMajorIteration, on the
    device-resident cube: the planes taken as original channels averaged to n_dec deconvolution
    channels, the MAD noise estimate of the integrated cube, the auto-mask from the model, the
    model interpolated back. Milliseconds per step (CUDA events on torch's stream, which the
    handle's own blocking stream orders with; the MAD and the interpolation synchronise)."""
    import torch
    c, n, _ = planes.shape
    dev = planes.device

    def timed(fn, reps=3):
        fn()
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(reps):
            fn()
        b.record()
        torch.cuda.synchronize()
        return a.elapsed_time(b) / reps

    index = [ch * n_dec // c for ch in range(c)]
    w64 = [float(w) for w in (weights if weights is not None else [1.0] * c)]
    averaged = torch.zeros((n_dec, n, n), device=dev)
    integrated = torch.zeros((n, n), device=dev)
    mask = torch.zeros((n, n), dtype=torch.uint8, device=dev)
    out = {"frame": n, "original_channels": c, "deconvolution_channels": n_dec}
    out["load_and_average_ms"] = timed(lambda: alg.load_and_average(list(planes), w64, index,
                                                                    list(averaged)))
    out["integrate_ms"] = timed(lambda: alg.integrate(cube, planes, integrated))
    out["median_and_stddev_from_mad_ms"] = timed(lambda: alg.median_and_stddev_from_mad(integrated))
    out["stddev"] = alg.median_and_stddev_from_mad(integrated)[1]
    out["auto_mask_from_model_ms"] = timed(lambda: alg.auto_mask_from_model(list(averaged), mask))
    fit = [sum(freqs[ch] for ch in range(c) if index[ch] == d) / index.count(d)
           for d in range(n_dec)]
    alg.set_spectral_fitter(rdl.FIT_POLYNOMIAL, 2, fit, None)
    outputs = [torch.empty((n, n), device=dev) for _ in range(c)]
    out["interpolate_model_ms"] = timed(lambda: alg.interpolate_model(list(averaged), list(freqs),
                                                                      outputs))
    out["note"] = ("device planes only; parity of each step and of the whole sequence against "
                   "radler::Radler::Perform of the reference's own sources: "
                   "tests/test_gpu_perform_steps.py")
    return out


# --------------------------------------------------------------------------------------
# multiscale workload (BASELINE.json: "multiscale major-iter ms at 8192^2 x 16ch")
# --------------------------------------------------------------------------------------
def ms_algorithm(rdl, local, c, freqs, weights, iters):
    alg = rdl.MultiScaleAlgorithm(6.0, 1.0, max_scales=5, device=local)
    alg.apply(max_iterations=iters, clean_border_ratio=0.0, minor_loop_gain=0.1,
              major_loop_gain=0.8, threshold=0.0, trace_capacity=0)
    if c > 1:
        alg.set_spectral_fitter(rdl.FIT_POLYNOMIAL, 2, freqs, weights, 1)
    return alg


def fft_roofline(rdl, n_pad, dev):
    """The dominant kernel group of a multiscale major iteration, timed alone: one circular FFT
    convolution of an n_pad^2 frame (RowForward, ColumnStrided, ColumnBlock, ColumnStrided,
    RowInverse). Algorithmic bytes = SURVEY U5 = 28 B per padded pixel."""
    import torch
    alg = rdl.GenericClean(device=dev.index)
    side = torch.cuda.Stream(device=dev)
    n_img, reps = 4, 3
    with torch.cuda.stream(side):
        imgs = torch.randn((n_img, n_pad, n_pad), device=dev)
        ker = torch.randn((n_pad, n_pad), device=dev)
        alg.convolve(imgs, ker, stream=side.cuda_stream)
        side.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(side)
        for _ in range(reps):
            alg.convolve(imgs, ker, stream=side.cuda_stream)
        e1.record(side)
        side.synchronize()
    ms = e0.elapsed_time(e1) / reps
    per_conv_s = ms * 1e-3 / (n_img + 16.0 / 28.0)  # + one kernel spectrum (16 B/px) per call
    peak, src = measured_peaks()
    bytes_ = 28.0 * n_pad * n_pad
    t8192 = profile_traffic("r02_ncu_full_fft_8192.json")
    return {"bound": "hbm", "kernel": f"FFT convolution {n_pad}^2 (FastRowForward + FastColumnStrided + "
            "FastColumnBlock + FastColumnStrided + FastRowInverse), timed alone on its stream",
            "achieved": bytes_ / per_conv_s / 1e9, "peak": peak, "unit": "GB/s",
            "frac": bytes_ / per_conv_s / 1e9 / peak,
            "traffic": None if t8192 is None else t8192 * (n_pad / 8192.0) ** 2,
            "traffic_note": "ncu dram bytes of the five kernels of one 8192^2 convolution (profiles/"
                            "r02_ncu_full_fft_8192.json: 2.72 GB = 40.6 B/px; the four-step column "
                            "transform moves 44 B/px by design) scaled by the pixel count; SURVEY U5 "
                            "counts 28 B/px",
            "peak_source": src, "bytes_per_launch": bytes_ / 5.0,
            "avg_launch_us": per_conv_s * 1e6 / 5.0, "us_per_convolution": per_conv_s * 1e6}


STEP_MS = []  # this rank's time of each timed step of the last run_multiscale_device


def run_multiscale_device(rdl, alg, cube, res0, psfs, steps, warmup, barrier, stream):
    """Device-resident timing of `steps` major iterations; returns (ms_total, last result,
    minor iterations, launches)."""
    import torch
    res = torch.empty_like(res0)
    model = torch.empty_like(res0)

    def one():
        res.copy_(res0)
        model.zero_()
        alg.set_iteration_number(0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        r = alg.execute_major_iteration(cube, res, model, psfs, stream=stream)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1), r

    for _ in range(warmup):
        one()
    barrier()
    l0 = alg.kernel_launch_count
    total, minor, r = 0.0, 0, None
    STEP_MS.clear()
    for _ in range(steps):
        ms, r = one()
        STEP_MS.append(ms)
        total += ms
        minor += alg.minor_iteration_count
    barrier()
    del res, model
    return total, r, minor, alg.kernel_launch_count - l0


def multiscale_e2e(alg, cube, res0, psfs, barrier, steps):
    """The same major iteration through rdl_execute_major_iteration with pinned HOST buffers:
    residual, model and PSFs copied to HBM, residual and model copied back, inside the timer."""
    import torch
    c, n = res0.shape[0], res0.shape[-1]
    h_res = torch.empty((c, n, n), dtype=torch.float32).pin_memory()
    h_model = torch.zeros((c, n, n), dtype=torch.float32).pin_memory()
    h_psf = torch.empty((c, n, n), dtype=torch.float32).pin_memory()
    h_psf.copy_(psfs)
    total = 0.0
    for i in range(1 + steps):
        h_res.copy_(res0)
        h_model.zero_()
        alg.set_iteration_number(0)
        barrier()
        t0 = time.perf_counter()
        alg.execute_major_iteration(cube, h_res.numpy(), h_model.numpy(), h_psf.numpy())
        barrier()
        if i > 0:  # the first call warms the staging buffers
            total += time.perf_counter() - t0
    plane = n * n * 4
    return {"ms_per_step": total / steps * 1e3, "value": total / steps * 1e3,
            "unit": "ms per major iteration", "h2d_bytes_per_step": 3 * c * plane,
            "d2h_bytes_per_step": 2 * c * plane}


def run_ours_multiscale(args):
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
    n, c, iters = args.size, args.channels, args.ms_iters
    # N > 1: ONE cube (every rank builds the same one), its CHANNELS sharded over the ranks: strong
    # scaling. Every rank convolves its own channels and the joined steps run replicated.
    res0, psfs, weights, freqs = make_ms_cube(n, c, dev, seed=1)
    cube = rdl.Cube(n, n, c, 1, weights)
    alg = ms_algorithm(rdl, local, c, freqs, weights, iters)
    if world > 1:
        from ska_sdp_func_radler_b200 import distributed as rd
        alg.dist_init(rd.share_unique_id(rdl.dist_unique_id, rank, world, device=dev), rank, world)
    stream = torch.cuda.current_stream().cuda_stream

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local)
    # warm-up happens inside; the sampler also covers it (clocks are read as a median)
    if rank == 0:
        sampler.start()
    total_ms, r, minor, launches = run_multiscale_device(rdl, alg, cube, res0, psfs, args.steps,
                                                         args.warmup, barrier, stream)
    clocks = sampler.stop() if rank == 0 else None
    t = torch.tensor([total_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_per_step = float(t[0]) / args.steps
    step_ms = list(STEP_MS)

    # ---- e2e: the same call with pinned HOST buffers --------------------------------------
    h_res = torch.empty((c, n, n), dtype=torch.float32).pin_memory()
    h_model = torch.zeros((c, n, n), dtype=torch.float32).pin_memory()
    h_psf = torch.empty((c, n, n), dtype=torch.float32).pin_memory()
    h_psf.copy_(psfs)
    e2e_steps = max(1, min(args.steps, 2))
    e2e_s = 0.0
    for i in range(1 + e2e_steps):
        h_res.copy_(res0)
        h_model.zero_()
        alg.set_iteration_number(0)
        barrier()
        t0 = time.perf_counter()
        alg.execute_major_iteration(cube, h_res.numpy(), h_model.numpy(), h_psf.numpy())
        barrier()
        if i > 0:  # first call warms the staging buffers
            e2e_s += time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_ms = float(t[0]) / e2e_steps * 1e3
    plane = n * n * 4

    if rank == 0:
        n_pad = int(rdl.convolution_size(alg.scale_infos()[-1]["scale"], n, 1.1))
        line = {
            "metric": "multiscale major-iteration time (MultiScaleAlgorithm::ExecuteMajorIteration)",
            "value": ms_per_step, "unit": "ms per major iteration",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_per_step, "step_ms_rank0": step_ms, "higher_is_better": False,
            "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"multiscale {n}^2 x {c} channels joined, 5 scales, polynomial "
                                   f"fit 2 terms, mgain 0.8, max {iters} components per major "
                                   "iteration (step)", "size": n, "channels": c,
                       "max_iterations": iters, "components_cleaned": int(r.iteration_number),
                       "minor_iterations_per_step": minor / args.steps,
                       "l2": "inputs larger than L2" if n * n * c * 4 > 4 * 126e6 else
                       "cube comparable to L2",
                       "parallelism": f"channels of one cube sharded over {world} GPUs (contiguous "
                                      "blocks; per-channel FFT convolutions local, joined steps "
                                      "replicated, planes broadcast by their owners over NCCL)"
                       if world > 1 else "single GPU"},
            "e2e": {"value": e2e_ms, "unit": "ms per major iteration",
                    "ms_per_step": e2e_ms, "h2d_bytes_per_step": 3 * c * plane,
                    "d2h_bytes_per_step": 2 * c * plane},
            "gpu_launches": int(launches), "clocks": clocks,
            "roofline": fft_roofline(rdl, n_pad, dev),
        }
        if not args.no_cpu:
            # identical call on both sides: max_iterations = cpu_iters (setup + first minor pass)
            a2 = ms_algorithm(rdl, local, c, freqs, weights, args.cpu_iters)
            ms_same, r2, _, _ = run_multiscale_device(rdl, a2, cube, res0, psfs, 1, 1, barrier,
                                                      stream)
            del a2
            dt, cores, it, kind = cpu_multiscale(res0.cpu().numpy(), h_psf.numpy(), weights, freqs,
                                                 args.cpu_iters)
            line["cpu_baseline"] = {
                "value": dt * 1e3, "unit": "ms", "cores": cores, "kind": kind,
                "sample": f"the same {n}^2 x {c} major iteration capped at max_iterations="
                          f"{args.cpu_iters} (scale setup + PSF convolutions + first minor pass; "
                          f"{it} components)",
                "gpu_same_sample_ms": ms_same, "gpu_same_sample_components": int(r2.iteration_number)}
        print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


def run_reference_multiscale(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n, c = args.size, args.channels
    res, psfs, weights, freqs = make_ms_cube(n, c, "cpu")
    res, psfs = res.numpy(), psfs.numpy()
    times = []
    steps = max(1, min(args.steps, 2))
    for _ in range(steps):
        dt, cores, it, kind = cpu_multiscale(res.copy(), psfs, weights, freqs, args.cpu_iters)
        times.append(dt * 1e3)
    value = float(np.mean(times))
    print(json.dumps({
        "impl": "reference",
        "metric": "multiscale major-iteration time (MultiScaleAlgorithm::ExecuteMajorIteration)",
        "value": value, "unit": "ms per major iteration (job average)", "n_gpus": args.gpus,
        "steps": steps, "warmup": 0, "ms_per_step": value, "higher_is_better": False,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"multiscale {n}^2 x {c} channels joined, 5 scales, capped at "
                               f"max_iterations={args.cpu_iters} (bounded sample)", "size": n,
                   "channels": c, "max_iterations": args.cpu_iters},
        "cpu_baseline": {"value": value, "unit": "ms", "cores": cores, "kind": kind,
                         "sample": f"max_iterations={args.cpu_iters}, {it} components"},
        "e2e": {"value": value, "unit": "ms per major iteration (job average)",
                "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


def main():
    # stdout carries the one JSON line: NCCL's version banner / debug output goes to stderr
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="hogbom", choices=["hogbom", "multiscale"])
    ap.add_argument("--ms-iters", type=int, default=500,
                    help="multiscale: max components per major iteration")
    ap.add_argument("--size", type=int, default=8192)
    ap.add_argument("--channels", type=int, default=16)
    ap.add_argument("--iters", type=int, default=256, help="minor iterations per step")
    ap.add_argument("--cpu-iters", type=int, default=None,
                    help="CPU sample: minor iterations (hogbom, default 6) / max_iterations "
                         "(multiscale, default 1)")
    ap.add_argument("--no-cpu", action="store_true")
    ap.add_argument("--no-ms", action="store_true",
                    help="hogbom: skip the extra multiscale measurement")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3 if args.impl == "ours" else args.warmup
    if args.cpu_iters is None:
        args.cpu_iters = 6 if args.workload == "hogbom" else 1
    if args.workload == "multiscale":
        (run_reference_multiscale if args.impl == "reference" else run_ours_multiscale)(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
