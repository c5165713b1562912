"""Inputs shared by the CPU and GPU tests of Radler::Perform's steps around the major iteration
(LoadAndAverage, MedianAndStdDevFromMAD, auto-mask from the model)."""
import numpy as np


def output_index_of(n_original, n_pol, n_deconvolution):
    """The output plane of every [original channel][pol] entry and the order the reference visits
    them in: WorkTable's grouping (work_table.cc:28-33) under LoadAndAverage's loops
    (image_set.cc:113-135). Entries are already in that order."""
    n_dec = min(n_original, n_deconvolution) if n_deconvolution else n_original
    index = []
    for ch in range(n_original):
        d = ch * n_dec // n_original
        index += [d * n_pol + p for p in range(n_pol)]
    return index, n_dec


def average_case(n_original, n_pol, h, w, seed, zero_weight_nan=True):
    rng = np.random.default_rng(seed)
    planes = rng.standard_normal((n_original * n_pol, h, w)).astype(np.float32)
    weights = rng.uniform(0.2, 3.0, n_original * n_pol)
    if zero_weight_nan and n_original > 2:
        # a flagged channel: weight zero, pixels NaN (image_set.cc:126-127)
        for p in range(n_pol):
            weights[1 * n_pol + p] = 0.0
            planes[1 * n_pol + p] = np.nan
    return planes, weights


def average_numpy(planes, weights, index, n_outputs):
    """fma accumulation in float32 is not expressible in numpy; float64 accumulate is the
    tolerance check (the bit check is against the reference's own image_set.cc)."""
    out = np.zeros((n_outputs,) + planes.shape[1:], np.float64)
    sums = np.zeros(n_outputs)
    for i, o in enumerate(index):
        if weights[i] != 0.0:
            out[o] += planes[i].astype(np.float64) * np.float32(weights[i])
            sums[o] += weights[i]
    with np.errstate(all="ignore"):
        return out * np.float32(1.0 / sums)[:, None, None]


def mad_numpy(a):
    v = a[np.isfinite(a)].astype(np.float32)
    if v.size == 0:
        return np.float32(0), np.float32(0)

    def median(x):
        x = np.sort(x)
        mid = x.size // 2
        return x[mid] if x.size % 2 else np.float32((x[mid - 1] + x[mid]) * np.float32(0.5))

    m = median(v)
    return np.float32(m), np.float32(np.float32(1.4826) * median(np.abs(v - m)))


def noise_images():
    """name -> image: odd / even counts, NaN and Inf pixels, ties, constants, one pixel."""
    rng = np.random.default_rng(17)
    cases = {}
    cases["odd"] = rng.standard_normal((33, 35)).astype(np.float32)
    cases["even"] = rng.standard_normal((64, 96)).astype(np.float32)
    a = rng.standard_normal((100, 130)).astype(np.float32) * 1e-3 + 0.25
    a[::7, ::5] = np.nan
    a[3, 4] = np.inf
    a[5, 6] = -np.inf
    cases["nan_inf"] = a
    cases["ties"] = rng.integers(-3, 4, (50, 60)).astype(np.float32)
    cases["constant"] = np.full((16, 16), 2.5, np.float32)
    cases["one"] = np.array([[-4.0]], np.float32)
    cases["two"] = np.array([[1.0, 4.0]], np.float32)
    cases["all_nan"] = np.full((4, 5), np.nan, np.float32)
    s = rng.standard_normal((257, 509)).astype(np.float32)
    s[100:110, 200:220] += 80.0  # a bright source on the noise
    cases["source"] = s
    cases["wide_range"] = (rng.standard_normal((64, 64)) * 10.0 ** rng.uniform(-30, 30, (64, 64))
                           ).astype(np.float32)
    return cases


def interpolate_case(n_dec, n_orig, h, w, seed):
    """Model planes of a few components per deconvolution channel, zero elsewhere; frequencies
    of the deconvolution channels (band centres of groups of original channels) and weights."""
    rng = np.random.default_rng(seed)
    orig = 1.2e8 + 2.0e6 * np.arange(n_orig) + rng.uniform(-2e5, 2e5, n_orig)
    groups = [[ch for ch in range(n_orig) if ch * n_dec // n_orig == d] for d in range(n_dec)]
    fit_freqs = np.array([orig[g].mean() for g in groups])
    fit_weights = rng.uniform(0.5, 2.0, n_dec).astype(np.float32)
    models = np.zeros((n_dec, h, w), np.float32)
    ys, xs = rng.integers(0, h, 40), rng.integers(0, w, 40)
    amp = rng.standard_normal(40)
    for d in range(n_dec):
        models[d, ys, xs] = (amp * (fit_freqs[d] / fit_freqs[0]) ** -0.7).astype(np.float32)
    models[0, 0, 0] = 0.5  # a component in one channel only
    models[-1, 1, 1] = -0.0
    return models, fit_freqs, fit_weights, orig


def interpolate_numpy(models, fit_freqs, fit_weights, n_terms, out_freqs):
    """Weighted polynomial least squares in (nu / nu_ref - 1), nu_ref = the middle fit channel,
    evaluated at out_freqs: float64 throughout (tolerance check)."""
    used = fit_weights > 0
    t = min(n_terms, int(used.sum()))
    ref = fit_freqs[len(fit_freqs) // 2]
    x = fit_freqs / ref - 1.0
    a = np.vander(x, t, increasing=True)
    sw = np.sqrt(np.where(used, fit_weights, 0.0).astype(np.float64))
    solve = np.linalg.pinv(a * sw[:, None]) * sw[None, :]  # t x n
    terms = np.tensordot(solve, models.astype(np.float64), axes=(1, 0))
    xo = out_freqs / ref - 1.0
    return np.tensordot(np.vander(xo, t, increasing=True), terms, axes=(1, 0))


def deconvolution_channels(frequencies, weights, n_deconvolution):
    """ImageSet::CalculateDeconvolutionFrequencies (image_set.cc:464-497): per deconvolution
    channel the weighted mean frequency (double) and the summed weight (accumulated in float)."""
    n = len(frequencies)
    n_dec = min(n, n_deconvolution) if n_deconvolution else n
    freqs, sums = np.zeros(n_dec), np.zeros(n_dec, np.float32)
    plain, counts = np.zeros(n_dec), np.zeros(n_dec, int)
    for i in range(n):
        d = i * n_dec // n
        freqs[d] += frequencies[i] * weights[i]
        sums[d] = np.float32(np.float64(sums[d]) + weights[i])
        plain[d] += frequencies[i]
        counts[d] += 1
    for d in range(n_dec):
        freqs[d] = freqs[d] / np.float64(sums[d]) if sums[d] > 0 else plain[d] / counts[d]
    return freqs, sums
