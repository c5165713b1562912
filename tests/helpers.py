"""Synthetic inputs shared by the parity tests, bench.py and smoke()."""
import numpy as np


def gaussian_psf(n, fwhm=4.0, sidelobe=0.01, h=None):
    """Circular Gaussian, peak exactly 1 at (w/2, h/2), plus a small damped sidelobe ring;
    values below 1e-7 are truncated to 0 so the PSF is compact."""
    h = n if h is None else h
    y, x = np.mgrid[0:h, 0:n].astype(np.float64)
    r = np.hypot(x - n // 2, y - h // 2)
    sigma = fwhm / 2.3548200450309493
    p = np.exp(-0.5 * (r / sigma) ** 2) + sidelobe * np.cos(r / 8.0) * np.exp(-r / 200.0) * (r > 0)
    p[np.abs(p) < 1e-7] = 0.0
    p[h // 2, n // 2] = 1.0
    return p.astype(np.float32)


def fft_convolve_same(sky, psf):
    """sky (*) psf with the PSF centre at (w/2, h/2); zero-padded (no wrap), double precision."""
    h, w = sky.shape
    ph, pw = 2 * h, 2 * w
    a = np.zeros((ph, pw))
    a[:h, :w] = sky
    b = np.zeros((ph, pw))
    b[:h, :w] = psf
    c = np.fft.irfft2(np.fft.rfft2(a) * np.fft.rfft2(b), s=(ph, pw))
    return c[h // 2:h // 2 + h, w // 2:w // 2 + w]


def point_source_sky(n, n_src, seed=1, margin=16, flux=(0.01, 10.0), h=None):
    h = n if h is None else h
    rng = np.random.default_rng(seed)
    sky = np.zeros((h, n))
    xs = rng.integers(margin, n - margin, n_src)
    ys = rng.integers(margin, h - margin, n_src)
    fl = np.exp(rng.uniform(np.log(flux[0]), np.log(flux[1]), n_src))
    np.add.at(sky, (ys, xs), fl)
    return sky


def add_blobs(sky, n_blobs, seed=3, sigmas=(6, 12, 24, 48), peak=(0.05, 0.5)):
    h, w = sky.shape
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w].astype(np.float64)
    for _ in range(n_blobs):
        s = float(rng.choice(sigmas))
        s = min(s, w / 16)
        cx, cy = rng.uniform(4 * s, w - 4 * s), rng.uniform(4 * s, h - 4 * s)
        sky += rng.uniform(*peak) * np.exp(-0.5 * ((x - cx) ** 2 + (y - cy) ** 2) / s ** 2)
    return sky


def dirty_image(sky, psf, noise=1e-3, seed=6):
    rng = np.random.default_rng(seed)
    d = fft_convolve_same(sky, psf.astype(np.float64))
    d += rng.normal(0.0, noise, d.shape)
    return d.astype(np.float32)


def hogbom_case(n, n_channels=1, seed=1, n_src=None, fwhm=4.0, noise=1e-3, alpha=True):
    """K1-shaped problem: point sources (*) Gaussian PSF + noise; joined channels share the
    sky with a power-law spectrum and per-channel PSF widths."""
    n_src = max(4, n * n // 4096) if n_src is None else n_src
    sky = point_source_sky(n, n_src, seed=seed)
    rng = np.random.default_rng(seed + 100)
    freqs = np.linspace(100e6, 250e6, n_channels) if n_channels > 1 else np.array([150e6])
    res = np.zeros((n_channels, n, n), np.float32)
    psfs = np.zeros((n_channels, n, n), np.float32)
    spectral_index = rng.uniform(-1.2, -0.2) if alpha else 0.0
    for c in range(n_channels):
        psfs[c] = gaussian_psf(n, fwhm * (freqs[0] / freqs[c]) ** 0.25)
        scale = (freqs[c] / freqs[0]) ** spectral_index
        res[c] = dirty_image(sky * scale, psfs[c], noise, seed=6 + c)
    weights = rng.uniform(0.5, 1.5, n_channels).astype(np.float32) if n_channels > 1 else None
    return res, psfs, weights, freqs


def multiscale_case(n, n_channels=1, seed=1, fwhm=6.0, noise=1e-3):
    sky = point_source_sky(n, max(4, n * n // 8192), seed=seed, margin=n // 8)
    sky = add_blobs(sky, max(2, n * n // 65536), seed=seed + 2)
    rng = np.random.default_rng(seed + 100)
    freqs = np.linspace(100e6, 250e6, n_channels) if n_channels > 1 else np.array([150e6])
    res = np.zeros((n_channels, n, n), np.float32)
    psfs = np.zeros((n_channels, n, n), np.float32)
    alpha = rng.uniform(-1.2, -0.2)
    for c in range(n_channels):
        psfs[c] = gaussian_psf(n, fwhm * (freqs[0] / freqs[c]) ** 0.25)
        res[c] = dirty_image(sky * (freqs[c] / freqs[0]) ** alpha, psfs[c], noise, seed=6 + c)
    weights = rng.uniform(0.5, 1.5, n_channels).astype(np.float32) if n_channels > 1 else None
    return res, psfs, weights, freqs


def plus_psf_fixture(shift_x=0, shift_y=0, factor=2.5, n=64):
    """The 5-pixel 'plus' PSF and residual = factor * shifted PSF of cpp/test/test_radler.cc:54-82."""
    c = n // 2 * n + n // 2
    psf = np.zeros(n * n, np.float32)
    for off, v in [(0, 1.0), (-1, 0.25), (1, 0.5), (-n, 0.4), (n, 0.6)]:
        psf[c + off] = v
    sc = (n // 2 + shift_y) * n + (n // 2 + shift_x)
    res = np.zeros(n * n, np.float32)
    for off, v in [(0, 1.0), (-1, 0.25), (1, 0.5), (-n, 0.4), (n, 0.6)]:
        res[sc + off] = np.float32(v * factor)
    return psf.reshape(1, n, n), res.reshape(1, n, n), sc
