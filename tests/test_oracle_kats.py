"""Pins the CPU oracle against the reference's own known-answer tests
and against oracle/_ref (the reference's leaf sources compiled where they lie)."""
import math

import numpy as np
import pytest

from helpers import plus_psf_fixture


# ---- cpp/math/test/test_peak_finder.cc:48-184 ---------------------------------------------
PEAK_KATS = [
    # (n, {index: value}, width, height, expected x, y)
    (16, {0: 1}, 4, 2, 0, 0),
    (16, {0: 1, 1: 2}, 4, 2, 1, 0),
    (16, {0: 1, 1: 2, 4: 3}, 4, 2, 0, 1),
    (16, {0: 1, 1: 2, 4: 3, 7: 4}, 4, 2, 3, 1),
    (16, {0: 1, 1: 2, 4: 3, 7: 4, 15: 6}, 4, 4, 3, 3),
    (16, {0: 1, 1: 2, 4: 3, 7: 4, 15: 6, 14: 5}, 3, 5, 2, 4),
    (36, {0: 1}, 6, 6, 0, 0),
    (36, {0: 1, 1: 2}, 6, 6, 1, 0),
    (36, {0: 1, 1: 2, 6: 3}, 6, 6, 0, 1),
    (36, {0: 1, 1: 2, 6: 3, 9: 4}, 6, 6, 3, 1),
    (36, {0: 1, 1: 2, 6: 3, 9: 4, 35: 6}, 6, 6, 5, 5),
    (36, {0: 1, 1: 2, 6: 3, 9: 4, 35: 6}, 2, 18, 1, 17),
    (38, {0: 1, 1: 2, 6: 3, 9: 4, 35: 6, 37: 7}, 6, 6, 5, 5),  # value past the end is ignored
]


@pytest.mark.parametrize("n,vals,w,h,ex,ey", PEAK_KATS)
def test_peak_finder_kats(oracle, n, vals, w, h, ex, ey):
    img = np.zeros(n, np.float32)
    for i, v in vals.items():
        img[i] = v
    v, x, y = oracle.find_peak(img[:w * h].reshape(h, w), allow_negative=True)
    assert (x, y) == (ex, ey)
    assert v == img[ey * w + ex]


def test_peak_finder_semantics(oracle):
    # FLT_MIN start value, strict '>', first index wins, NaN never wins (peak_finder.cc:19-56)
    img = np.zeros((4, 4), np.float32)
    assert oracle.find_peak(img)[0] is None
    img[1, 2] = img[3, 0] = 5.0
    assert oracle.find_peak(img)[1:] == (2, 1)
    img[0, 0] = np.nan
    assert oracle.find_peak(img)[1:] == (2, 1)
    img[2, 2] = -7.0
    assert oracle.find_peak(img, allow_negative=True) == (-7.0, 2, 2)
    assert oracle.find_peak(img, allow_negative=False) == (5.0, 2, 1)
    # borders / row range
    assert oracle.find_peak(img, h_border=1, v_border=1)[1:] == (2, 2)
    assert oracle.find_peak(img, start_y=3, end_y=4) == (5.0, 0, 3)
    mask = np.zeros((4, 4), np.uint8)
    mask[3, 0] = 1
    assert oracle.find_peak(img, mask=mask) == (5.0, 0, 3)


def test_peak_finder_matches_reference_source(oracle):
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    rng = np.random.default_rng(7)
    for _ in range(300):
        h, w = rng.integers(1, 70, 2)
        img = rng.standard_normal((h, w)).astype(np.float32)
        if rng.integers(0, 4) == 0:
            img[rng.integers(0, h), rng.integers(0, w)] = np.nan
        an = bool(rng.integers(0, 2))
        # borders the reference code can take without reading out of bounds
        hb, vb = int(rng.integers(0, 4)), int(rng.integers(0, 4))
        hb, vb = min(hb, (w - 1) // 2), min(vb, (h - 1) // 2)
        ys, ye = sorted(int(v) for v in rng.integers(0, h + 1, 2))
        got = oracle.find_peak(img, an, ys, ye, hb, vb)
        assert got == oracle.ref_find_peak(img, 0, an, ys, ye, hb, vb)
        mask = rng.integers(0, 2, (h, w)).astype(np.uint8)
        assert oracle.find_peak(img, an, ys, ye, hb, vb, mask) == \
            oracle.ref_find_peak(img, 2, an, ys, ye, hb, vb, mask)
        if 2 * hb < w and 2 * vb < h and ye > ys:
            ref_avx = oracle.ref_find_peak(img, 1, an, ys, ye, hb, vb)
            if got[0] is not None:  # Avx differs only in the "nothing found" return
                assert got == ref_avx


def test_subtract_image_matches_reference_source(oracle):
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built (needs /root/reference)")
    rng = np.random.default_rng(8)
    for _ in range(200):
        h, w = rng.integers(1, 50, 2)
        img = rng.standard_normal((h, w)).astype(np.float32)
        psf = rng.standard_normal((h, w)).astype(np.float32)
        x, y = int(rng.integers(0, w)), int(rng.integers(0, h))
        f = float(np.float32(rng.standard_normal()))
        a, b = img.copy(), img.copy()
        oracle.subtract_image(a, psf, x, y, f)           # threaded PartialSubtractImage
        oracle.ref_partial_subtract_image(b, psf, x, y, f, 0, h)
        assert np.array_equal(a, b)
        a, b = img.copy(), img.copy()
        oracle.subtract_image_whole(a, psf, x, y, f)     # drops the odd column (quirk A7)
        oracle.ref_subtract_image(b, psf, x, y, f)
        assert np.array_equal(a, b)
        ph, pw = int(rng.integers(1, h + 1)), int(rng.integers(1, w + 1))
        small = np.ascontiguousarray(psf[:ph, :pw])
        ys, ye = sorted(int(v) for v in rng.integers(0, h + 1, 2))
        if x + pw // 2 >= pw // 2 and x >= 0:
            a, b = img.copy(), img.copy()
            oracle.partial_subtract_image(a, small, x, y, f, ys, ye)
            oracle.ref_partial_subtract_image(b, small, x, y, f, ys, ye)
            assert np.array_equal(a, b)


# ---- cpp/utils/test/test_fft_size_calculations.cc:9-50 ------------------------------------
def test_fft_size_kats(oracle):
    for n in range(1154, 1177):
        assert oracle.good_fft_size(n) == 1176
    assert oracle.good_fft_size(1) == 2 and oracle.good_fft_size(2) == 2
    assert oracle.convolution_size(10, 1010, 1.1) == 1134
    assert oracle.convolution_size(0, 1010, 1.1) == 1120
    assert oracle.convolution_size(0, 1154, 1.0) == 1176
    # sizes the bench configs hit
    assert [oracle.convolution_size(s, 4096, 1.1) for s in (0, 24, 48, 96, 192)] == \
        [4536, 4608, 4608, 4704, 4860]
    assert [oracle.convolution_size(s, 8192, 1.1) for s in (0, 24, 48, 96, 192)] == \
        [9072, 9072, 9216, 9216, 9408]
    if oracle.ref_lib() is not None:
        rng = np.random.default_rng(3)
        for n in rng.integers(1, 20000, 200):
            assert oracle.good_fft_size(int(n)) == oracle.ref_lib().ref_good_fft_size(int(n))
        for _ in range(100):
            s, n, p = rng.uniform(0, 300), int(rng.integers(8, 20000)), rng.uniform(1.0, 2.0)
            assert oracle.convolution_size(s, n, p) == oracle.ref_lib().ref_convolution_size(s, n, p)


# ---- cpp/test/test_image_set.cc:123-386 (integration values) ------------------------------
def _cube(oracle, n_channels, n_pol, weights=None, linked=None, pol_norm=1.0, squared=False):
    return oracle.Cube(2, 2, n_channels, n_pol, None if weights is None else np.float32(weights),
                       None if linked is None else np.uint8(linked), pol_norm, squared)


def _check(oracle, cube, pixel_values, index, linear, squared):
    imgs = np.zeros((cube.n_channels * cube.n_pol, 2, 2), np.float32)
    for i, v in enumerate(pixel_values):
        imgs[i].flat[index] = v
    if linear is not None:
        assert oracle.integrate(cube, imgs, square=False).flat[index] == pytest.approx(linear, rel=1e-6)
    if squared is not None:
        assert oracle.integrate(cube, imgs, square=True).flat[index] == pytest.approx(squared, rel=1e-6)


def test_image_set_integration_kats(oracle):
    c = lambda *a, **k: _cube(oracle, *a, **k)
    _check(oracle, c(1, 1), [5.0], 1, 5.0, 5.0)                                  # xxNormalization
    _check(oracle, c(1, 1), [6.0], 2, 6.0, 6.0)                                  # iNormalization
    _check(oracle, c(2, 1, [1, 1]), [12.0, 13.0], 0, 12.5, 12.5)                 # i_2channel
    _check(oracle, c(2, 1, [0, 1]), [np.nan, 42.0], 0, 42.0, 42.0)               # i_2channel_NaNs
    _check(oracle, c(1, 2, pol_norm=0.5), [7.0, 8.0], 3, 7.5, None)              # xxyy
    _check(oracle, c(1, 2, pol_norm=0.5), [-7.0, 8.0], 3, None, math.sqrt((49 + 64) * 0.5))
    _check(oracle, c(1, 2), [6.0, -1.0], 0, 5.0, math.sqrt(37.0))                # iq
    _check(oracle, c(1, 2, linked=[1, 0]), [3.0, -1.0], 0, 3.0, 3.0)             # linkedI
    _check(oracle, c(1, 4), [9.0, 0.2, 0.2, 0.2], 0, 9.6, math.sqrt(81 + 3 * 0.04))  # iquv
    _check(oracle, c(1, 4, pol_norm=0.5), [10.0, 0.25, 0.25, 10.0], 1, 10.25,
           math.sqrt((200 + 2 * 0.0625) * 0.5))                                  # xx_xy_yx_yy
    v = [5.0, 0.1, 0.2, 6.0, 7.0, 0.3, 0.4, 8.0]
    s1, s2 = sum(np.float32(x) ** 2 for x in v[:4]), sum(np.float32(x) ** 2 for x in v[4:])
    _check(oracle, c(2, 4, [1, 1], pol_norm=0.5), v, 2, 27.0 * 0.25,
           (math.sqrt(s1 * 0.5) + math.sqrt(s2 * 0.5)) * 0.5)                    # ..._2channel
    _check(oracle, c(2, 2, [1, 1], pol_norm=0.5, squared=True), [5.0, 6.0, 7.0, 8.0], 2, None,
           math.sqrt((25 + 36 + 49 + 64) / 4.0))                                 # qu_squared
    v = [7.5, 0.1, -0.2, 6.5, 8.5, 0.3, -0.4, 9.5]
    _check(oracle, c(2, 4, [1, 1], linked=[1, 0, 0, 1], pol_norm=0.5), v, 2, 32.0 * 0.25,
           (math.sqrt((7.5 ** 2 + 6.5 ** 2) * 0.5) + math.sqrt((8.5 ** 2 + 9.5 ** 2) * 0.5)) * 0.5)
    _check(oracle, c(2, 4, [1, 1], linked=[1, 0, 0, 0]), v, 2, 32.0 * 0.25, (7.5 + 8.5) * 0.5)
    _check(oracle, c(2, 1, [2, 4]), [10.0, 13.0], 0, 12.0, 12.0)                 # deconvchannels
    _check(oracle, c(2, 1, [1, 4]), [10.0, 5.0], 0, 6.0, 6.0)                    # zeroweight


# ---- spectral fitter pins: test_image_set.cc:622-675, python/test/test_radler.py:474-596 --
def test_spectral_fit_pins(oracle):
    freqs = np.array([100e6, 200e6])
    # 1 term == weighted mean
    out = oracle.spectral_fit([3.0, 7.0], freqs, [1.0, 1.0], 1)
    assert np.allclose(out, 5.0, rtol=1e-6)
    out = oracle.spectral_fit([2.0, 6.0], freqs, [0.5, 4.2], 1)
    assert np.allclose(out, (2.0 * 0.5 + 6.0 * 4.2) / 4.7, rtol=1e-6)
    # as many terms as points reproduces the data
    out = oracle.spectral_fit([2.0, 6.0], freqs, [0.5, 4.2], 2)
    assert np.allclose(out, [2.0, 6.0], rtol=1e-5)
    # no fitting leaves the values untouched
    out = oracle.spectral_fit([2.0, 6.0], freqs, [1, 1], 0, mode=0)
    assert np.array_equal(out, np.float32([2.0, 6.0]))


# ---- cpp/test/test_radler.cc:98-172 point-source fixture, python/test/test_radler.py:240-282
@pytest.mark.parametrize("shift", [(0, 0), (7, -11), (-9, 15), (-9, 23)])
@pytest.mark.parametrize("algo", ["generic_a", "generic_b", "multiscale"])
def test_point_source_fixture(oracle, algo, shift):
    psf, res, src = plus_psf_fixture(*shift)
    model = np.zeros_like(res)
    if algo == "multiscale":
        a = oracle.MultiScale(beam_size=0.0, pixel_scale=1.0 / 60.0 * math.pi / 180.0)
    else:
        a = oracle.GenericClean(use_sub_minor_optimization=(algo == "generic_a"))
    s = oracle.Settings(threshold=1e-8, max_iterations=1000, clean_border_ratio=0.0,
                        minor_loop_gain=0.1, major_loop_gain=1.0)
    a.execute(oracle.Cube(64, 64), s, res, model, psf)
    assert np.abs(res).max() < 2e-6
    assert model.flat[src] == pytest.approx(2.5, rel=1e-6)
    m = model.copy()
    m.flat[src] = 0
    assert np.abs(m).max() < 2e-6


def test_fft_convolve_against_numpy(oracle):
    rng = np.random.default_rng(11)
    for (h, w) in [(16, 16), (30, 42), (94, 70), (128, 18)]:  # 47 | 94: Bluestein path
        img = rng.standard_normal((h, w)).astype(np.float32)
        ker = rng.standard_normal((h, w)).astype(np.float32)
        want = np.fft.irfft2(np.fft.rfft2(img.astype(np.float64)) * np.fft.rfft2(ker.astype(np.float64)), s=(h, w))
        got = oracle.convolve(img, ker)
        assert np.abs(got - want).max() < 2e-5 * np.abs(want).max()
