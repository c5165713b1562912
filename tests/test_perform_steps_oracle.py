"""The checkers of csrc/image_statistics.cu: ImageSet::LoadAndAverage of the reference's own
image_set.cc (oracle/_ref) against a numpy statement of its grouping, and the aocommon stand-in's
MedianAndStdDevFromMAD against sorted-array medians. CPU only."""
import numpy as np
import pytest

import oracle
from perform_steps_cases import (average_case, average_numpy, mad_numpy, noise_images,
                                 output_index_of)

needs_ref = pytest.mark.skipif(oracle.ref_lib() is None, reason="oracle/_ref not built")


@needs_ref
@pytest.mark.parametrize("n_original,n_pol,n_deconvolution", [(1, 1, 1), (8, 1, 2), (8, 1, 3),
                                                              (6, 2, 3), (5, 4, 2), (4, 1, 4),
                                                              (7, 1, 0), (3, 2, 9)])
def test_load_and_average_grouping(n_original, n_pol, n_deconvolution):
    planes, weights = average_case(n_original, n_pol, 12, 20, seed=n_original * 10 + n_pol)
    index, n_dec = output_index_of(n_original, n_pol, n_deconvolution)
    got = oracle.ref_load_and_average(planes, weights, n_pol, n_deconvolution or n_original)
    assert got.shape[0] == n_dec * n_pol
    want = average_numpy(planes, weights, index, n_dec * n_pol)
    if n_dec == n_original and n_original > 2:  # the flagged channel alone: 0 * (1 / 0)
        assert np.isnan(got[n_pol:2 * n_pol]).all()
        got, want = np.delete(got, np.s_[n_pol:2 * n_pol], 0), np.delete(want, np.s_[n_pol:2 * n_pol], 0)
    assert np.isfinite(got).all()
    assert np.allclose(got, want, rtol=2e-6, atol=1e-6)


@needs_ref
@pytest.mark.parametrize("name", sorted(noise_images()))
def test_median_and_stddev_from_mad(name):
    a = noise_images()[name]
    m, s = oracle.ref_median_and_stddev_from_mad(a)
    wm, ws = mad_numpy(a)
    assert m == wm and s == ws


@needs_ref
@pytest.mark.parametrize("n_dec,n_orig,n_terms", [(4, 16, 2), (3, 7, 3), (2, 8, 4), (8, 24, 1)])
def test_interpolate_model(n_dec, n_orig, n_terms):
    from perform_steps_cases import interpolate_case, interpolate_numpy
    models, ff, fw, orig = interpolate_case(n_dec, n_orig, 24, 30, seed=n_dec + n_orig)
    got = oracle.ref_interpolate_model(models, 1, ff, fw, n_terms, orig)
    want = interpolate_numpy(models, ff, fw, n_terms, orig)
    assert got.shape == want.shape
    assert np.abs(got - want).max() <= 2e-5 * np.abs(want).max()
    empty = ~models.any(axis=0)
    assert np.all(got[:, empty] == 0.0)


@needs_ref
def test_load_and_average_psfs():
    rng = np.random.default_rng(0)
    psfs = rng.standard_normal((8, 32, 40)).astype(np.float32)
    w = rng.uniform(0.5, 2, 8)
    w[3] = 0.0
    got = oracle.ref_load_and_average_psfs(psfs, w, 3)
    index, n_dec = output_index_of(8, 1, 3)
    want, sums = np.zeros((n_dec, 32, 40), np.float32), np.zeros(n_dec)
    for i, d in enumerate(index):  # image_set.cc:186-190: the product and the sum in double
        want[d] = (want[d].astype(np.float64) + psfs[i].astype(np.float64) * w[i]).astype(np.float32)
        sums[d] += w[i]
    for d in range(n_dec):
        want[d] = (want[d].astype(np.float64) * (1.0 / sums[d])).astype(np.float32)
    assert np.mean(got != want) < 1e-4  # numpy cannot fuse the multiply-add: rare last-bit cases
    assert np.allclose(got, want, rtol=2e-7, atol=0)


@needs_ref
def test_radler_perform_is_the_sum_of_its_steps():
    """radler::Radler::Perform (the reference's own radler.cc) against the same major iteration
    assembled from the pieces the device tests use as checkers: LoadAndAverage(+Psfs),
    MedianAndStdDevFromMAD, the restated GenericClean, InterpolateAndStoreModel."""
    from helpers import hogbom_case
    from perform_steps_cases import deconvolution_channels
    n, n_orig, n_dec = 128, 8, 2
    res, psfs, w_orig, f_orig = hogbom_case(n, n_orig, seed=12)
    w_orig = w_orig.astype(np.float64)
    r_res, r_model = res.copy(), np.zeros_like(res)
    another, iterations = oracle.ref_radler_perform(
        r_res, r_model, psfs, f_orig, w_orig, n_dec, auto_threshold_sigma=5.0,
        minor_loop_gain=0.1, major_loop_gain=0.9, minor_iteration_count=300, fit_terms=2)
    index, _ = output_index_of(n_orig, 1, n_dec)
    f_dec, w_dec = deconvolution_channels(f_orig, w_orig, n_dec)
    p_res = oracle.ref_load_and_average(res, w_orig, 1, n_dec)
    p_psf = oracle.ref_load_and_average_psfs(psfs, w_orig, n_dec)
    cube = oracle.Cube(n, n, n_dec, 1, w_dec)
    _, std = oracle.ref_median_and_stddev_from_mad(oracle.integrate(cube, p_res))
    s = oracle.Settings(threshold=float(np.float32(float(std) * 5.0)), max_iterations=300,
                        minor_loop_gain=0.1, major_loop_gain=0.9, clean_border_ratio=0.0)
    s.spectral_mode, s.n_terms, s.frequencies, s.fit_weights = 1, 2, f_dec, w_dec
    p_model = np.zeros_like(p_res)
    o = oracle.GenericClean(use_sub_minor_optimization=False).execute(cube, s, p_res, p_model, p_psf)
    assert (o.another_iteration_required, o.iteration_number) == (another, iterations)
    assert iterations > 10
    assert np.array_equal(oracle.ref_interpolate_model(p_model, 1, f_dec, w_dec, 2, f_orig), r_model)
    for ch, d in enumerate(index):
        assert np.array_equal(p_res[d], r_res[ch])
