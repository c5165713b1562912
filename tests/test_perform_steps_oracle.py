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
