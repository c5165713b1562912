"""Pins the IUWT restatement (oracle/iuwt.cc) bit-exactly against the reference's own
iuwt_decomposition.{h,cc} compiled in oracle/_ref, and writes/reads the golden fixture the GPU
tests fall back on when oracle/_ref is absent."""
import os

import numpy as np
import pytest

import oracle

GOLDEN = os.path.join(os.path.dirname(__file__), "golden", "iuwt_decomposition.npz")


def image(h, w, seed):
    rng = np.random.default_rng(seed)
    img = rng.normal(0, 1e-2, (h, w)).astype(np.float32)
    yy, xx = np.mgrid[0:h, 0:w]
    for _ in range(6):
        x0, y0, s = rng.uniform(0, w), rng.uniform(0, h), rng.uniform(1.5, 12)
        img += (rng.uniform(0.1, 3) * np.exp(-0.5 * ((xx - x0) ** 2 + (yy - y0) ** 2) / s ** 2)).astype(np.float32)
    return img


needs_ref = pytest.mark.skipif(oracle.ref_lib() is None, reason="oracle/_ref not built")


@needs_ref
@pytest.mark.parametrize("h,w,n_scales", [(64, 64, 2), (96, 80, 3), (70, 131, 4), (256, 256, 5),
                                          (30, 29, 3)])
@pytest.mark.parametrize("include_largest", [True, False])
def test_decompose_matches_reference_sources(h, w, n_scales, include_largest):
    img = image(h, w, h * 1000 + w)
    ours = oracle.iuwt_decompose(img, n_scales, include_largest)
    ref = oracle.ref_iuwt_decompose(img, n_scales, include_largest)
    assert ours.shape == ref.shape
    assert np.array_equal(ours, ref)


@needs_ref
@pytest.mark.parametrize("h,w,n_scales", [(64, 64, 2), (96, 80, 3), (70, 131, 4), (256, 256, 5)])
@pytest.mark.parametrize("masked", [False, True])
@pytest.mark.parametrize("include_largest", [True, False])
def test_recompose_matches_reference_sources(h, w, n_scales, masked, include_largest):
    img = image(h, w, 7 + h)
    coef = oracle.iuwt_decompose(img, n_scales, True)
    masks = None
    if masked:
        masks = (np.random.default_rng(3).random((n_scales, h, w)) > 0.4).astype(np.uint8)
    ours = oracle.iuwt_recompose(coef, n_scales, include_largest, masks)
    ref = oracle.ref_iuwt_recompose(coef, n_scales, include_largest, masks)
    assert np.array_equal(ours, ref)


@needs_ref
def test_simple_smoothing_matches_reference_sources():
    img = image(90, 75, 11)
    ref = oracle.ref_iuwt_smooth(img, 3)  # IuwtDecomposition::Convolve: scales 1, 2, 3
    ours = img
    for s in (1, 2, 3):
        ours = oracle.iuwt_smooth(ours, s, mt=False)
    assert np.array_equal(ours, ref)


def test_recompose_inverts_decompose():
    # size-independent property: i0 = B(i1) + w at every scale, so Recompose(include_largest)
    # returns the input to round-off
    img = image(128, 128, 5)
    coef = oracle.iuwt_decompose(img, 4, True)
    rec = oracle.iuwt_recompose(coef, 4, True)
    assert np.abs(rec - img).max() <= 1e-5 * np.abs(img).max()


def test_golden_fixture():
    """tests/golden/iuwt_decomposition.npz was generated from the reference's own sources by
    tests/golden/make_iuwt_golden.py; the restatement must reproduce it bit for bit."""
    g = np.load(GOLDEN)
    img = g["image"]
    n_scales = int(g["n_scales"])
    assert np.array_equal(oracle.iuwt_decompose(img, n_scales, True), g["coefficients"])
    assert np.array_equal(oracle.iuwt_recompose(g["coefficients"], n_scales, False, g["masks"]),
                          g["recomposed_masked"])
