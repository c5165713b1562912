"""Oracle restatement of the IUWT structure selection (oracle/iuwt_analysis.cc) pinned against
the reference's own image_analysis.cc and the private Mad / GetMaxAbs members of
IuwtDeconvolutionAlgorithm, compiled where they lie into oracle/_ref. Masks, areas, positions and
values must be identical (all of it is compare/select work, no rounding involved except the
final Mad division, which both sides do in double)."""
import numpy as np
import pytest

import oracle

needs_ref = pytest.mark.skipif(oracle.ref_lib() is None, reason="oracle/_ref not built")


def coefficient_planes(seed, n_scales, h, w, blobs=6):
    """Noise plus a few positive and negative blobs whose footprint grows with scale, so that the
    structures connect through the scale axis as well as in the plane."""
    rng = np.random.default_rng(seed)
    c = rng.standard_normal((n_scales, h, w)).astype(np.float32)
    yy, xx = np.mgrid[0:h, 0:w]
    for _ in range(blobs):
        cx, cy = rng.integers(4, w - 4), rng.integers(4, h - 4)
        amp = rng.choice([-1.0, 1.0]) * rng.uniform(3, 9)
        for s in range(n_scales):
            r = 1.5 * (s + 1)
            c[s] += (amp * np.exp(-((xx - cx) ** 2 + (yy - cy) ** 2) / (2 * r * r))).astype(np.float32)
    return c


CASES = [
    # n_scales, h, w, min_scale, end_scale, clean_border, threshold sign, prior mask, pre-mask
    (4, 40, 56, 0, 4, 0.0, +1, False, False),
    (4, 40, 56, 1, 3, 0.1, +1, False, True),
    (5, 64, 48, 0, 5, 0.05, -1, True, False),
    (3, 33, 47, 0, 3, 0.2, -1, True, True),
    (1, 30, 30, 0, 1, 0.0, +1, False, False),
]


def setup_case(case, seed):
    n_scales, h, w, lo, hi, border, sign, with_prior, with_pre = case
    rng = np.random.default_rng(seed + 100)
    c = coefficient_planes(seed, n_scales, h, w)
    thr = (sign * rng.uniform(1.2, 2.2, n_scales)).astype(np.float32)
    prior = (rng.random((h, w)) < 0.8).astype(np.uint8) if with_prior else None
    masks = ((rng.random((n_scales, h, w)) < 0.05).astype(np.uint8) if with_pre
             else np.zeros((n_scales, h, w), np.uint8))
    return c, masks, thr, lo, hi, border, prior


@needs_ref
@pytest.mark.parametrize("case", CASES)
def test_select_structures_matches_reference_source(case):
    sel, _, _, _, _ = oracle.iuwt_analysis()
    rsel, _, _, _, _ = oracle.ref_iuwt_analysis()
    for seed in range(3):
        c, masks, thr, lo, hi, border, prior = setup_case(case, seed)
        got, area = sel(c, masks, thr, lo, hi, border, prior)
        want, rarea = rsel(c, masks, thr, lo, hi, border, prior)
        assert area == rarea and area > 0
        assert np.array_equal(got, want)


@needs_ref
@pytest.mark.parametrize("case", CASES)
def test_seeded_floodfill_matches_reference_source(case):
    _, fill, _, _, _ = oracle.iuwt_analysis()
    _, rfill, _, _, _ = oracle.ref_iuwt_analysis()
    n_scales, h, w = case[:3]
    for seed in range(3):
        c, masks, thr, lo, hi, border, prior = setup_case(case, seed)
        rng = np.random.default_rng(seed)
        # seeds: the strongest voxel, a random one, and one outside the clean border
        s0, y0, x0 = np.unravel_index(np.argmax(np.abs(c[lo:hi])), c[lo:hi].shape)
        seeds = [(int(x0), int(y0), int(s0) + lo),
                 (int(rng.integers(0, w)), int(rng.integers(0, h)), int(rng.integers(lo, hi))),
                 (0, 0, lo), (w - 1, h - 1, hi - 1)]
        for sd in seeds:
            got, area = fill(c, masks, thr, lo, hi, sd, border, prior)
            want, rarea = rfill(c, masks, thr, lo, hi, sd, border, prior)
            assert area == rarea, sd
            assert np.array_equal(got, want), sd


@needs_ref
def test_floodfill_2d_matches_reference_source():
    _, _, fill2, _, _ = oracle.iuwt_analysis()
    _, _, rfill2, _, _ = oracle.ref_iuwt_analysis()
    for seed in range(4):
        img = coefficient_planes(seed, 1, 48, 64)[0]
        rng = np.random.default_rng(seed)
        pre = (rng.random(img.shape) < 0.03).astype(np.uint8)
        y0, x0 = np.unravel_index(np.argmax(np.abs(img)), img.shape)
        for thr in (1.0, -1.0, 0.3):
            for use_abs in (False, True):
                for sd in [(int(x0), int(y0)), (0, 0), (63, 47)]:
                    got, area = fill2(img, pre, thr, sd, use_abs)
                    want, rarea = rfill2(img, pre, thr, sd, use_abs)
                    assert area == rarea and np.array_equal(got, want)


@needs_ref
def test_mad_and_max_abs_match_reference_source():
    _, _, _, mad, max_abs = oracle.iuwt_analysis()
    _, _, _, rmad, rmax_abs = oracle.ref_iuwt_analysis()
    rng = np.random.default_rng(3)
    for (h, w) in [(31, 45), (64, 64), (50, 51)]:
        img = rng.standard_normal((h, w)).astype(np.float32)
        img[rng.integers(0, h), rng.integers(0, w)] = -7.5
        assert mad(img) == rmad(img)
        mask = (rng.random((h, w)) < 0.3).astype(np.uint8)
        for border in (0.0, 0.1):
            for neg in (True, False):
                for m in (None, mask, np.zeros_like(mask)):
                    assert max_abs(img, border, neg, m) == rmax_abs(img, border, neg, m)
    # ties: the first pixel in row-major order wins; an all-negative plane without abs
    tie = np.zeros((8, 8), np.float32)
    tie[2, 5] = tie[2, 3] = tie[6, 1] = 4.0
    assert max_abs(tie) == rmax_abs(tie) == (4.0, 3, 2)
    neg = -np.abs(rng.standard_normal((16, 16)).astype(np.float32)) - 1.0
    assert max_abs(neg, 0.0, False) == rmax_abs(neg, 0.0, False)


def test_select_structures_is_the_thresholded_set():
    """Every voxel that passes the test is itself a seed, so the union of the fills is exactly
    the set {exceeds, not yet masked, inside prior mask, inside the border}: the property the
    device kernel is built on."""
    sel, _, _, _, _ = oracle.iuwt_analysis()
    for case in CASES:
        c, masks, thr, lo, hi, border, prior = setup_case(case, 7)
        n_scales, h, w = c.shape
        got, area = sel(c, masks, thr, lo, hi, border, prior)
        bx, by = int(np.float32(border) * w), int(np.float32(border) * h)
        want = masks.copy()
        for s in range(lo, hi):
            t = thr[s]
            over = (c[s] > t) if t >= 0 else ((c[s] < t) | (c[s] > -t))
            if prior is not None:
                over &= prior != 0
            inside = np.zeros((h, w), bool)
            inside[by:h - by, bx:w - bx] = True
            want[s] |= (over & inside).astype(np.uint8)
        assert np.array_equal(got, want)
        assert area == int(want.sum()) - int(masks.sum())
