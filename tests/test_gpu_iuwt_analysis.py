"""GPU parity of the IUWT structure selection kernels (csrc/iuwt_analysis.cu) through the C ABI
against the CPU oracle restatement (oracle/iuwt_analysis.cc, itself pinned to the reference's own
image_analysis.cc): masks, areas, positions and values must be IDENTICAL. Covers the cases of the
oracle test (positive / sign-aware thresholds, prior mask, pre-marked voxels, clean border, seeds
outside the border), a structure that winds across many tiles, and size-independent properties
at the K3 size (2048^2, 6 scales)."""
import numpy as np
import pytest

from test_iuwt_analysis_oracle import CASES, coefficient_planes, setup_case

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    import torch
    import oracle
    import ska_sdp_func_radler_b200 as rdl
    return torch, oracle, rdl, rdl.DeconvolutionAlgorithm(rdl.IUWT)


def dev(torch, a):
    return None if a is None else torch.from_numpy(np.ascontiguousarray(a)).cuda()


@pytest.mark.parametrize("case", CASES)
def test_select_structures_identical(env, case):
    torch, oracle, _, alg = env
    sel = oracle.iuwt_analysis()[0]
    for seed in range(3):
        c, masks, thr, lo, hi, border, prior = setup_case(case, seed)
        want, warea = sel(c, masks, thr, lo, hi, border, prior)
        m = dev(torch, masks)
        area = alg.iuwt_select_structures(dev(torch, c), m, thr, lo, hi, border, dev(torch, prior))
        assert area == warea
        assert np.array_equal(m.cpu().numpy(), want)


@pytest.mark.parametrize("case", CASES)
def test_seeded_floodfill_identical(env, case):
    torch, oracle, _, alg = env
    fill = oracle.iuwt_analysis()[1]
    n_scales, h, w = case[:3]
    for seed in range(3):
        c, masks, thr, lo, hi, border, prior = setup_case(case, seed)
        rng = np.random.default_rng(seed)
        s0, y0, x0 = np.unravel_index(np.argmax(np.abs(c[lo:hi])), c[lo:hi].shape)
        seeds = [(int(x0), int(y0), int(s0) + lo),
                 (int(rng.integers(0, w)), int(rng.integers(0, h)), int(rng.integers(lo, hi))),
                 (0, 0, lo), (w - 1, h - 1, hi - 1)]
        dc, dp = dev(torch, c), dev(torch, prior)
        for sd in seeds:
            want, warea = fill(c, masks, thr, lo, hi, sd, border, prior)
            m = dev(torch, masks)
            area = alg.iuwt_floodfill(dc, m, thr, lo, hi, sd, border, dp)
            assert area == warea, sd
            assert np.array_equal(m.cpu().numpy(), want), sd


def test_floodfill_2d_identical(env):
    torch, oracle, _, alg = env
    fill2 = oracle.iuwt_analysis()[2]
    for seed in range(3):
        img = coefficient_planes(seed, 1, 48, 64)[0]
        rng = np.random.default_rng(seed)
        pre = (rng.random(img.shape) < 0.03).astype(np.uint8)
        y0, x0 = np.unravel_index(np.argmax(np.abs(img)), img.shape)
        d = dev(torch, img)
        for thr in (1.0, -1.0, 0.3):
            for use_abs in (False, True):
                for sd in [(int(x0), int(y0)), (0, 0), (63, 47)]:
                    want, warea = fill2(img, pre, thr, sd, use_abs)
                    m = dev(torch, pre)
                    assert alg.iuwt_floodfill_2d(d, m, thr, sd, use_abs) == warea
                    assert np.array_equal(m.cpu().numpy(), want)


def spiral(n, gap=3):
    """One-pixel-wide square spiral: a single structure that crosses every tile many times."""
    img = np.zeros((n, n), np.float32)
    x0, y0, x1, y1 = 1, 1, n - 2, n - 2
    while x1 - x0 > 2 * gap and y1 - y0 > 2 * gap:
        img[y0, x0:x1 + 1] = 1
        img[y0:y1 + 1, x1] = 1
        img[y1, x0 + gap:x1 + 1] = 1
        img[y0 + gap:y1 + 1, x0 + gap] = 1
        img[y0 + gap, x0 + gap:x1 - gap + 1] = 1  # connects to the next turn
        x0 += gap
        y0 += gap
        x1 -= gap
        y1 -= gap
        # the connecting stroke above is the top edge of the next turn
    return img


def test_winding_structure_across_tiles(env):
    torch, oracle, _, alg = env
    fill, fill2 = oracle.iuwt_analysis()[1:3]
    n = 333  # odd width: the scalar code path
    img = spiral(n)
    want, warea = fill2(img, np.zeros((n, n), np.uint8), 0.5, (1, 1))
    assert warea == int((img > 0).sum())  # the spiral is one connected piece
    m = torch.zeros((n, n), dtype=torch.uint8, device="cuda")
    assert alg.iuwt_floodfill_2d(dev(torch, img), m, 0.5, (1, 1)) == warea
    assert np.array_equal(m.cpu().numpy(), want)
    # the same spiral, alternating between two scales every few pixels
    c = np.zeros((2, n, n), np.float32)
    yy, xx = np.mgrid[0:n, 0:n]
    upper = ((xx // 5 + yy // 5) % 2 == 0)
    c[0] = img * upper
    c[1] = img * ~upper
    # overlap the two planes by one pixel so the structure stays connected through scale moves
    both = img > 0
    c[0][both & (xx % 5 == 0)] = 1
    c[1][both & (xx % 5 == 0)] = 1
    c[0][both & (yy % 5 == 0)] = 1
    c[1][both & (yy % 5 == 0)] = 1
    thr = np.array([0.5, 0.5], np.float32)
    zeros = np.zeros((2, n, n), np.uint8)
    want3, warea3 = fill(c, zeros, thr, 0, 2, (1, 1, 0))
    m3 = torch.zeros((2, n, n), dtype=torch.uint8, device="cuda")
    assert alg.iuwt_floodfill(dev(torch, c), m3, thr, 0, 2, (1, 1, 0)) == warea3
    assert warea3 > warea
    assert np.array_equal(m3.cpu().numpy(), want3)


def test_serpentine_needs_many_rounds(env):
    """Rows two pixels apart joined at alternating ends: one structure whose length is ~n^2/2, so
    the tile rounds must keep going until nothing changes (hundreds of rounds)."""
    torch, oracle, _, alg = env
    fill2 = oracle.iuwt_analysis()[2]
    n = 200
    img = np.zeros((n, n), np.float32)
    img[0::2, :] = -2.0  # found through the sign-aware test with a negative threshold
    for k, y in enumerate(range(1, n, 2)):
        img[y, n - 1 if k % 2 == 0 else 0] = -2.0
    want, warea = fill2(img, np.zeros((n, n), np.uint8), -1.0, (0, 0))
    assert warea == int((img != 0).sum())
    m = torch.zeros((n, n), dtype=torch.uint8, device="cuda")
    assert alg.iuwt_floodfill_2d(dev(torch, img), m, -1.0, (0, 0)) == warea
    assert np.array_equal(m.cpu().numpy(), want)


def test_mad_and_max_abs_identical(env):
    torch, oracle, _, alg = env
    mad, max_abs = oracle.iuwt_analysis()[3:5]
    rng = np.random.default_rng(3)
    for (h, w) in [(31, 45), (64, 64), (50, 51), (512, 768)]:
        img = rng.standard_normal((h, w)).astype(np.float32)
        img[rng.integers(0, h), rng.integers(0, w)] = -7.5
        d = dev(torch, img)
        assert alg.iuwt_mad(d) == mad(img)
        mask = (rng.random((h, w)) < 0.3).astype(np.uint8)
        for border in (0.0, 0.1):
            for neg in (True, False):
                for m in (None, mask, np.zeros_like(mask)):
                    assert alg.iuwt_max_abs(d, border, neg, dev(torch, m)) == max_abs(img, border, neg, m)
    # many equal values: the median falls inside a run of duplicates
    q = np.round(rng.standard_normal((40, 40)) * 2).astype(np.float32)
    assert alg.iuwt_mad(dev(torch, q)) == mad(q)
    assert alg.iuwt_mad(dev(torch, np.zeros((8, 8), np.float32))) == 0.0
    tie = np.zeros((8, 8), np.float32)
    tie[2, 5] = tie[2, 3] = tie[6, 1] = 4.0
    assert alg.iuwt_max_abs(dev(torch, tie)) == (4.0, 3, 2)
    neg = -np.abs(rng.standard_normal((16, 16)).astype(np.float32)) - 1.0
    assert alg.iuwt_max_abs(dev(torch, neg), 0.0, False) == max_abs(neg, 0.0, False)
    zero = np.zeros((6, 6), np.float32)
    zero[0, 0] = -0.0
    assert alg.iuwt_max_abs(dev(torch, zero)) == max_abs(zero)


def test_full_size_properties(env):
    """K3 size (2048^2, 6 scales), no oracle: the selected set equals the torch expression of
    the threshold test; the seeded fill is idempotent, contained in the selected set, and
    filling from every selected voxel's structure cannot exceed it; Mad equals torch's k-th
    value; max_abs equals torch's arg-max."""
    torch, _, _, alg = env
    g = torch.Generator(device="cuda").manual_seed(11)
    s, n = 6, 2048
    c = torch.randn((s, n, n), device="cuda", generator=g)
    c[:, 1000:1040, 900:1100] += 6.0
    thr = [3.0, 3.2, -3.1, 3.0, -2.9, 3.3]
    border = 0.05
    b = int(np.float32(border) * n)
    masks = torch.zeros((s, n, n), dtype=torch.uint8, device="cuda")
    area = alg.iuwt_select_structures(c, masks, thr, 1, 6, border)
    want = torch.zeros_like(masks)
    for k in range(1, 6):
        t = thr[k]
        over = (c[k] > t) if t >= 0 else ((c[k] < t) | (c[k] > -t))
        want[k, b:n - b, b:n - b] = over[b:n - b, b:n - b].to(torch.uint8)
    assert torch.equal(masks, want) and area == int(want.sum())
    # selecting again adds nothing
    assert alg.iuwt_select_structures(c, masks, thr, 1, 6, border) == 0
    # the seeded fill from the bright box stays inside the selected set and is idempotent
    fill = torch.zeros_like(masks)
    a1 = alg.iuwt_floodfill(c, fill, thr, 1, 6, (1000, 1020, 3), border)
    assert a1 >= 40 * 200 and a1 == int(fill.sum())
    assert bool(((fill == 1) & (want == 0)).sum() == 0)
    again = fill.clone()
    assert alg.iuwt_floodfill(c, again, thr, 1, 6, (1000, 1020, 3), border) == 1  # only the seed
    assert torch.equal(again, fill)
    # measurements
    plane = c[2]
    kth = torch.kthvalue(plane.abs().flatten(), n * n // 2 + 1).values.item()
    assert alg.iuwt_mad(plane) == float(np.float32(np.float64(np.float32(kth)) / 0.674559))
    v, x, y = alg.iuwt_max_abs(plane, 0.0, True)
    flat = plane.abs().flatten()
    assert v == flat.max().item()
    first = int(torch.nonzero(flat == flat.max())[0])
    assert (x, y) == (first % n, first // n)
