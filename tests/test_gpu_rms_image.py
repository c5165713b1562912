"""GPU parity of the local-RMS map kernels (csrc/rms_image.cu) through the C ABI against the
reference's OWN cpp/math/rms_image.cc (oracle/_ref): SlidingMinimum / SlidingMaximum and
MakeRmsFactorImage BIT-EXACT; Make / MakeWithNegativityLimit against a float64 restatement of
the reference's lines (the Gaussian window, schaapcommon's RestoreImage, is not in the reference
tree: restated, unpinned -- DESIGN.md)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _image(h, w, seed):
    rng = np.random.default_rng(seed)
    a = rng.standard_normal((h, w)).astype(np.float32)
    a[h // 3, w // 2] = -25.0
    a[h // 2, w // 4] = 40.0
    return a


@pytest.mark.parametrize("h,w,window", [(37, 53, 7), (64, 64, 2), (100, 300, 25), (257, 129, 64),
                                        (512, 512, 201), (90, 70, 140)])
@pytest.mark.parametrize("maximum", [False, True])
def test_sliding_minimum_bit_exact(rdl, oracle, torch_cuda, h, w, window, maximum):
    torch = torch_cuda
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built")
    a = _image(h, w, h + w + window)
    want = oracle.ref_rms_sliding_minimum(a, window, maximum)
    alg = rdl.GenericClean()
    d, out = torch.from_numpy(a).cuda(), torch.full((h, w), 9.0, device="cuda")
    alg.rms_sliding_minimum(d, out, window, maximum)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy().view(np.uint32), want.view(np.uint32))


def test_sliding_minimum_rejects_bad_windows(rdl, torch_cuda):
    torch = torch_cuda
    alg = rdl.GenericClean()
    d, out = torch.zeros((32, 32), device="cuda"), torch.zeros((32, 32), device="cuda")
    for window in (0, 1, 80):
        with pytest.raises(rdl.RadlerError):
            alg.rms_sliding_minimum(d, out, window)
    with pytest.raises(rdl.RadlerError):
        alg.rms_sliding_minimum(d, d, 4)


@pytest.mark.parametrize("zeros", [False, True])
@pytest.mark.parametrize("strength", [1.0, 0.0, 0.5, 2.0])
def test_make_rms_factor_image(rdl, oracle, torch_cuda, strength, zeros):
    torch = torch_cuda
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(4)
    rms = rng.uniform(0.2, 3.0, (96, 130)).astype(np.float32)
    if zeros:  # a zero pixel is the minimum: stddev 0, every factor 0 (strength 0: all ones)
        rms[5, 7] = 0.0
        rms[50, 60] = 0.0
    want, want_std = oracle.ref_rms_make_factor_image(rms, strength)
    alg = rdl.GenericClean()
    d = torch.from_numpy(rms.copy()).cuda()
    got_std = alg.rms_make_factor_image(d, strength)
    got = d.cpu().numpy()
    assert got_std == want_std
    if strength in (0.0, 1.0):
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    else:  # std::pow on the host vs pow on the device: an ulp
        assert np.allclose(got, want, rtol=3e-7, atol=0.0)
    neg = torch.from_numpy(-np.abs(rms) - 1.0).cuda()
    with pytest.raises(rdl.RadlerError):
        alg.rms_make_factor_image(neg, 1.0)


def _gaussian_window(h, w, s_maj, s_min, pa):
    y, x = np.mgrid[0:h, 0:w].astype(np.float64)
    dx, dy = x - w // 2, y - h // 2
    u = dx * np.sin(pa) + dy * np.cos(pa)
    v = dx * np.cos(pa) - dy * np.sin(pa)
    return np.exp(-0.5 * (u * u / s_maj ** 2 + v * v / s_min ** 2))


@pytest.mark.parametrize("h,w,negativity", [(128, 128, False), (200, 160, True), (131, 97, True)])
def test_make_follows_the_reference_lines(rdl, oracle, torch_cuda, h, w, negativity):
    torch = torch_cuda
    a = _image(h, w, 11)
    window, bmaj, bmin, pa, scale = 5.0, 4.0e-4, 3.0e-4, 0.6, 1.0e-4  # beam of 4 x 3 pixels
    alg = rdl.GenericClean()
    d, out = torch.from_numpy(a).cuda(), torch.zeros((h, w), device="cuda")
    alg.rms_make(d, out, window, bmaj, bmin, pa, scale, scale, negativity)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    # rms_image.cc:16-37 in float64; RestoreImage = circular convolution with a unit-peak Gaussian
    k = 1.0 / (2.0 * np.sqrt(2.0 * np.log(2.0)))
    g = _gaussian_window(h, w, bmaj * window * k / scale, bmin * window * k / scale, pa)
    g = np.roll(g, (-(h // 2), -(w // 2)), axis=(0, 1))
    sq = a.astype(np.float64) ** 2
    conv = np.fft.irfft2(np.fft.rfft2(sq) * np.fft.rfft2(g), s=(h, w))
    s = np.sqrt(2.0 * np.pi)
    norm = 1.0 / (s * (bmaj * k) / scale * window * s * (bmin * k) / scale * window)
    want = np.sqrt(np.maximum(conv, 0.0) * norm)
    if negativity:
        if oracle.ref_lib() is None:
            pytest.skip("oracle/_ref not built")
        beam_px = max(bmaj / scale, 1.0)
        smin = oracle.ref_rms_sliding_minimum(a, int(window * beam_px))
        want = np.maximum(want, np.abs(smin.astype(np.float64)) * (1.5 / 5.0))
    assert np.abs(got - want).max() <= 2e-5 * want.max()
