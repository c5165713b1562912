"""GPU parity of the hand-written FFT convolution (schaapcommon::math::Convolve restated,
called at multiscale_transforms.cc:21 and subminor_loop.cc:210) and MultiScaleTransforms::
Transform against the CPU oracle and a float64 numpy FFT. FFT results are not bit-comparable
between implementations: tolerance 1e-5 of the result's peak (north_star tolerance)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu
TOL = 1e-5

SIZES = [(64, 64), (30, 42), (18, 128), (250, 256), (512, 384), (1176, 1134),
         (1128, 1128),   # GenericClean padding of 1024: 2^3 * 3 * 47 (generic prime radix)
         (2254, 282),    # 2 * 7^2 * 23
         (4608, 4536)]   # multiscale padded sizes of 4096


def _f64_convolve(img, ker):
    h, w = img.shape
    return np.fft.irfft2(np.fft.rfft2(img.astype(np.float64)) * np.fft.rfft2(ker.astype(np.float64)),
                         s=(h, w))


@pytest.mark.parametrize("h,w", SIZES)
def test_convolve_matches_float64_fft(rdl, torch_cuda, h, w):
    torch = torch_cuda
    rng = np.random.default_rng(h * 100003 + w)
    imgs = rng.standard_normal((2, h, w)).astype(np.float32)
    imgs[1] = 0
    imgs[1, h // 3, w // 5] = 1.0  # impulse: result is the shifted kernel
    ker = rng.standard_normal((h, w)).astype(np.float32)
    alg = rdl.GenericClean()
    dev = torch.from_numpy(imgs).cuda()
    alg.convolve(dev, torch.from_numpy(ker).cuda())
    torch.cuda.synchronize()
    got = dev.cpu().numpy()
    for k in range(2):
        want = _f64_convolve(imgs[k], ker)
        err = np.abs(got[k] - want).max() / np.abs(want).max()
        assert err < TOL, (h, w, k, err)


def test_convolve_matches_oracle(rdl, oracle, torch_cuda):
    torch = torch_cuda
    rng = np.random.default_rng(3)
    for (h, w) in [(64, 64), (282, 94), (512, 640)]:
        img = rng.standard_normal((1, h, w)).astype(np.float32)
        ker = rng.standard_normal((h, w)).astype(np.float32)
        dev = torch.from_numpy(img).cuda()
        rdl.GenericClean().convolve(dev, torch.from_numpy(ker).cuda())
        want = oracle.convolve(img[0], ker)
        got = dev.cpu().numpy()[0]
        assert np.abs(got - want).max() / np.abs(want).max() < TOL


def test_odd_width_is_rejected_loudly(rdl, torch_cuda):
    torch = torch_cuda
    dev = torch.zeros((1, 16, 15), dtype=torch.float32, device="cuda")
    with pytest.raises(rdl.RadlerError) as e:
        rdl.GenericClean().convolve(dev, torch.zeros((16, 15), dtype=torch.float32, device="cuda"))
    assert e.value.status == rdl.RDL_ERR_UNSUPPORTED


@pytest.mark.parametrize("shape", [0, 1])
@pytest.mark.parametrize("n,scale", [(64, 4.0), (256, 12.0), (256, 48.0), (500, 24.0), (1024, 96.0)])
def test_multiscale_transform_matches_oracle(rdl, oracle, torch_cuda, n, scale, shape):
    torch = torch_cuda
    rng = np.random.default_rng(int(n + scale))
    imgs = rng.standard_normal((3, n, n)).astype(np.float32)
    imgs[2] = 0
    imgs[2, n // 2, n // 2] = 1.0
    dev = torch.from_numpy(imgs).cuda()
    rdl.GenericClean().multiscale_transform(dev, scale, shape)
    torch.cuda.synchronize()
    want = oracle.ms_transform(imgs, scale, shape)
    got = dev.cpu().numpy()
    for k in range(3):
        assert np.abs(got[k] - want[k]).max() / np.abs(want[k]).max() < TOL, (n, scale, shape, k)
    # the transformed impulse is the (unit-sum) scale kernel centred on the pixel
    assert got[2].sum() == pytest.approx(1.0, rel=1e-4)
    k = oracle.make_shape_function(scale, n, shape)
    assert got[2, n // 2, n // 2] == pytest.approx(k[k.shape[0] // 2, k.shape[0] // 2], rel=1e-4)
