"""GPU parity of csrc/image_statistics.cu (Radler::Perform's steps either side
of ExecuteMajorIteration) through the C ABI: LoadAndAverage BIT-EXACT against the reference's own
image_set.cc (oracle/_ref); MedianAndStdDevFromMAD equal to the aocommon stand-in the reference
sources are compiled against (restated: unpinned) and to sorted-array medians; the auto-mask from
the model against radler.cc:254-266 stated in numpy."""
import numpy as np
import pytest

from perform_steps_cases import average_case, mad_numpy, noise_images, output_index_of

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("n_original,n_pol,n_deconvolution,h,w",
                         [(1, 1, 1, 16, 16), (8, 1, 2, 33, 47), (8, 1, 3, 64, 64),
                          (6, 2, 3, 40, 52), (5, 4, 2, 31, 29), (4, 1, 4, 128, 96),
                          (40, 1, 2, 64, 80), (3, 2, 9, 20, 24)])
def test_load_and_average_bit_exact(rdl, oracle, torch_cuda, n_original, n_pol, n_deconvolution,
                                    h, w):
    torch = torch_cuda
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built")
    planes, weights = average_case(n_original, n_pol, h, w, seed=n_original + 7 * n_pol)
    index, n_dec = output_index_of(n_original, n_pol, n_deconvolution)
    want = oracle.ref_load_and_average(planes, weights, n_pol, n_deconvolution)
    alg = rdl.GenericClean()
    d_in = [torch.from_numpy(p).cuda() for p in planes]
    d_out = [torch.full((h, w), 7.0, device="cuda") for _ in range(n_dec * n_pol)]
    alg.load_and_average(d_in, weights, index, d_out)
    torch.cuda.synchronize()
    got = np.stack([t.cpu().numpy() for t in d_out])
    assert np.array_equal(np.isnan(got), np.isnan(want))
    ok = ~np.isnan(want)
    assert np.array_equal(got[ok].view(np.uint32), want[ok].view(np.uint32))


def test_load_and_average_unaligned_planes_and_errors(rdl, oracle, torch_cuda):
    torch = torch_cuda
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built")
    planes, weights = average_case(6, 1, 17, 19, seed=3, zero_weight_nan=False)
    index, n_dec = output_index_of(6, 1, 2)
    want = oracle.ref_load_and_average(planes, weights, 1, 2)
    alg = rdl.GenericClean()
    # planes at odd float offsets of one buffer: the scalar instance
    flat = torch.zeros(6 * 17 * 19 + 8, device="cuda")
    d_in = []
    for i, p in enumerate(planes):
        view = flat[1 + i * 17 * 19:1 + (i + 1) * 17 * 19].view(17, 19)
        view.copy_(torch.from_numpy(p))
        d_in.append(view)
    d_out = [torch.zeros((17, 19), device="cuda") for _ in range(2)]
    alg.load_and_average(d_in, weights, index, d_out)
    got = np.stack([t.cpu().numpy() for t in d_out])
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    with pytest.raises(rdl.RadlerError):
        alg.load_and_average(d_in, weights, [0, 0, 0, 1, 1, 2], d_out)
    with pytest.raises(rdl.RadlerError):
        alg.load_and_average(d_in, weights, index, [d_in[0], d_out[1]])


@pytest.mark.parametrize("name", sorted(noise_images()))
def test_median_and_stddev_from_mad(rdl, oracle, torch_cuda, name):
    torch = torch_cuda
    a = noise_images()[name]
    alg = rdl.GenericClean()
    m, s, n = alg.median_and_stddev_from_mad(torch.from_numpy(a).cuda())
    wm, ws = mad_numpy(a)
    assert n == int(np.isfinite(a).sum())
    assert np.float32(m) == wm and np.float32(s) == ws
    if oracle.ref_lib() is not None:
        rm, rs = oracle.ref_median_and_stddev_from_mad(a)
        assert np.float32(m) == rm and np.float32(s) == rs


def test_median_and_stddev_from_mad_of_the_integrated_residual(rdl, oracle, torch_cuda):
    """The call Radler::Perform makes (radler.cc:162-166): GetLinearIntegrated of the residual
    cube, then MedianAndStdDevFromMAD -- at 2048^2 x 4, unaligned view included."""
    torch = torch_cuda
    n, c = 2048, 4
    rng = np.random.default_rng(5)
    res = (rng.standard_normal((c, n, n)) * 2e-3).astype(np.float32)
    res[:, 1000:1010, 500:520] += 1.0
    weights = np.array([1.0, 2.0, 0.5, 1.5])
    alg = rdl.GenericClean()
    cube = rdl.Cube(n, n, c, 1, weights)
    d = torch.from_numpy(res).cuda()
    integrated = torch.zeros((n, n), device="cuda")
    alg.integrate(cube, d, integrated)
    torch.cuda.synchronize()
    host = integrated.cpu().numpy()
    m, s, count = alg.median_and_stddev_from_mad(integrated)
    wm, ws = mad_numpy(host)
    assert count == n * n and np.float32(m) == wm and np.float32(s) == ws
    assert abs(s - host[:900].std()) < 0.02 * s  # and it is the noise level
    odd = integrated.view(-1)[1:n * n - 2]
    m, s, count = alg.median_and_stddev_from_mad(odd)
    wm, ws = mad_numpy(host.ravel()[1:n * n - 2])
    assert count == n * n - 3 and np.float32(m) == wm and np.float32(s) == ws


@pytest.mark.parametrize("n_models,h,w", [(1, 33, 35), (3, 64, 64), (40, 16, 24), (2, 17, 19)])
def test_auto_mask_from_model(rdl, torch_cuda, n_models, h, w):
    torch = torch_cuda
    rng = np.random.default_rng(n_models)
    models = np.zeros((n_models, h, w), np.float32)
    for m in models:
        ys, xs = rng.integers(0, h, 12), rng.integers(0, w, 12)
        m[ys, xs] = rng.standard_normal(12).astype(np.float32)
    models[0, 1, 2] = np.nan
    models[-1, 2, 3] = np.inf
    models[0, 3, 4] = -0.0
    want = np.zeros((h, w), bool)
    for m in models:  # radler.cc:258-263
        want |= np.isfinite(m) & (m != 0.0)
    alg = rdl.GenericClean()
    d = [torch.from_numpy(m).cuda() for m in models]
    mask = torch.full((h, w), 9, dtype=torch.uint8, device="cuda")
    alg.auto_mask_from_model(d, mask)
    assert np.array_equal(mask.cpu().numpy(), want.astype(np.uint8))
    alg.auto_mask_from_model([], mask)
    assert int(mask.sum()) == 0
