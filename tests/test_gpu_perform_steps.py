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


@pytest.mark.parametrize("n_dec,n_orig,n_terms,h,w", [(4, 16, 2, 48, 64), (3, 7, 3, 33, 35),
                                                      (2, 8, 4, 20, 24), (8, 40, 1, 16, 16),
                                                      (5, 70, 8, 12, 20)])
def test_interpolate_model_bit_exact(rdl, oracle, torch_cuda, n_dec, n_orig, n_terms, h, w):
    """InterpolateAndStoreModel of the reference's own image_set.cc through the polynomial
    fitter the reference sources are compiled against (schaapcommon's is not in the tree:
    restated, unpinned) -- terms clamped to the channels (n_terms 4 on 2 channels), more outputs
    than one launch takes (40, 70), a zero-weight fit channel."""
    from perform_steps_cases import interpolate_case
    torch = torch_cuda
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built")
    models, ff, fw, orig = interpolate_case(n_dec, n_orig, h, w, seed=n_dec * n_orig)
    if n_dec == 5:
        fw[2] = 0.0
    want = oracle.ref_interpolate_model(models, 1, ff, fw, n_terms, orig)
    alg = rdl.GenericClean()
    alg.set_spectral_fitter(rdl.FIT_POLYNOMIAL, n_terms, ff, fw)
    d_models = [torch.from_numpy(m).cuda() for m in models]
    d_out = [torch.full((h, w), 3.0, device="cuda") for _ in range(n_orig)]
    alg.interpolate_model(d_models, orig, d_out)
    got = np.stack([t.cpu().numpy() for t in d_out])
    assert np.count_nonzero(want) > 30 * n_orig
    assert np.array_equal(got, want)  # zeros of either sign are zeros


def test_interpolate_model_needs_the_polynomial_fitter(rdl, torch_cuda):
    torch = torch_cuda
    alg = rdl.GenericClean()
    m = [torch.zeros((8, 8), device="cuda") for _ in range(2)]
    o = [torch.zeros((8, 8), device="cuda") for _ in range(3)]
    with pytest.raises(rdl.RadlerError):
        alg.interpolate_model(m, [1e8, 1.1e8, 1.2e8], o)
    alg.set_spectral_fitter(rdl.FIT_POLYNOMIAL, 2, [1e8, 1.2e8], None)
    with pytest.raises(rdl.RadlerError):
        alg.interpolate_model(m + m, [1e8, 1.1e8, 1.2e8], o)
    with pytest.raises(rdl.RadlerError):
        alg.interpolate_model(m, [1e8, 1.1e8], [m[0], o[1]])
    alg.interpolate_model(m, [1e8, 1.1e8, 1.2e8], o)


def test_one_major_iteration_of_perform_on_the_device(rdl, oracle, torch_cuda):
    """The sequence Radler::Perform runs (radler.cc:150-273, image_set.cc): LoadAndAverage of 8
    original channels into 2 deconvolution channels, GetLinearIntegrated + MedianAndStdDevFromMAD,
    the threshold from the noise, GenericClean::ExecuteMajorIteration with a 2-term polynomial
    fit, InterpolateAndStoreModel back to the 8 channels -- every step on device planes, one copy
    back at the end, against the same sequence through the reference's own sources / the
    oracle."""
    from helpers import hogbom_case
    from perform_steps_cases import output_index_of
    torch = torch_cuda
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built")
    n, n_orig, n_dec = 256, 8, 2
    res, psfs, w_orig, f_orig = hogbom_case(n, n_orig, seed=12)
    w_orig = w_orig.astype(np.float64)
    index, _ = output_index_of(n_orig, 1, n_dec)
    groups = [[i for i, o in enumerate(index) if o == d] for d in range(n_dec)]
    w_dec = np.array([w_orig[g].sum() for g in groups], np.float32)
    f_dec = np.array([(f_orig[g] * w_orig[g]).sum() / w_orig[g].sum() for g in groups])
    kw = dict(max_iterations=300, minor_loop_gain=0.1, major_loop_gain=0.8, clean_border_ratio=0.0)

    # host: the reference's own image_set.cc around the oracle's minor loop
    r_res = oracle.ref_load_and_average(res, w_orig, 1, n_dec)
    r_psf = oracle.ref_load_and_average(psfs, w_orig, 1, n_dec)
    ocube = oracle.Cube(n, n, n_dec, 1, w_dec)
    _, r_std = oracle.ref_median_and_stddev_from_mad(oracle.integrate(ocube, r_res))
    os_ = oracle.Settings(threshold=float(5.0 * r_std), **kw)
    os_.spectral_mode, os_.n_terms, os_.frequencies, os_.fit_weights = 1, 2, f_dec, w_dec
    r_model = np.zeros_like(r_res)
    o = oracle.GenericClean(use_sub_minor_optimization=False).execute(ocube, os_, r_res, r_model,
                                                                      r_psf)
    assert 20 < o.iteration_number < 300  # stopped by the noise threshold
    r_out = oracle.ref_interpolate_model(r_model, 1, f_dec, w_dec, 2, f_orig)

    # device
    alg = rdl.GenericClean(use_sub_minor_optimization=False)
    d_orig = [torch.from_numpy(p).cuda() for p in res]
    d_psf_orig = [torch.from_numpy(p).cuda() for p in psfs]
    d_res = torch.zeros((n_dec, n, n), device="cuda")
    d_psf = torch.zeros((n_dec, n, n), device="cuda")
    d_model = torch.zeros((n_dec, n, n), device="cuda")
    alg.load_and_average(d_orig, w_orig, index, list(d_res))
    alg.load_and_average(d_psf_orig, w_orig, index, list(d_psf))
    cube = rdl.Cube(n, n, n_dec, 1, w_dec)
    integrated = torch.zeros((n, n), device="cuda")
    alg.integrate(cube, d_res, integrated)
    _, g_std, _ = alg.median_and_stddev_from_mad(integrated)
    assert np.float32(g_std) == r_std
    alg.apply(threshold=float(5.0 * np.float32(g_std)), **kw)
    alg.set_spectral_fitter(rdl.FIT_POLYNOMIAL, 2, f_dec, w_dec)
    g = alg.execute_major_iteration(cube, d_res, d_model, d_psf)
    assert g.iteration_number == o.iteration_number
    d_out = [torch.zeros((n, n), device="cuda") for _ in range(n_orig)]
    alg.interpolate_model(list(d_model), f_orig, d_out)
    assert np.array_equal(d_res.cpu().numpy(), r_res)
    assert np.array_equal(d_model.cpu().numpy(), r_model)
    assert np.array_equal(np.stack([t.cpu().numpy() for t in d_out]), r_out)


def test_interpolate_model_on_float_rounding_ties(rdl, oracle, torch_cuda):
    """Evenly spaced channels with unit weights give a term-solve operator of round numbers
    (0.1, 0.2, ...): a tenth of all fitted terms then sit on a float rounding tie that the last bit
    of the operator decides. Both sides build it with the same uncontracted operations."""
    torch = torch_cuda
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built")
    n_dec, n_orig, n = 4, 16, 96
    rng = np.random.default_rng(1)
    orig = 1.2e8 + 2.0e6 * np.arange(n_orig)
    fit = orig.reshape(n_dec, -1).mean(axis=1)
    models = (rng.standard_normal((n_dec, n, n)) + 3.0).astype(np.float32)
    for n_terms in (2, 3):
        want = oracle.ref_interpolate_model(models, 1, fit, np.ones(n_dec, np.float32), n_terms, orig)
        alg = rdl.GenericClean()
        alg.set_spectral_fitter(rdl.FIT_POLYNOMIAL, n_terms, fit, None)
        d_out = [torch.zeros((n, n), device="cuda") for _ in range(n_orig)]
        alg.interpolate_model([torch.from_numpy(m).cuda() for m in models], orig, d_out)
        got = np.stack([t.cpu().numpy() for t in d_out])
        assert np.array_equal(got, want), n_terms


@pytest.mark.parametrize("n_original,n_deconvolution,h,w", [(8, 3, 33, 47), (5, 5, 64, 64),
                                                            (40, 2, 32, 40), (6, 0, 17, 19)])
def test_load_and_average_psfs_bit_exact(rdl, oracle, torch_cuda, n_original, n_deconvolution, h, w):
    """ImageSet::LoadAndAveragePsfs (image_set.cc:142-205): double products, zero-weight
    channels included, against the reference's own source."""
    torch = torch_cuda
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(n_original)
    psfs = rng.standard_normal((n_original, h, w)).astype(np.float32)
    weights = rng.uniform(0.2, 3.0, n_original)
    weights[n_original // 2] = 0.0
    if n_deconvolution == n_original:
        weights[0] = 0.0  # a deconvolution channel without weight: scaled by 0
    index, n_dec = output_index_of(n_original, 1, n_deconvolution)
    want = oracle.ref_load_and_average_psfs(psfs, weights, n_deconvolution)
    alg = rdl.GenericClean()
    d_out = [torch.full((h, w), 5.0, device="cuda") for _ in range(n_dec)]
    alg.load_and_average([torch.from_numpy(p).cuda() for p in psfs], weights, index, d_out,
                         psfs=True)
    got = np.stack([t.cpu().numpy() for t in d_out])
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("n_orig,n_dec,terms", [(8, 2, 2), (6, 6, 2), (9, 3, 3)])
def test_radler_perform_major_iteration(rdl, oracle, torch_cuda, n_orig, n_dec, terms):
    """radler::Radler::Perform of the reference's OWN radler.cc (oracle/_ref), one major
    iteration: LoadAndAverage of residual and model, MedianAndStdDevFromMAD of the integrated
    residual, threshold = 5 sigma, LoadAndAveragePsfs, GenericClean::ExecuteMajorIteration with a
    polynomial fit, AssignAndStoreResidual, InterpolateAndStoreModel -- against the same steps
    on device planes through the C ABI. Stored residuals and models bit for bit."""
    from helpers import hogbom_case
    from perform_steps_cases import deconvolution_channels
    torch = torch_cuda
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built")
    n = 256
    res, psfs, w_orig, f_orig = hogbom_case(n, n_orig, seed=20 + n_orig)
    w_orig = w_orig.astype(np.float64)
    kw = dict(minor_loop_gain=0.1, major_loop_gain=0.95)
    r_res, r_model = res.copy(), np.zeros_like(res)
    r_model[:, 5, 7] = 0.25  # a component of an earlier major iteration
    start_model = r_model.copy()
    another, iterations = oracle.ref_radler_perform(
        r_res, r_model, psfs, f_orig, w_orig, n_dec, auto_threshold_sigma=5.0,
        minor_iteration_count=400, fit_terms=terms, **kw)
    assert 30 < iterations <= 400

    index, _ = output_index_of(n_orig, 1, n_dec)
    f_dec, w_dec = deconvolution_channels(f_orig, w_orig, n_dec)
    alg = rdl.GenericClean(use_sub_minor_optimization=False)
    upload = lambda planes: [torch.from_numpy(p).cuda() for p in planes]
    d_res = torch.zeros((n_dec, n, n), device="cuda")
    d_model = torch.zeros((n_dec, n, n), device="cuda")
    d_psf = torch.zeros((n_dec, n, n), device="cuda")
    alg.load_and_average(upload(res), w_orig, index, list(d_res))
    alg.load_and_average(upload(start_model), w_orig, index, list(d_model))
    alg.load_and_average(upload(psfs), w_orig, index, list(d_psf), psfs=True)
    cube = rdl.Cube(n, n, n_dec, 1, w_dec)
    integrated = torch.zeros((n, n), device="cuda")
    alg.integrate(cube, d_res, integrated)
    _, stddev, _ = alg.median_and_stddev_from_mad(integrated)
    # radler.cc:233-237, then DeconvolutionAlgorithm::SetThreshold(float)
    alg.apply(threshold=max(float(np.float32(stddev)) * 5.0, 0.0), max_iterations=400,
              clean_border_ratio=0.0, **kw)
    alg.set_spectral_fitter(rdl.FIT_POLYNOMIAL, terms, f_dec, w_dec)
    g = alg.execute_major_iteration(cube, d_res, d_model, d_psf)
    assert g.iteration_number == iterations
    assert g.another_iteration_required == another
    g_res = d_res.cpu().numpy()
    for ch, d in enumerate(index):  # AssignAndStoreResidual
        assert np.array_equal(g_res[d], r_res[ch]), ch
    if n_dec == n_orig:  # stored as they are (image_set.cc:210-215)
        g_model = d_model.cpu().numpy()
    else:
        d_out = [torch.zeros((n, n), device="cuda") for _ in range(n_orig)]
        alg.interpolate_model(list(d_model), f_orig, d_out)
        g_model = np.stack([t.cpu().numpy() for t in d_out])
    assert np.count_nonzero(r_model) > 4 * n_orig  # several sources, every channel
    assert np.array_equal(g_model, r_model)


def test_radler_perform_two_phase_auto_mask(rdl, oracle, torch_cuda):
    """Two major iterations of radler::Radler::Perform (the reference's own radler.cc) with
    auto-masking: the first cleans to auto_mask_sigma, the second doubles the minor-loop gain
    (:171-177), cleans to auto_threshold_sigma (:233-237) under the mask derived from the
    averaged model planes (:252-268). The device side keeps everything in device memory: the
    mask goes from rdl_auto_mask_from_model straight into rdl_set_clean_mask."""
    from helpers import hogbom_case
    from perform_steps_cases import deconvolution_channels
    torch = torch_cuda
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built")
    n, n_orig, n_dec, terms = 256, 8, 2, 2
    res, psfs, w_orig, f_orig = hogbom_case(n, n_orig, seed=28)
    w_orig = w_orig.astype(np.float64)
    r_res, r_model = res.copy(), np.zeros_like(res)
    per_major = oracle.ref_radler_perform(
        r_res, r_model, psfs, f_orig, w_orig, n_dec, auto_threshold_sigma=3.0,
        minor_loop_gain=0.1, major_loop_gain=1.0, minor_iteration_count=2000, fit_terms=terms,
        auto_mask_sigma=8.0, n_major_iterations=2)
    assert per_major[0][0] and per_major[0][1] > 50 and per_major[1][1] > per_major[0][1] + 50

    index, _ = output_index_of(n_orig, 1, n_dec)
    f_dec, w_dec = deconvolution_channels(f_orig, w_orig, n_dec)
    alg = rdl.GenericClean(use_sub_minor_optimization=False)
    alg.apply(max_iterations=2000, major_loop_gain=1.0, clean_border_ratio=0.0)
    alg.set_spectral_fitter(rdl.FIT_POLYNOMIAL, terms, f_dec, w_dec)
    cube = rdl.Cube(n, n, n_dec, 1, w_dec)
    # the caller's planes of the original channels, kept on the device across major iterations
    s_res = [torch.from_numpy(p).cuda() for p in res]
    s_model = [torch.zeros((n, n), device="cuda") for _ in range(n_orig)]
    s_psf = [torch.from_numpy(p).cuda() for p in psfs]
    d_res = torch.zeros((n_dec, n, n), device="cuda")
    d_model = torch.zeros((n_dec, n, n), device="cuda")
    d_psf = torch.zeros((n_dec, n, n), device="cuda")
    integrated = torch.zeros((n, n), device="cuda")
    mask = torch.zeros((n, n), dtype=torch.uint8, device="cuda")
    finished = False
    for major in range(2):
        alg.load_and_average(s_res, w_orig, index, list(d_res))
        alg.load_and_average(s_model, w_orig, index, list(d_model))
        alg.integrate(cube, d_res, integrated)
        _, stddev, _ = alg.median_and_stddev_from_mad(integrated)
        stddev = float(np.float32(stddev))
        if finished:
            alg.apply(minor_loop_gain=min(1.0, 0.1 * 2.0), threshold=max(stddev * 3.0, 0.0))
            alg.auto_mask_from_model(list(d_model), mask)
            alg.set_clean_mask(mask)
        else:
            alg.apply(minor_loop_gain=0.1, threshold=max(stddev * 8.0, 0.0))
        alg.load_and_average(s_psf, w_orig, index, list(d_psf), psfs=True)
        g = alg.execute_major_iteration(cube, d_res, d_model, d_psf)
        another = g.another_iteration_required
        if not another and not finished:
            finished, another = True, True
        assert (another, g.iteration_number) == per_major[major], major
        for ch, d in enumerate(index):  # AssignAndStoreResidual
            s_res[ch].copy_(d_res[d])
        alg.interpolate_model(list(d_model), f_orig, s_model)
    assert finished and 0 < int(mask.sum()) < 100
    assert np.array_equal(np.stack([t.cpu().numpy() for t in s_res]), r_res)
    assert np.array_equal(np.stack([t.cpu().numpy() for t in s_model]), r_model)


def test_rms_factor_image_from_a_device_plane(rdl, torch_cuda):
    """rdl_set_rms_factor_image / rdl_set_clean_mask take device planes as well: same peaks as
    with the host copies."""
    from helpers import hogbom_case
    torch = torch_cuda
    n = 128
    res, psfs, _, _ = hogbom_case(n, 1, seed=3)
    rng = np.random.default_rng(3)
    factor = rng.uniform(0.5, 1.0, (n, n)).astype(np.float32)
    mask = (rng.uniform(size=(n, n)) > 0.3).astype(np.uint8)
    traces = []
    for device in (False, True):
        alg = rdl.GenericClean(use_sub_minor_optimization=False)
        alg.apply(max_iterations=60, clean_border_ratio=0.0, trace_capacity=1 << 10)
        alg.set_rms_factor_image(torch.from_numpy(factor).cuda() if device else factor)
        alg.set_clean_mask(torch.from_numpy(mask).cuda() if device else mask)
        d_res, d_model = torch.from_numpy(res.copy()).cuda(), torch.zeros((1, n, n), device="cuda")
        alg.execute_major_iteration(rdl.Cube(n, n, 1, 1, None), d_res, d_model,
                                    torch.from_numpy(psfs).cuda())
        traces.append(alg.trace())
    assert len(traces[0]) == 60
    for k in ("x", "y", "value"):
        assert np.array_equal(traces[0][k], traces[1][k])


def test_radler_perform_with_the_sub_minor_loop(rdl, oracle, torch_cuda):
    """The same major iteration through the reference's DEFAULT GenericClean (sub-minor loop,
    CorrectResidualDirty through FFT convolutions): components and model bit for bit, the stored
    residuals within 1e-5 of the peak (FFT round-off, as in the path-A tests)."""
    from helpers import hogbom_case
    from perform_steps_cases import deconvolution_channels
    torch = torch_cuda
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built")
    n, n_orig, n_dec, terms = 256, 8, 2, 2
    res, psfs, w_orig, f_orig = hogbom_case(n, n_orig, seed=28)
    w_orig = w_orig.astype(np.float64)
    kw = dict(minor_loop_gain=0.1, major_loop_gain=0.9)
    r_res, r_model = res.copy(), np.zeros_like(res)
    another, iterations = oracle.ref_radler_perform(
        r_res, r_model, psfs, f_orig, w_orig, n_dec, auto_threshold_sigma=5.0,
        minor_iteration_count=2000, fit_terms=terms, use_sub_minor_optimization=True, **kw)
    assert iterations > 50

    index, _ = output_index_of(n_orig, 1, n_dec)
    f_dec, w_dec = deconvolution_channels(f_orig, w_orig, n_dec)
    alg = rdl.GenericClean(use_sub_minor_optimization=True)
    upload = lambda planes: [torch.from_numpy(p).cuda() for p in planes]
    d_res = torch.zeros((n_dec, n, n), device="cuda")
    d_model = torch.zeros((n_dec, n, n), device="cuda")
    d_psf = torch.zeros((n_dec, n, n), device="cuda")
    alg.load_and_average(upload(res), w_orig, index, list(d_res))
    alg.load_and_average(upload(psfs), w_orig, index, list(d_psf), psfs=True)
    cube = rdl.Cube(n, n, n_dec, 1, w_dec)
    integrated = torch.zeros((n, n), device="cuda")
    alg.integrate(cube, d_res, integrated)
    _, stddev, _ = alg.median_and_stddev_from_mad(integrated)
    alg.apply(threshold=max(float(np.float32(stddev)) * 5.0, 0.0), max_iterations=2000,
              clean_border_ratio=0.0, **kw)
    alg.set_spectral_fitter(rdl.FIT_POLYNOMIAL, terms, f_dec, w_dec)
    g = alg.execute_major_iteration(cube, d_res, d_model, d_psf)
    assert (g.another_iteration_required, g.iteration_number) == (another, iterations)
    d_out = [torch.zeros((n, n), device="cuda") for _ in range(n_orig)]
    alg.interpolate_model(list(d_model), f_orig, d_out)
    assert np.array_equal(np.stack([t.cpu().numpy() for t in d_out]), r_model)
    g_res, peak = d_res.cpu().numpy(), np.abs(res).max()
    for ch, d in enumerate(index):
        assert np.abs(g_res[d] - r_res[ch]).max() <= 1e-5 * peak, ch


def test_radler_perform_multiscale(rdl, oracle, torch_cuda):
    """One major iteration of radler::Radler::Perform with AlgorithmType::kMultiscale (4
    original channels averaged to 2, polynomial fit) against the device steps around
    MultiScaleAlgorithm: same component count and stop reason, stored residuals and interpolated
    models within 1e-4 of their peak (the scale convolutions are FFTs on both sides)."""
    from helpers import multiscale_case
    from perform_steps_cases import deconvolution_channels
    torch = torch_cuda
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built")
    n, n_orig, n_dec, terms, beam = 256, 4, 2, 2, 6.0
    res, psfs, w_orig, f_orig = multiscale_case(n, n_orig, seed=5)
    w_orig = w_orig.astype(np.float64)
    kw = dict(minor_loop_gain=0.1, major_loop_gain=0.8)
    r_res, r_model = res.copy(), np.zeros_like(res)
    another, iterations = oracle.ref_radler_perform(
        r_res, r_model, psfs, f_orig, w_orig, n_dec, auto_threshold_sigma=5.0,
        minor_iteration_count=500, fit_terms=terms, multiscale=True, beam_size=beam, **kw)
    assert iterations > 40

    index, _ = output_index_of(n_orig, 1, n_dec)
    f_dec, w_dec = deconvolution_channels(f_orig, w_orig, n_dec)
    alg = rdl.MultiScaleAlgorithm(beam, 1.0)
    upload = lambda planes: [torch.from_numpy(p).cuda() for p in planes]
    d_res = torch.zeros((n_dec, n, n), device="cuda")
    d_model = torch.zeros((n_dec, n, n), device="cuda")
    d_psf = torch.zeros((n_dec, n, n), device="cuda")
    alg.load_and_average(upload(res), w_orig, index, list(d_res))
    alg.load_and_average(upload(psfs), w_orig, index, list(d_psf), psfs=True)
    cube = rdl.Cube(n, n, n_dec, 1, w_dec)
    integrated = torch.zeros((n, n), device="cuda")
    alg.integrate(cube, d_res, integrated)
    _, stddev, _ = alg.median_and_stddev_from_mad(integrated)
    alg.apply(threshold=max(float(np.float32(stddev)) * 5.0, 0.0), max_iterations=500,
              clean_border_ratio=0.0, **kw)
    alg.set_spectral_fitter(rdl.FIT_POLYNOMIAL, terms, f_dec, w_dec)
    g = alg.execute_major_iteration(cube, d_res, d_model, d_psf)
    assert (g.another_iteration_required, g.iteration_number) == (another, iterations)
    d_out = [torch.zeros((n, n), device="cuda") for _ in range(n_orig)]
    alg.interpolate_model(list(d_model), f_orig, d_out)
    g_model = np.stack([t.cpu().numpy() for t in d_out])
    assert np.abs(g_model - r_model).max() <= 1e-4 * np.abs(r_model).max()
    g_res, peak = d_res.cpu().numpy(), np.abs(res).max()
    for ch, d in enumerate(index):
        assert np.abs(g_res[d] - r_res[ch]).max() <= 1e-4 * peak, ch


def test_radler_perform_iuwt(rdl, oracle, torch_cuda):
    """One major iteration of radler::Radler::Perform with AlgorithmType::kIuwt on one channel:
    same iteration count and flag, images within 1e-5 of the peak (tests/
    test_gpu_iuwt_deconvolution.py says why not bit for bit)."""
    from helpers import multiscale_case
    torch = torch_cuda
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built")
    n = 256
    res, psfs, _, freqs = multiscale_case(n, 1, seed=3)
    kw = dict(minor_loop_gain=0.1, major_loop_gain=0.8)
    r_res, r_model = res.copy(), np.zeros_like(res)
    another, iterations = oracle.ref_radler_perform(
        r_res, r_model, psfs, freqs, np.ones(1), 1, auto_threshold_sigma=3.0,
        minor_iteration_count=12, fit_terms=0, iuwt=True, **kw)
    assert iterations > 3 and np.abs(r_model).max() > 0

    alg = rdl.DeconvolutionAlgorithm(rdl.IUWT)
    d_res = torch.zeros((1, n, n), device="cuda")
    d_psf = torch.zeros((1, n, n), device="cuda")
    d_model = torch.zeros((1, n, n), device="cuda")
    alg.load_and_average([torch.from_numpy(res[0]).cuda()], [1.0], [0], list(d_res))
    alg.load_and_average([torch.from_numpy(psfs[0]).cuda()], [1.0], [0], list(d_psf), psfs=True)
    cube = rdl.Cube(n, n, 1, 1, None)
    integrated = torch.zeros((n, n), device="cuda")
    alg.integrate(cube, d_res, integrated)
    _, stddev, _ = alg.median_and_stddev_from_mad(integrated)
    alg.apply(threshold=max(float(np.float32(stddev)) * 3.0, 0.0), max_iterations=12,
              clean_border_ratio=0.0, **kw)
    g = alg.execute_major_iteration(cube, d_res, d_model, d_psf)
    assert (g.another_iteration_required, g.iteration_number) == (another, iterations)
    peak = np.abs(res).max()
    assert np.abs(d_res.cpu().numpy() - r_res).max() <= 1e-5 * peak
    assert np.abs(d_model.cpu().numpy() - r_model).max() <= 1e-5 * peak
