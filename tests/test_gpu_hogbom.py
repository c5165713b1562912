"""GPU parity: peak finder, SubtractImage and GenericClean path B (FindPeak -> SubtractImage
loop, generic_clean.cc:163-207) through the C ABI vs the CPU oracle on identical inputs.
Peak positions / iteration counts bit-exact; images within 1e-5 (in practice bit-equal)."""
import numpy as np
import pytest

from helpers import hogbom_case, plus_psf_fixture

pytestmark = pytest.mark.gpu
REL_TOL = 1e-5  # north_star: model and residual agree within 1e-5 relative (fp32)


def _rel_err(a, b):
    return float(np.abs(a.astype(np.float64) - b).max() / max(np.abs(b).max(), 1e-30))


def test_find_peak_kernel_matches_oracle(rdl, oracle, torch_cuda):
    torch = torch_cuda
    rng = np.random.default_rng(5)
    alg = rdl.GenericClean(use_sub_minor_optimization=False)
    for (h, w, c) in [(1, 1, 1), (3, 5, 1), (6, 6, 1), (2, 18, 1), (64, 64, 1), (37, 129, 3),
                      (256, 256, 4), (1024, 1024, 1), (512, 2048, 2)]:
        imgs = rng.standard_normal((c, h, w)).astype(np.float32)
        weights = rng.uniform(0.5, 1.5, c).astype(np.float32) if c > 1 else None
        ocube = oracle.Cube(w, h, c, 1, weights)
        cube = rdl.Cube(w, h, c, 1, weights)
        dev = torch.from_numpy(imgs).cuda()
        for allow_negative in (True, False):
            for (hb, vb) in [(0, 0), (min(2, (w - 1) // 2), min(3, (h - 1) // 2))]:
                mask = rng.integers(0, 2, (h, w)).astype(np.uint8)
                integ = oracle.integrate(ocube, imgs, square=False)
                for m in (None, mask):
                    want = oracle.find_peak(integ, allow_negative, 0, h, hb, vb, m)
                    got = alg.find_peak(cube, dev, allow_negative=allow_negative, h_border=hb,
                                        v_border=vb,
                                        mask=None if m is None else torch.from_numpy(m).cuda())
                    assert got == want, (h, w, c, allow_negative, hb, vb, m is not None)


def test_find_peak_ties_and_nan(rdl, oracle, torch_cuda):
    torch = torch_cuda
    alg = rdl.GenericClean(use_sub_minor_optimization=False)
    img = np.zeros((1, 300, 500), np.float32)
    cube = rdl.Cube(500, 300)
    assert alg.find_peak(cube, torch.from_numpy(img).cuda())[0] is None  # nothing above FLT_MIN
    img[0, 250, 17] = img[0, 10, 400] = img[0, 10, 401] = 3.0            # first index wins
    img[0, 0, 0] = np.nan
    img[0, 299, 499] = -3.0
    assert alg.find_peak(cube, torch.from_numpy(img).cuda()) == (3.0, 400, 10)
    img[0, 5, 5] = -3.0
    assert alg.find_peak(cube, torch.from_numpy(img).cuda()) == (-3.0, 5, 5)
    assert alg.find_peak(cube, torch.from_numpy(img).cuda(), allow_negative=False) == (3.0, 400, 10)


def test_integrate_kernel_matches_oracle(rdl, oracle, torch_cuda):
    torch = torch_cuda
    rng = np.random.default_rng(6)
    alg = rdl.GenericClean(use_sub_minor_optimization=False)
    cases = [dict(n_channels=1, n_pol=1), dict(n_channels=3, n_pol=1, weights=[1.0, 0.0, 2.5]),
             dict(n_channels=1, n_pol=2, pol_norm=0.5), dict(n_channels=2, n_pol=4, pol_norm=0.5,
                                                            pol_linked=[1, 0, 0, 1], weights=[0.7, 1.3]),
             dict(n_channels=2, n_pol=2, pol_norm=0.5, squared_joins=True, weights=[1.0, 2.0])]
    for kw in cases:
        for (h, w) in [(33, 65), (128, 128)]:
            kw2 = dict(kw)
            for k in ("weights", "pol_linked"):
                if k in kw2:
                    kw2[k] = np.asarray(kw2[k], np.float32 if k == "weights" else np.uint8)
            cube, ocube = rdl.Cube(w, h, **kw2), oracle.Cube(w, h, **kw2)
            imgs = rng.standard_normal((cube.n_images, h, w)).astype(np.float32)
            dev = torch.from_numpy(imgs).cuda()
            for square in (False, True):
                dest = torch.zeros((h, w), dtype=torch.float32, device="cuda")
                alg.integrate(cube, dev, dest, square=square)
                torch.cuda.synchronize()
                want = oracle.integrate(ocube, imgs, square=square)
                assert np.array_equal(dest.cpu().numpy(), want), (kw, square)


def test_subtract_image_kernel_matches_oracle(rdl, oracle, torch_cuda):
    torch = torch_cuda
    rng = np.random.default_rng(9)
    alg = rdl.GenericClean(use_sub_minor_optimization=False)
    for (h, w, c) in [(5, 7, 1), (64, 64, 2), (130, 258, 3), (512, 512, 4), (97, 33, 1)]:
        imgs = rng.standard_normal((c, h, w)).astype(np.float32)
        psfs = rng.standard_normal((c, h, w)).astype(np.float32)
        cube = rdl.Cube(w, h, c, 1)
        for (x, y) in [(0, 0), (w - 1, h - 1), (w // 2, h // 2), (w // 3, h - 2), (w - 2, 1)]:
            f = rng.standard_normal(c).astype(np.float32)
            dev = torch.from_numpy(imgs).cuda()
            alg.subtract_image(cube, dev, torch.from_numpy(psfs).cuda(), x, y, f)
            torch.cuda.synchronize()
            want = imgs.copy()
            for i in range(c):
                oracle.subtract_image(want[i], psfs[i], x, y, float(f[i]))
            assert np.array_equal(dev.cpu().numpy(), want), (h, w, c, x, y)


def _run_both(rdl, oracle, res, psfs, weights, settings, sub_minor=False, device_resident=False,
              fit=None, n_pol=1, pol_norm=1.0, squared_joins=False, pol_linked=None):
    """res: [n_channels * n_pol, H, W] ordered [channel][pol]; psfs: [n_channels, H, W]."""
    c, h, w = res.shape
    c //= n_pol
    cube_args = (w, h, c, n_pol, weights, pol_linked, pol_norm, squared_joins)
    o_res, o_model = res.copy(), np.zeros_like(res)
    oalg = oracle.GenericClean(use_sub_minor_optimization=sub_minor)
    os_ = oracle.Settings(record_trace=True, **settings)
    if fit:
        os_.spectral_mode, os_.n_terms, os_.frequencies, os_.fit_weights = 1, fit[0], fit[1], fit[2]
    o = oalg.execute(oracle.Cube(*cube_args), os_, o_res, o_model, psfs)

    g_res, g_model = res.copy(), np.zeros_like(res)
    alg = rdl.GenericClean(use_sub_minor_optimization=sub_minor)
    alg.apply(threshold=settings.get("threshold", 0.0),
              minor_loop_gain=settings.get("minor_loop_gain", 0.1),
              major_loop_gain=settings.get("major_loop_gain", 1.0),
              clean_border_ratio=settings.get("clean_border_ratio", 0.05),
              max_iterations=settings.get("max_iterations", 500),
              allow_negative_components=settings.get("allow_negative", True),
              stop_on_negative_components=settings.get("stop_on_negative", False),
              divergence_limit=settings.get("divergence_limit", 4.0),
              trace_capacity=1 << 20)
    if "clean_mask" in settings:
        alg.set_clean_mask(settings["clean_mask"])
    if settings.get("rms_factor") is not None:
        alg.set_rms_factor_image(settings["rms_factor"])
    if fit:
        alg.set_spectral_fitter(rdl.FIT_POLYNOMIAL, fit[0], fit[1], fit[2], n_pol)
    cube = rdl.Cube(*cube_args)
    if device_resident:
        import torch
        d_res, d_model = torch.from_numpy(g_res).cuda(), torch.from_numpy(g_model).cuda()
        g = alg.execute_major_iteration(cube, d_res, d_model, torch.from_numpy(psfs).cuda())
        g_res, g_model = d_res.cpu().numpy(), d_model.cpu().numpy()
    else:
        g = alg.execute_major_iteration(cube, g_res, g_model, psfs)
    return o, (o_res, o_model), g, (g_res, g_model), alg


def _assert_same(o, o_imgs, g, g_imgs, alg):
    assert g.iteration_number == o.iteration_number
    tr = alg.trace()
    assert len(tr) == len(o.trace)
    for k in ("iteration", "x", "y"):
        assert np.array_equal(tr[k], o.trace[k]), k
    assert np.array_equal(tr["value"].view(np.uint32), o.trace["value"].view(np.uint32))
    assert (g.starting_peak_value is None) == (o.starting_peak_value is None)
    if o.starting_peak_value is not None:
        assert g.starting_peak_value == o.starting_peak_value
    assert g.final_peak_value == o.final_peak_value
    assert g.another_iteration_required == o.another_iteration_required
    assert g.is_diverging == o.is_diverging
    assert _rel_err(g_imgs[0], o_imgs[0]) <= REL_TOL
    assert _rel_err(g_imgs[1], o_imgs[1]) <= REL_TOL


@pytest.mark.parametrize("shift", [(0, 0), (7, -11), (-9, 15), (-9, 23)])
def test_path_b_point_source_fixture(rdl, oracle, torch_cuda, hogbom_loop_flavour, shift):
    # cpp/test/test_radler.cc:98-172 through the CUDA path
    psf, res, src = plus_psf_fixture(*shift)
    st = dict(threshold=1e-8, max_iterations=1000, clean_border_ratio=0.0)
    o, oi, g, gi, alg = _run_both(rdl, oracle, res, psf, None, st)
    _assert_same(o, oi, g, gi, alg)
    assert np.abs(gi[0]).max() < 2e-6
    assert gi[1].flat[src] == pytest.approx(2.5, rel=1e-6)


@pytest.mark.parametrize("n,c,iters,device_resident", [(256, 1, 500, False), (250, 1, 300, True),
                                                       (512, 4, 400, True), (1024, 1, 2000, True)])
def test_path_b_matches_oracle(rdl, oracle, torch_cuda, hogbom_loop_flavour, n, c, iters, device_resident):
    res, psfs, weights, freqs = hogbom_case(n, c)
    st = dict(threshold=0.0, max_iterations=iters, clean_border_ratio=0.0, minor_loop_gain=0.1)
    o, oi, g, gi, alg = _run_both(rdl, oracle, res, psfs, weights, st,
                                  device_resident=device_resident)
    assert o.iteration_number == iters
    _assert_same(o, oi, g, gi, alg)


def test_path_b_variants(rdl, oracle, torch_cuda, hogbom_loop_flavour):
    res, psfs, weights, freqs = hogbom_case(256, 3)
    rng = np.random.default_rng(2)
    mask = (rng.uniform(size=(256, 256)) < 0.5).astype(np.uint8)
    for st in [dict(threshold=0.05, max_iterations=5000, clean_border_ratio=0.05),
               dict(max_iterations=300, major_loop_gain=0.8, clean_border_ratio=0.1),
               dict(max_iterations=300, allow_negative=False, clean_border_ratio=0.0),
               dict(max_iterations=300, stop_on_negative=True, clean_border_ratio=0.0),
               dict(max_iterations=200, clean_mask=mask, clean_border_ratio=0.0),
               dict(max_iterations=0, clean_border_ratio=0.0)]:
        o, oi, g, gi, alg = _run_both(rdl, oracle, res, psfs, weights, st)
        _assert_same(o, oi, g, gi, alg)
    # joined channels with a 2-term polynomial fit (deconvolution_algorithm.cc:29-46)
    st = dict(max_iterations=300, clean_border_ratio=0.0)
    o, oi, g, gi, alg = _run_both(rdl, oracle, res, psfs, weights, st,
                                  fit=(2, freqs, weights))
    _assert_same(o, oi, g, gi, alg)


def test_path_b_divergence(rdl, oracle, torch_cuda, hogbom_loop_flavour):
    # a PSF with peak -1.5 makes every subtraction grow the peak (test_divergence.cc idea)
    res, psfs, weights, _ = hogbom_case(128, 1)
    psfs = psfs * np.float32(-1.5)
    st = dict(max_iterations=1000, clean_border_ratio=0.0, minor_loop_gain=0.5)
    o, oi, g, gi, alg = _run_both(rdl, oracle, res, psfs, None, st)
    assert o.is_diverging
    _assert_same(o, oi, g, gi, alg)


@pytest.mark.parametrize("c", [1, 3])
def test_path_b_host_buffers_keep_the_model_on_the_host(rdl, oracle, torch_cuda, hogbom_loop_flavour, c):
    # RDL_MEM_HOST, GenericClean without the sub-minor loop: the model planes are not uploaded; the
    # call logs its components on the device and applies them to the caller's planes in order.
    # Same bits as the device-resident call, from a non-zero model, over two major iterations
    # (components falling on the same pixel several times).
    torch = torch_cuda
    n = 192
    res, psfs, weights, _ = hogbom_case(n, c, seed=17)
    rng = np.random.default_rng(3)
    model0 = rng.normal(0, 0.1, res.shape).astype(np.float32)
    cube = rdl.Cube(n, n, c, 1, weights)

    def run(host):
        alg = rdl.GenericClean(use_sub_minor_optimization=False)
        if host:
            r, m = res.copy(), model0.copy()
            p = psfs
        else:
            r, m = torch.from_numpy(res.copy()).cuda(), torch.from_numpy(model0.copy()).cuda()
            p = torch.from_numpy(psfs).cuda()
        out = []
        for call in range(2):
            alg.apply(max_iterations=150 * (call + 1), clean_border_ratio=0.0, major_loop_gain=0.9,
                      trace_capacity=1000)
            g = alg.execute_major_iteration(cube, r, m, p)
            out.append(g.iteration_number)
        if not host:
            r, m = r.cpu().numpy(), m.cpu().numpy()
        return out, r, m

    it_h, r_h, m_h = run(True)
    it_d, r_d, m_d = run(False)
    assert it_h == it_d and it_h[1] > it_h[0] > 5
    assert np.array_equal(r_h, r_d)
    assert np.array_equal(m_h, m_d)
    assert np.count_nonzero(m_h != model0) < it_h[1] * c  # some pixels were hit more than once
