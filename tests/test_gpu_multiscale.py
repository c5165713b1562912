"""GPU parity of MultiScaleAlgorithm::ExecuteMajorIteration (multiscale_algorithm.cc:183-576)
through the C ABI vs the CPU oracle: scale selection, component positions and iteration
counts identical; component values, model and residual within 1e-5 of the image peak (every
non-zero scale goes through FFT convolutions, which are not bit-reproducible across
implementations)."""
import math

import numpy as np
import pytest

from helpers import multiscale_case, plus_psf_fixture

pytestmark = pytest.mark.gpu
REL_TOL = 1e-5
NEAR_TIE_RTOL = 5e-5      # value gap under which two adjacent pixels count as tied
NEAR_TIE_IMAGE_TOL = 2e-3  # image tolerance once a tie was broken differently


def run_both(rdl, oracle, res, psfs, weights, settings, beam=6.0, ms_kwargs=None, fit=None,
             automask=None, device_resident=True, algs=None, n_pol=1, pol_norm=1.0,
             squared_joins=False):
    import torch
    ms_kwargs = ms_kwargs or {}
    c, h, w = res.shape
    n_img = c
    c //= n_pol
    cube_args = (w, h, c, n_pol, weights, None, pol_norm, squared_joins)
    if algs is None:
        oalg = oracle.MultiScale(beam_size=beam, pixel_scale=1.0, **ms_kwargs)
        galg = rdl.MultiScaleAlgorithm(beam, 1.0, **ms_kwargs)
    else:
        oalg, galg = algs
    os_ = oracle.Settings(record_trace=True, **settings)
    if fit:
        os_.spectral_mode, os_.n_terms, os_.frequencies, os_.fit_weights = 1, fit[0], fit[1], fit[2]
    if automask:
        oalg.set_automask_mode(*automask)
        galg.set_auto_mask_mode(*automask)
    o_res, o_model = res.copy(), np.zeros_like(res)
    o = oalg.execute(oracle.Cube(*cube_args), os_, o_res, o_model, psfs)

    galg.apply(threshold=settings.get("threshold", 0.0),
               minor_loop_gain=settings.get("minor_loop_gain", 0.1),
               major_loop_gain=settings.get("major_loop_gain", 1.0),
               clean_border_ratio=settings.get("clean_border_ratio", 0.05),
               max_iterations=settings.get("max_iterations", 500),
               allow_negative_components=settings.get("allow_negative", True),
               stop_on_negative_components=settings.get("stop_on_negative", False),
               divergence_limit=settings.get("divergence_limit", 4.0),
               iteration_number=settings.get("iteration_number", 0),
               trace_capacity=1 << 20)
    if "clean_mask" in settings:
        galg.set_clean_mask(settings["clean_mask"])
    if settings.get("rms_factor") is not None:
        galg.set_rms_factor_image(settings["rms_factor"])
    if fit:
        galg.set_spectral_fitter(rdl.FIT_POLYNOMIAL, fit[0], fit[1], fit[2], n_pol)
    cube = rdl.Cube(*cube_args)
    if device_resident:
        d_res, d_model = torch.from_numpy(res.copy()).cuda(), torch.zeros((n_img, h, w), device="cuda")
        g = galg.execute_major_iteration(cube, d_res, d_model, torch.from_numpy(psfs).cuda())
        g_res, g_model = d_res.cpu().numpy(), d_model.cpu().numpy()
    else:
        g_res, g_model = res.copy(), np.zeros_like(res)
        g = galg.execute_major_iteration(cube, g_res, g_model, psfs)
    return o, (o_res, o_model), g, (g_res, g_model), (oalg, galg)


def check(o, oi, g, gi, algs, value_rtol=2e-4, minor_before=0, image_scale=None):
    oalg, galg = algs
    assert g.iteration_number == o.iteration_number
    # the oracle's counter runs over the object's lifetime, the C ABI's getter per call
    assert galg.minor_iteration_count == oalg.minor_iterations() - minor_before
    tr = galg.trace()
    assert len(tr) == len(o.trace)
    assert np.array_equal(tr["iteration"], o.trace["iteration"])
    assert np.array_equal(tr["scale"], o.trace["scale"])
    # Positions are bit-exact except at VERIFIED near-ties: two adjacent pixels of a
    # scale-convolved image whose values differ by less than the FFT round-off of the two
    # implementations (the reference's FFTW would not reproduce the oracle's bits either).
    bad = np.nonzero((tr["x"] != o.trace["x"]) | (tr["y"] != o.trace["y"]))[0]
    for b in bad:
        dx = abs(int(tr["x"][b]) - int(o.trace["x"][b]))
        dy = abs(int(tr["y"][b]) - int(o.trace["y"][b]))
        dv = abs(float(tr["value"][b]) - float(o.trace["value"][b]))
        assert tr["scale"][b] > 0 and dx <= 1 and dy <= 1 and \
            dv <= NEAR_TIE_RTOL * abs(float(o.trace["value"][b])), \
            f"first divergence at trace entry {b}: gpu {tr[b]} oracle {o.trace[b]}"
    # an outer-pass record and the first component of its sub-minor run are one event
    near_ties = len({int(tr["iteration"][b]) for b in bad})
    assert near_ties <= max(4, len(tr) // 50), f"{near_ties} near-ties in {len(tr)} entries"
    assert np.allclose(tr["value"], o.trace["value"], rtol=value_rtol, atol=1e-7)
    assert g.another_iteration_required == o.another_iteration_required
    assert g.is_diverging == o.is_diverging
    assert g.final_peak_value == pytest.approx(o.final_peak_value, rel=value_rtol, abs=1e-7)
    scale = image_scale or max(np.abs(oi[0]).max(), np.abs(oi[1]).max(), 1e-30)
    # a near-tie moves one component by one pixel: the images then differ by that
    # component's (scale-kernel shaped) footprint, not by round-off
    tol = REL_TOL if near_ties == 0 else NEAR_TIE_IMAGE_TOL
    assert np.abs(gi[0].astype(np.float64) - oi[0]).max() / scale <= tol
    assert np.abs(gi[1].astype(np.float64) - oi[1]).max() / scale <= tol
    gs, os_ = galg.scale_infos(), oalg.scale_infos()
    assert len(gs) == len(os_)
    for a, b in zip(gs, os_):
        assert a["scale"] == b["scale"] and a["kernel_peak"] == b["kernel_peak"]
        assert a["bias_factor"] == b["bias_factor"]
        assert a["n_components_cleaned"] == b["n_components_cleaned"]
        assert a["psf_peak"] == pytest.approx(b["psf_peak"], rel=1e-5)
        assert a["total_flux_cleaned"] == pytest.approx(b["total_flux_cleaned"], rel=1e-3, abs=1e-6)
        assert bool(a["is_active"]) == bool(b["is_active"])
    return near_ties


@pytest.mark.parametrize("shift", [(0, 0), (7, -11), (-9, 15), (-9, 23)])
def test_multiscale_point_source_fixture(rdl, oracle, torch_cuda, shift):
    # cpp/test/test_radler.cc:98-172 (beam size 0 -> 1 pixel), python/test/test_radler.py:240-282
    psf, res, src = plus_psf_fixture(*shift)
    st = dict(threshold=1e-8, max_iterations=1000, clean_border_ratio=0.0)
    o, oi, g, gi, algs = run_both(rdl, oracle, res, psf, None, st, beam=0.0, device_resident=False)
    assert np.abs(gi[0]).max() < 2e-6
    assert gi[1].flat[src] == pytest.approx(2.5, rel=1e-6)
    m = gi[1].copy()
    m.flat[src] = 0
    assert np.abs(m).max() < 2e-6
    assert g.iteration_number == o.iteration_number


@pytest.mark.parametrize("n,c,iters,mgain,max_scales",
                         [(256, 1, 2000, 0.8, 4), (384, 2, 3000, 0.8, 0), (512, 1, 5000, 0.85, 5)])
def test_multiscale_matches_oracle(rdl, oracle, torch_cuda, n, c, iters, mgain, max_scales):
    res, psfs, weights, freqs = multiscale_case(n, c)
    st = dict(threshold=0.0, max_iterations=iters, clean_border_ratio=0.0, major_loop_gain=mgain)
    o, oi, g, gi, algs = run_both(rdl, oracle, res, psfs, weights, st,
                                  ms_kwargs=dict(max_scales=max_scales))
    assert o.iteration_number > 20
    assert len({int(s) for s in o.trace["scale"]}) > 1  # several scales were cleaned
    check(o, oi, g, gi, algs)


def test_multiscale_joined_polynomial_fit(rdl, oracle, torch_cuda):
    # K4-shaped: joined channels + 2-term polynomial spectral fit
    res, psfs, weights, freqs = multiscale_case(256, 4)
    st = dict(threshold=0.0, max_iterations=1500, clean_border_ratio=0.0, major_loop_gain=0.8)
    o, oi, g, gi, algs = run_both(rdl, oracle, res, psfs, weights, st, fit=(2, freqs, weights),
                                  ms_kwargs=dict(max_scales=4))
    check(o, oi, g, gi, algs)


def test_multiscale_variants(rdl, oracle, torch_cuda):
    res, psfs, weights, freqs = multiscale_case(256, 1, seed=4)
    rng = np.random.default_rng(2)
    mask = (rng.uniform(size=(256, 256)) < 0.6).astype(np.uint8)
    for st, kw in [
            (dict(threshold=0.02, max_iterations=5000, clean_border_ratio=0.05, major_loop_gain=0.9), {}),
            (dict(max_iterations=400, allow_negative=False, clean_border_ratio=0.0, major_loop_gain=0.8), {}),
            (dict(max_iterations=400, clean_mask=mask, clean_border_ratio=0.0, major_loop_gain=0.8), {}),
            (dict(max_iterations=600, clean_border_ratio=0.0, major_loop_gain=0.8),
             dict(shape=1, max_scales=3)),
            (dict(max_iterations=600, clean_border_ratio=0.0, major_loop_gain=0.8),
             dict(scale_list=[0.0, 10.0, 30.0])),
            (dict(max_iterations=0, clean_border_ratio=0.0), {})]:
        o, oi, g, gi, algs = run_both(rdl, oracle, res, psfs, weights, st, ms_kwargs=kw)
        check(o, oi, g, gi, algs)


def test_multiscale_automask_two_phase(rdl, oracle, torch_cuda):
    # K2-shaped: phase 1 tracks per-scale masks, phase 2 cleans inside them (radler.cc:254-266)
    res, psfs, weights, freqs = multiscale_case(256, 1, seed=7)
    sigma = 1.4826 * np.median(np.abs(res - np.median(res)))
    st1 = dict(threshold=float(4 * sigma), max_iterations=3000, clean_border_ratio=0.0,
               major_loop_gain=0.8)
    o, oi, g, gi, algs = run_both(rdl, oracle, res, psfs, weights, st1, automask=(True, False),
                                  ms_kwargs=dict(max_scales=4))
    check(o, oi, g, gi, algs)
    oalg, galg = algs
    n_scales = len(oalg.scale_infos())
    for i in range(n_scales):
        assert np.array_equal(galg.scale_mask(i, 256, 256), oalg.scale_mask(i, 256, 256)), i
    # phase 2 continues from the oracle's phase-1 residual so both start identically
    minor_before = oalg.minor_iterations()
    st2 = dict(threshold=float(1 * sigma), max_iterations=6000, clean_border_ratio=0.0,
               major_loop_gain=0.8, iteration_number=o.iteration_number)
    o2, oi2, g2, gi2, _ = run_both(rdl, oracle, oi[0], psfs, weights, st2, automask=(False, True),
                                   algs=algs)
    check(o2, oi2, g2, gi2, algs, minor_before=minor_before)


def test_multiscale_clone_keeps_state(rdl, oracle, torch_cuda):
    res, psfs, weights, freqs = multiscale_case(128, 1, seed=9)
    st = dict(max_iterations=300, clean_border_ratio=0.0, major_loop_gain=0.8)
    o, oi, g, gi, algs = run_both(rdl, oracle, res, psfs, weights, st, automask=(True, False),
                                  ms_kwargs=dict(max_scales=3))
    galg = algs[1]
    clone = galg.clone()
    assert clone.iteration_number == galg.iteration_number
    assert clone.scale_infos() == galg.scale_infos()
    for i in range(clone.scale_count):
        assert np.array_equal(clone.scale_mask(i, 128, 128), galg.scale_mask(i, 128, 128))


@pytest.mark.parametrize("n,c,iters,fit", [(192, 1, 600, False), (160, 3, 400, True)])
def test_multiscale_without_sub_minor_loop(rdl, oracle, torch_cuda, n, c, iters, fit):
    """settings.fast_sub_minor_loop = false (multiscale_algorithm.cc:462-520): one component per
    iteration, subtracted from the data and from the scale-convolved images, model built from
    shape functions; joined channels with the polynomial fit in the second case."""
    res, psfs, weights, freqs = multiscale_case(n, c, seed=3)
    st = dict(threshold=0.0, max_iterations=iters, clean_border_ratio=0.0, major_loop_gain=0.98)
    o, oi, g, gi, algs = run_both(rdl, oracle, res, psfs, weights, st,
                                  fit=(2, freqs, weights) if fit else None,
                                  ms_kwargs=dict(max_scales=4, fast_sub_minor_loop=False))
    assert o.iteration_number > 20
    assert len({int(s) for s in o.trace["scale"]}) > 1
    # cleaned to 2 % of the peak: the 1e-5 image tolerance refers to the image that went in
    check(o, oi, g, gi, algs, image_scale=float(np.abs(res).max()))


def test_multiscale_without_sub_minor_loop_automask(rdl, oracle, torch_cuda):
    """The per-iteration loop marks the per-scale mask at every component (:692-697) and phase 2
    searches inside the masks."""
    res, psfs, weights, freqs = multiscale_case(160, 1, seed=1)
    slow = dict(max_scales=4, fast_sub_minor_loop=False)
    st1 = dict(threshold=0.0, max_iterations=250, clean_border_ratio=0.0, major_loop_gain=0.98)
    o, oi, g, gi, algs = run_both(rdl, oracle, res, psfs, weights, st1, automask=(True, False),
                                  ms_kwargs=slow)
    scale = float(np.abs(res).max())
    check(o, oi, g, gi, algs, image_scale=scale)
    oalg, galg = algs
    masks = [oalg.scale_mask(i, 160, 160) for i in range(len(oalg.scale_infos()))]
    assert sum(int(m.sum()) for m in masks) > 20
    for i, m in enumerate(masks):
        assert np.array_equal(galg.scale_mask(i, 160, 160), m), i
    minor_before = oalg.minor_iterations()
    st2 = dict(threshold=0.0, max_iterations=400, clean_border_ratio=0.0, major_loop_gain=0.9,
               iteration_number=o.iteration_number)
    o2, oi2, g2, gi2, _ = run_both(rdl, oracle, oi[0], psfs, weights, st2, automask=(False, True),
                                   algs=algs)
    assert o2.iteration_number > o.iteration_number
    check(o2, oi2, g2, gi2, algs, minor_before=minor_before, image_scale=scale)


@pytest.mark.parametrize("fast", [True, False])
def test_multiscale_component_list(rdl, oracle, torch_cuda, fast):
    """settings.save_source_list: MultiScaleAlgorithm keeps every component (position, scale,
    one value per image) over the major iterations; after ComponentList::MergeDuplicates the
    list must equal the reference's own (oracle/_ref): same positions in the same order, values
    within the FFT tolerance. Joined channels with the polynomial fit, two major iterations."""
    torch = torch_cuda
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built")
    n, c = 128, 3
    res, psfs, weights, freqs = multiscale_case(n, c, seed=3)
    kw = dict(max_iterations=400 if fast else 150, major_loop_gain=0.9, clean_border_ratio=0.0)
    ref = oracle.Reference("multiscale", n, n, c, weights=weights, frequencies=freqs, beam_size=6.0,
                           max_scales=4, track_components=True, fast_sub_minor_loop=fast,
                           fit_terms=2, **kw)
    alg = rdl.MultiScaleAlgorithm(6.0, 1.0, max_scales=4, track_components=True,
                                  fast_sub_minor_loop=fast)
    alg.apply(**kw)
    alg.set_spectral_fitter(rdl.FIT_POLYNOMIAL, 2, freqs, weights, 1)
    r_res, r_model = res.copy(), np.zeros_like(res)
    d_res, d_model = torch.from_numpy(res.copy()).cuda(), torch.zeros((c, n, n), device="cuda")
    d_psf = torch.from_numpy(psfs).cuda()
    for call in range(2):
        w = ref.execute(r_res, r_model, psfs)
        g = alg.execute_major_iteration(rdl.Cube(n, n, c, 1, weights), d_res, d_model, d_psf)
        assert g.iteration_number == w.iteration_number, call
    want, got = ref.component_list(c), alg.component_list(c)
    assert len(got) == len(want) and sum(len(x) for x, _, _ in want) > 20
    peak = float(np.abs(res).max())
    for scale, ((wx, wy, wv), (gx, gy, gv)) in enumerate(zip(want, got)):
        assert np.array_equal(gx, wx) and np.array_equal(gy, wy), scale
        assert np.abs(gv - wv).max(initial=0.0) <= 1e-5 * peak, scale
    # the model is what the components add up to (scale 0 components are single pixels)
    alg.clear_component_list()
    assert alg.component_list(c) == []


def _major_iterations(rdl, res, psf_sets, weights, constant, host, n_calls=3):
    """n_calls major iterations on one algorithm object; psf_sets[k] are call k's PSFs."""
    import torch
    c, h, w = res.shape
    alg = rdl.MultiScaleAlgorithm(6.0, 1.0, max_scales=4)
    alg.set_psfs_constant(constant)
    cube = rdl.Cube(w, h, c, 1, weights)
    if host:
        r, m = res.copy(), np.zeros_like(res)
    else:
        r, m = torch.from_numpy(res.copy()).cuda(), torch.zeros((c, h, w), device="cuda")
    dev_psfs = {}
    out = []
    for k in range(n_calls):
        alg.apply(max_iterations=300 * (k + 1), clean_border_ratio=0.0, major_loop_gain=0.7,
                  trace_capacity=1 << 16)
        p = psf_sets[k]
        if not host:
            p = dev_psfs.setdefault(id(p), torch.from_numpy(p).cuda())
        g = alg.execute_major_iteration(cube, r, m, p)
        rr, mm = (r.copy(), m.copy()) if host else (r.cpu().numpy(), m.cpu().numpy())
        out.append((g.iteration_number, alg.trace().copy(), rr, mm, alg.scale_infos()))
    return out


@pytest.mark.parametrize("host", [False, True])
@pytest.mark.parametrize("c", [1, 3])
def test_multiscale_psf_cache_keeps_results(rdl, torch_cuda, host, c):
    # rdl_set_psfs_constant: what a call derives from the PSFs (convolved PSFs, spectra, twice
    # convolved PSFs, PSF peaks) is kept from one major iteration to the next; results must be the
    # bits of a handle that recomputes everything, call after call
    res, psfs, weights, _ = multiscale_case(192, c, seed=11)
    plain = _major_iterations(rdl, res, [psfs] * 3, weights, False, host)
    kept = _major_iterations(rdl, res, [psfs] * 3, weights, True, host)
    assert plain[-1][0] > plain[0][0] > 20
    for a, b in zip(plain, kept):
        assert a[0] == b[0]
        assert np.array_equal(a[1], b[1])
        assert np.array_equal(a[2], b[2]) and np.array_equal(a[3], b[3])
        assert a[4] == b[4]


@pytest.mark.parametrize("host", [False, True])
def test_multiscale_psf_cache_dropped_when_psfs_change(rdl, torch_cuda, host):
    # other planes in the second call (other pointers): the host path uploads them into the SAME
    # staging addresses, so the kept PSF products have to go
    res, psfs, weights, _ = multiscale_case(160, 2, seed=12)
    other = (psfs * np.float32(0.5) + np.roll(psfs, 3, axis=2) * np.float32(0.5)).astype(np.float32)
    sets = [psfs, other, other]
    plain = _major_iterations(rdl, res, sets, weights, False, host)
    kept = _major_iterations(rdl, res, sets, weights, True, host)
    for a, b in zip(plain, kept):
        assert a[0] == b[0]
        assert np.array_equal(a[1], b[1])
        assert np.array_equal(a[2], b[2]) and np.array_equal(a[3], b[3])
