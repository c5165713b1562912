"""Pins the restated oracle (oracle/*.cc) against the reference's OWN algorithm sources.

oracle/_ref/libradler_ref.so compiles cpp/algorithms/{generic_clean, subminor_loop,
multiscale_algorithm, multiscale/multiscale_transforms, threaded_deconvolution_tools,
deconvolution_algorithm, parallel_deconvolution, iuwt*}.cc, cpp/image_set.cc, ... where they
lie under /root/reference (oracle/Makefile target `ref`; only the un-vendored aocommon /
schaapcommon leaves are stand-ins, oracle/ref_shim/). Both sides share those leaves, so the
restatement has to agree with the reference's control flow and arithmetic BIT FOR BIT:
iteration counts, result flags, residual and model images are compared with array_equal.
Also re-runs the reference's own known-answer fixture (cpp/test/test_radler.cc:98-172)
through the reference sources themselves."""
import numpy as np
import pytest

from helpers import hogbom_case, multiscale_case, plus_psf_fixture


@pytest.fixture(scope="module")
def ref(oracle):
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref is not built (needs /root/reference or a prebuilt .so)")
    return oracle


def same_images(a, b):
    return np.array_equal(a.view(np.uint32), b.view(np.uint32))


def run_generic(orc, res, psfs, weights, sub_minor, fit=None, freqs=None, mask=None, rms=None,
                **kw):
    c, h, w = res.shape
    r = orc.Reference("generic", w, h, c, weights=weights, frequencies=freqs,
                      use_sub_minor_optimization=sub_minor, fit_terms=fit or 0, **kw)
    if mask is not None:
        r.set_clean_mask(mask)
    if rms is not None:
        r.set_rms_factor_image(rms)
    a, m = res.copy(), np.zeros_like(res)
    out = r.execute(a, m, psfs)
    st = orc.Settings(threshold=kw.get("threshold", 0.0),
                      minor_loop_gain=kw.get("minor_loop_gain", 0.1),
                      major_loop_gain=kw.get("major_loop_gain", 1.0),
                      clean_border_ratio=kw.get("clean_border_ratio", 0.0),
                      max_iterations=kw.get("max_iterations", 500),
                      allow_negative=kw.get("allow_negative", True),
                      stop_on_negative=kw.get("stop_on_negative", False), clean_mask=mask,
                      rms_factor=rms)
    if fit:
        st.spectral_mode, st.n_terms, st.frequencies, st.fit_weights = 1, fit, freqs, weights
    a2, m2 = res.copy(), np.zeros_like(res)
    o = orc.GenericClean(sub_minor).execute(orc.Cube(w, h, c, 1, weights), st, a2, m2, psfs)
    assert out.iteration_number == o.iteration_number
    assert out.final_peak_value == o.final_peak_value
    assert out.starting_peak_value == o.starting_peak_value
    assert out.another_iteration_required == o.another_iteration_required
    assert same_images(a, a2) and same_images(m, m2)
    return out


@pytest.mark.parametrize("sub_minor", [False, True])
def test_generic_clean_restatement_is_bit_exact(ref, sub_minor):
    res, psfs, weights, freqs = hogbom_case(128, 3)
    rng = np.random.default_rng(5)
    mask = (rng.uniform(size=(128, 128)) < 0.7).astype(np.uint8)
    rms = rng.uniform(0.5, 1.5, (128, 128)).astype(np.float32)
    n = 0
    for kw in [dict(max_iterations=300, major_loop_gain=0.9),
               dict(max_iterations=2000, major_loop_gain=0.8, threshold=0.05,
                    clean_border_ratio=0.05),
               dict(max_iterations=200, allow_negative=False),
               dict(max_iterations=200, stop_on_negative=True),
               dict(max_iterations=250, mask=mask, major_loop_gain=0.85),
               dict(max_iterations=250, rms=rms, major_loop_gain=0.85),
               dict(max_iterations=250, fit=2, freqs=freqs, major_loop_gain=0.85),
               dict(max_iterations=0)]:
        n += run_generic(ref, res, psfs, weights, sub_minor, **kw).iteration_number
    assert n > 300
    single = hogbom_case(96, 1, seed=3)
    run_generic(ref, single[0], single[1], None, sub_minor, max_iterations=400,
                major_loop_gain=0.9)


def run_multiscale(orc, res, psfs, weights, freqs, fit=0, ms=None, phases=None, **kw):
    c, h, w = res.shape
    ms = ms or {}
    r = orc.Reference("multiscale", w, h, c, weights=weights, frequencies=freqs, beam_size=6.0,
                      fit_terms=fit, **ms, **kw)
    o = orc.MultiScale(beam_size=6.0, **ms)
    a, m = res.copy(), np.zeros_like(res)
    a2, m2 = res.copy(), np.zeros_like(res)
    outs = []
    for phase in phases or [dict()]:
        st = orc.Settings(threshold=phase.get("threshold", kw.get("threshold", 0.0)),
                          minor_loop_gain=kw.get("minor_loop_gain", 0.1),
                          major_loop_gain=kw.get("major_loop_gain", 1.0),
                          clean_border_ratio=kw.get("clean_border_ratio", 0.0),
                          max_iterations=kw.get("max_iterations", 500),
                          allow_negative=kw.get("allow_negative", True),
                          iteration_number=r.iteration_number)
        if fit:
            st.spectral_mode, st.n_terms, st.frequencies, st.fit_weights = 1, fit, freqs, weights
        if "automask" in phase:
            r.set_auto_mask_mode(*phase["automask"])
            o.set_automask_mode(*phase["automask"])
        if "threshold" in phase:
            r.set_threshold(phase["threshold"])
        out = r.execute(a, m, psfs)
        oo = o.execute(orc.Cube(w, h, c, 1, weights), st, a2, m2, psfs)
        assert out.iteration_number == oo.iteration_number
        assert out.final_peak_value == oo.final_peak_value
        assert out.another_iteration_required == oo.another_iteration_required
        assert same_images(a, a2) and same_images(m, m2)
        outs.append(out)
    assert r.scale_sizes() == [s["scale"] for s in o.scale_infos()]
    return r, o, outs


def test_multiscale_restatement_is_bit_exact(ref):
    res, psfs, weights, freqs = multiscale_case(128, 3)
    kw = dict(max_iterations=800, major_loop_gain=0.8)
    _, _, outs = run_multiscale(ref, res, psfs, weights, freqs, ms=dict(max_scales=4), **kw)
    assert outs[0].iteration_number > 30
    run_multiscale(ref, res, psfs, weights, freqs, fit=2, ms=dict(max_scales=4), **kw)
    run_multiscale(ref, res, psfs, weights, freqs, ms=dict(shape=1, max_scales=3), **kw)
    run_multiscale(ref, res, psfs, weights, freqs, ms=dict(scale_list=[0.0, 10.0, 30.0]), **kw)
    run_multiscale(ref, res, psfs, weights, freqs, max_iterations=300, allow_negative=False,
                   clean_border_ratio=0.05, threshold=0.02, major_loop_gain=0.9)
    single = multiscale_case(256, 1, seed=2)
    run_multiscale(ref, single[0], single[1], None, single[3], ms=dict(max_scales=5),
                   max_iterations=1500, major_loop_gain=0.85)


def test_multiscale_without_sub_minor_loop_bit_exact(ref):
    # settings.fast_sub_minor_loop = false: multiscale_algorithm.cc:462-520
    res, psfs, weights, freqs = multiscale_case(128, 3, seed=3)
    kw = dict(max_iterations=300, major_loop_gain=0.97)
    slow = dict(max_scales=4, fast_sub_minor_loop=False)
    _, _, outs = run_multiscale(ref, res, psfs, weights, freqs, ms=slow, **kw)
    assert outs[0].iteration_number > 30
    run_multiscale(ref, res, psfs, weights, freqs, fit=2, ms=slow, **kw)
    single = multiscale_case(160, 1, seed=1)
    run_multiscale(ref, single[0], single[1], None, single[3], ms=slow, max_iterations=400,
                   major_loop_gain=0.98,
                   phases=[dict(automask=(True, False)), dict(automask=(False, True))])


def test_multiscale_automask_phases_bit_exact(ref):
    # radler.cc:244-268: phase 1 tracks per-scale masks, phase 2 cleans inside them
    res, psfs, weights, freqs = multiscale_case(128, 1, seed=7)
    sigma = float(1.4826 * np.median(np.abs(res - np.median(res))))
    r, o, outs = run_multiscale(
        ref, res, psfs, None, freqs, ms=dict(max_scales=3), max_iterations=4000,
        major_loop_gain=0.8,
        phases=[dict(automask=(True, False), threshold=4 * sigma),
                dict(automask=(False, True), threshold=1 * sigma)])
    assert outs[1].iteration_number > outs[0].iteration_number > 0
    for i in range(len(r.scale_sizes())):
        assert np.array_equal(r.scale_mask(i), o.scale_mask(i, 128, 128)), i


@pytest.mark.parametrize("algorithm,kw", [("generic", dict(use_sub_minor_optimization=True)),
                                          ("generic", dict(use_sub_minor_optimization=False)),
                                          ("multiscale", dict())])
@pytest.mark.parametrize("shift", [(0, 0), (7, -11), (-9, 15)])
def test_reference_known_answer_through_reference_sources(ref, algorithm, kw, shift):
    # cpp/test/test_radler.cc:98-172 with the reference's own algorithm code: validates the
    # stand-in leaves (Convolve, PrepareConvolutionKernel, Trim/Untrim) that both sides share
    psf, res, src = plus_psf_fixture(*shift)
    r = ref.Reference(algorithm, 64, 64, threshold=1e-8, max_iterations=1000, **kw)
    model = np.zeros_like(res)
    r.execute(res, model, psf)
    assert np.abs(res).max() < 2e-6
    assert model.flat[src] == pytest.approx(2.5, rel=1e-6)
    model.flat[src] = 0
    assert np.abs(model).max() < 2e-6
