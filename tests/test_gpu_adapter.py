"""The drop-in, compiled: include/cuda_deconvolution.h (`radler::algorithms::CudaDeconvolution :
DeconvolutionAlgorithm`, the adapter of INTEGRATION.md) built against the reference's UNMODIFIED
headers and handed to the reference's OWN ParallelDeconvolution::SetAlgorithm
(parallel_deconvolution.cc:227-242) in oracle/_ref/libradler_ref_adapter.so. The reference code then
drives the CUDA library exactly as it drives its CPU algorithms: settings through the base class,
Clone() per facet, the non-owning clean-mask pointer, the per-facet SetMaxIterations(0) probe and
SetMajorIterationThreshold, host `aocommon::Image` buffers. Each case is run twice -- CPU algorithm
and CUDA adapter behind the same reference driver -- and compared."""
import numpy as np
import pytest

from helpers import hogbom_case, multiscale_case, plus_psf_fixture

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def adapter(oracle):
    if oracle.ref_lib() is None or oracle.adapter_lib() is None:
        pytest.skip("oracle/_ref (reference sources + adapter) is not built")
    return oracle


def run_pair(orc, algorithm, res, psfs, calls=1, mask=None, rms=None, **kw):
    n_img, h, w = res.shape
    n_pol = len(kw.get("pols", ("I",)))
    out = []
    for use_adapter in (False, True):
        r = orc.Reference(algorithm, w, h, n_img // n_pol, adapter=use_adapter, **kw)
        if mask is not None:
            r.set_clean_mask(mask)
        if rms is not None:
            r.set_rms_factor_image(rms)
        a, m = res.copy(), np.zeros_like(res)
        results = [r.execute(a, m, psfs) for _ in range(calls)]
        out.append((results, a, m, r.iteration_number))
    return out


def compare(pair, exact, peak):
    (want, w_res, w_model, w_it), (got, g_res, g_model, g_it) = pair
    assert g_it == w_it
    for w, g in zip(want, got):
        assert g.another_iteration_required == w.another_iteration_required
        assert (g.starting_peak_value is None) == (w.starting_peak_value is None)
        if exact:
            assert g.final_peak_value == w.final_peak_value
        else:
            assert g.final_peak_value == pytest.approx(w.final_peak_value, rel=2e-4, abs=1e-6 * peak)
    tol = 0.0 if exact else 1e-5 * peak
    assert np.abs(g_res - w_res).max() <= tol
    assert np.abs(g_model - w_model).max() <= tol


@pytest.mark.parametrize("algorithm,kw", [("generic", dict(use_sub_minor_optimization=True)),
                                          ("generic", dict(use_sub_minor_optimization=False)),
                                          ("multiscale", dict())])
@pytest.mark.parametrize("shift", [(0, 0), (7, -11), (-9, 23)])
def test_reference_known_answer_through_the_adapter(adapter, torch_cuda, algorithm, kw, shift):
    # cpp/test/test_radler.cc:98-172 with CudaDeconvolution as the algorithm
    psf, res, src = plus_psf_fixture(*shift)
    r = adapter.Reference(algorithm, 64, 64, threshold=1e-8, max_iterations=1000, adapter=True, **kw)
    model = np.zeros_like(res)
    r.execute(res, model, psf)
    assert np.abs(res).max() < 2e-6
    assert model.flat[src] == pytest.approx(2.5, rel=1e-6)
    model.flat[src] = 0
    assert np.abs(model).max() < 2e-6


@pytest.mark.parametrize("grid", [(1, 1), (3, 3)])
def test_generic_clean_facets_through_the_reference_parallel_deconvolution(adapter, torch_cuda, grid):
    n, c = 192, 2
    res, psfs, weights, freqs = hogbom_case(n, c)
    rng = np.random.default_rng(3)
    mask = (rng.uniform(size=(n, n)) < 0.9).astype(np.uint8)
    pair = run_pair(adapter, "generic", res, psfs, calls=2, mask=mask, weights=weights,
                    frequencies=freqs, use_sub_minor_optimization=False, grid=grid, threads=1,
                    max_iterations=200, major_loop_gain=0.7, fit_terms=2)
    # (the iteration counter read back is the first facet's; that facet may have nothing to clean)
    assert np.count_nonzero(pair[0][2]) > 4
    compare(pair, exact=True, peak=np.abs(res).max())


def test_generic_clean_polarisations_rms_and_path_a(adapter, torch_cuda):
    n, c, pols = 128, 2, ("XX", "YY")
    base, psfs, weights, freqs = hogbom_case(n, c)
    rng = np.random.default_rng(5)
    res = np.zeros((c * 2, n, n), np.float32)
    for ch in range(c):
        res[2 * ch] = base[ch]
        res[2 * ch + 1] = base[ch] * np.float32(0.8) + rng.normal(0, 1e-3, (n, n)).astype(np.float32)
    rms = rng.uniform(0.7, 1.3, (n, n)).astype(np.float32)
    for sub_minor in (False, True):
        pair = run_pair(adapter, "generic", res, psfs, rms=rms, pols=pols, weights=weights,
                        frequencies=freqs, use_sub_minor_optimization=sub_minor, max_iterations=300,
                        major_loop_gain=0.9, fit_terms=2)
        assert pair[0][3] > 20
        compare(pair, exact=not sub_minor, peak=np.abs(res).max())


def test_multiscale_and_iuwt_through_the_adapter(adapter, torch_cuda):
    n = 256
    res, psfs, weights, freqs = multiscale_case(n, 2)
    pair = run_pair(adapter, "multiscale", res, psfs, weights=weights, frequencies=freqs,
                    beam_size=6.0, max_scales=4, max_iterations=600, major_loop_gain=0.8, fit_terms=2)
    assert pair[0][3] > 30
    compare(pair, exact=False, peak=np.abs(res).max())
    res, psfs, _, _ = multiscale_case(n, 1, seed=3)
    pair = run_pair(adapter, "iuwt", res, psfs, max_iterations=10, major_loop_gain=0.8)
    assert pair[0][3] > 3
    compare(pair, exact=False, peak=np.abs(res).max())
