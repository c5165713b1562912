"""GPU parity of IuwtDeconvolution::ExecuteMajorIteration (csrc/iuwt_deconvolution.cu) against the
reference's OWN iuwt_deconvolution_algorithm.cc + iuwt/*.cc compiled in oracle/_ref (the reference
excludes IUWT from its own tests, cpp/test/test_radler.cc:101, so its sources are the only
checker there is).

Bit-exact: the input-aliased decomposition `Decompose(x, x, false)` that the conjugate-gradient loop
really computes. By tolerance: the major iteration -- its step sizes come from float32 sequential
dot products in the reference and from double-precision tree sums here, and every CG step goes
through FFT convolutions; the iteration count, the scale bookkeeping (hence the number of
successful / failed structure fits) and the result flag must agree exactly, images within 1e-5 of
the image peak."""
import numpy as np
import pytest

from helpers import multiscale_case

pytestmark = pytest.mark.gpu
REL_TOL = 1e-5


@pytest.mark.parametrize("h,w,scales", [(48, 64, 2), (256, 256, 4), (130, 250, 3), (512, 384, 5)])
def test_aliased_decomposition_is_bit_exact(rdl, oracle, torch_cuda, h, w, scales):
    torch = torch_cuda
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(h + w)
    x = rng.standard_normal((h, w)).astype(np.float32)
    want = oracle.ref_iuwt_decompose_aliased(x, scales)
    alg = rdl.DeconvolutionAlgorithm(rdl.IUWT)
    d_x = torch.from_numpy(x.copy()).cuda()
    coef = torch.zeros((scales, h, w), device="cuda")
    alg.iuwt_decompose_aliased(d_x, coef, scales)
    torch.cuda.synchronize()
    got = coef.cpu().numpy()
    for s in range(scales):
        assert np.array_equal(got[s], want[s]), s
    # and it is NOT the decomposition of x: scale 0 differs grossly from the un-aliased call
    assert np.abs(want[0] - oracle.ref_iuwt_decompose(x, scales, False)[0]).max() > 0.1


def run_both(rdl, oracle, torch, res, psfs, calls=1, mask=None, **kw):
    c, n, _ = res.shape
    ref = oracle.Reference("iuwt", n, n, c, **kw)
    alg = rdl.DeconvolutionAlgorithm(rdl.IUWT)
    alg.apply(max_iterations=kw.get("max_iterations", 500),
              minor_loop_gain=kw.get("minor_loop_gain", 0.1),
              major_loop_gain=kw.get("major_loop_gain", 1.0),
              clean_border_ratio=kw.get("clean_border_ratio", 0.0),
              threshold=kw.get("threshold", 0.0),
              allow_negative_components=kw.get("allow_negative", True))
    if mask is not None:
        ref.set_clean_mask(mask)
        alg.set_clean_mask(mask)
    r_res, r_model = res.copy(), np.zeros_like(res)
    d_res, d_model = torch.from_numpy(res.copy()).cuda(), torch.zeros((c, n, n), device="cuda")
    d_psf = torch.from_numpy(psfs).cuda()
    cube = rdl.Cube(n, n, c, 1, None)
    out = []
    for _ in range(calls):
        w = ref.execute(r_res, r_model, psfs)
        g = alg.execute_major_iteration(cube, d_res, d_model, d_psf)
        out.append((w, g))
    return out, (r_res, r_model), (d_res.cpu().numpy(), d_model.cpu().numpy())


def check(out, ref_images, gpu_images, res):
    for call, (w, g) in enumerate(out):
        assert g.iteration_number == w.iteration_number, call
        assert g.another_iteration_required == w.another_iteration_required, call
        assert g.final_peak_value == pytest.approx(w.final_peak_value, rel=1e-4, abs=1e-6), call
        assert g.starting_peak_value is None
    peak = np.abs(res).max()
    assert np.abs(gpu_images[0] - ref_images[0]).max() <= REL_TOL * peak
    assert np.abs(gpu_images[1] - ref_images[1]).max() <= REL_TOL * peak


@pytest.mark.parametrize("n,iters,mgain", [(256, 12, 0.8), (384, 10, 0.9), (512, 10, 0.8)])
def test_iuwt_major_iteration_matches_reference_sources(rdl, oracle, torch_cuda, n, iters, mgain):
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built")
    res, psfs, _, _ = multiscale_case(n, 1, seed=3)
    out, ri, gi = run_both(rdl, oracle, torch_cuda, res, psfs, max_iterations=iters,
                           major_loop_gain=mgain, minor_loop_gain=0.1)
    assert out[0][0].iteration_number > 3
    assert np.abs(ri[1]).max() > 0  # something was deconvolved
    check(out, ri, gi, res)


def test_iuwt_two_calls_mask_and_positive_only(rdl, oracle, torch_cuda):
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built")
    n = 256
    res, psfs, _, _ = multiscale_case(n, 1, seed=5)
    mask = np.zeros((n, n), np.uint8)
    mask[16:240, 24:232] = 1
    out, ri, gi = run_both(rdl, oracle, torch_cuda, res, psfs, calls=2, mask=mask, max_iterations=16,
                           major_loop_gain=0.5, minor_loop_gain=0.2, clean_border_ratio=0.02,
                           allow_negative=False)
    assert out[1][0].iteration_number > out[0][0].iteration_number or out[0][0].iteration_number == 16
    check(out, ri, gi, res)


@pytest.mark.parametrize("n,c,iters", [(256, 2, 8), (256, 3, 6), (384, 2, 6)])
def test_iuwt_joined_images_refit_per_image(rdl, oracle, torch_cuda, n, c, iters):
    # PerformSubImageFitAll (iuwt_deconvolution_algorithm.cc:628-801): the structure found in the
    # integrated image is split into connected components and refitted image by image. The
    # reference sums model and masked dirty over a component in float32, in flood-fill order; the
    # device in double, fixed order: the per-image factors agree to ~1e-6, the images to 1e-4.
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built")
    res, psfs, _, _ = multiscale_case(n, c, seed=8)
    out, ri, gi = run_both(rdl, oracle, torch_cuda, res, psfs, max_iterations=iters,
                           major_loop_gain=0.8, minor_loop_gain=0.1)
    assert out[0][0].iteration_number > 2
    assert all(np.abs(ri[1][i]).max() > 0 for i in range(c))  # every image got its share
    # the images are NOT copies of one another: the refit is per image
    assert np.abs(ri[1][0] - ri[1][c - 1]).max() > 1e-4 * np.abs(ri[1]).max()
    for call, (w, g) in enumerate(out):
        assert g.iteration_number == w.iteration_number, call
        assert g.another_iteration_required == w.another_iteration_required, call
        assert g.final_peak_value == pytest.approx(w.final_peak_value, rel=1e-3, abs=1e-6), call
    peak = np.abs(res).max()
    assert np.abs(gi[0] - ri[0]).max() <= 1e-4 * peak
    assert np.abs(gi[1] - ri[1]).max() <= 1e-4 * peak
