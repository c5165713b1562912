"""Parity at the sizes BASELINE.json quotes, bounded so the CPU side of
each test stays within a minute on the GPU box's host cores:

* K2: MultiScaleAlgorithm on 4096^2 x 1, 5 scales, two-phase auto-masking, against the restated
  oracle (bit-identical to the reference's own sources, tests/test_reference_sources.py), peak
  trace included; the number of verified FFT near-ties the check tolerated is recorded;
* K3: the IUWT kernels at 2048^2 x 6 scales against the reference's own iuwt_decomposition.cc
  and the restated image_analysis / Mad;
* K5: an 8 x 8 facet grid (reduced frame: 1024^2, 128^2 facets) against the reference's own
  ParallelDeconvolution + GenericClean, both paths.
"""
import json
import os

import numpy as np
import pytest

from helpers import add_blobs, dirty_image, gaussian_psf, hogbom_case, point_source_sky
from test_gpu_multiscale import check as ms_check
from test_gpu_multiscale import run_both as ms_run_both

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _record(name, payload):
    out = os.path.join(ROOT, "gpurun_out")
    if os.path.isdir(out):
        with open(os.path.join(out, name), "w") as f:
            json.dump(payload, f)


def test_k2_multiscale_4096_two_phase_automask(rdl, oracle, torch_cuda):
    n = 4096
    # (helpers.multiscale_case scales its blob count with the area: minutes of set-up at this size)
    sky = add_blobs(point_source_sky(n, 300, seed=21, margin=n // 8), 10, seed=23)
    psfs = gaussian_psf(n, 6.0)[None]
    res = dirty_image(sky, psfs[0], 1e-3, seed=27)[None]
    weights = None
    sigma = 1.4826 * np.median(np.abs(res - np.median(res)))
    st1 = dict(threshold=float(5 * sigma), max_iterations=150, clean_border_ratio=0.0,
               major_loop_gain=0.8)
    o, oi, g, gi, algs = ms_run_both(rdl, oracle, res, psfs, weights, st1, automask=(True, False),
                                     ms_kwargs=dict(max_scales=5))
    assert o.iteration_number > 50
    ties1 = ms_check(o, oi, g, gi, algs)
    oalg, galg = algs
    # a component that a verified near-tie moved by one pixel marks another pixel of the mask
    differing = sum(int(np.count_nonzero(galg.scale_mask(i, n, n) != oalg.scale_mask(i, n, n)))
                    for i in range(len(oalg.scale_infos())))
    assert differing <= 2 * ties1
    minor_before = oalg.minor_iterations()
    st2 = dict(threshold=float(1 * sigma), max_iterations=250, clean_border_ratio=0.0,
               major_loop_gain=0.8, iteration_number=o.iteration_number)
    o2, oi2, g2, gi2, _ = ms_run_both(rdl, oracle, oi[0], psfs, weights, st2, automask=(False, True),
                                      algs=algs)
    ties2 = ms_check(o2, oi2, g2, gi2, algs, minor_before=minor_before)
    _record("near_ties_k2_4096.json",
            {"frame": n, "scales": len(oalg.scale_infos()),
             "phase1": {"components": int(o.iteration_number), "trace_entries": len(o.trace),
                        "near_ties": ties1, "mask_pixels_differing": differing},
             "phase2": {"components": int(o2.iteration_number - o.iteration_number),
                        "trace_entries": len(o2.trace), "near_ties": ties2}})


def test_k3_iuwt_kernels_2048(rdl, oracle, torch_cuda):
    torch = torch_cuda
    n, n_scales = 2048, 6
    rng = np.random.default_rng(31)
    img = rng.standard_normal((n, n)).astype(np.float32)
    y, x = np.mgrid[0:n, 0:n]
    for cx, cy, s, a in [(700, 900, 30.0, 40.0), (1500, 400, 6.0, 80.0), (300, 1700, 90.0, 15.0)]:
        img += (a * np.exp(-((x - cx) ** 2 + (y - cy) ** 2) / (2 * s * s))).astype(np.float32)
    alg = rdl.DeconvolutionAlgorithm(rdl.IUWT)
    d = torch.from_numpy(img).cuda()
    coef = torch.zeros((n_scales + 1, n, n), device="cuda")
    alg.iuwt_decompose(d, coef, n_scales, True)
    got = coef.cpu().numpy()
    # the reference's own iuwt_decomposition.cc where it is built, else its restatement
    want = (oracle.ref_iuwt_decompose(img, n_scales, True) if oracle.ref_lib() is not None
            else oracle.iuwt_decompose(img, n_scales, True))
    assert np.array_equal(got.view(np.uint32), np.asarray(want, np.float32).view(np.uint32))
    out = torch.zeros((n, n), device="cuda")
    alg.iuwt_recompose(coef, out, n_scales, True)
    assert np.array_equal(out.cpu().numpy(), oracle.iuwt_recompose(want, n_scales, True))
    # Mad and arg-max of every plane (the reference's own source where it is built)
    mad, max_abs = (oracle.ref_iuwt_analysis() if oracle.ref_lib() is not None
                    else oracle.iuwt_analysis())[3:5]
    for s in range(n_scales):
        plane = np.ascontiguousarray(want[s])
        assert alg.iuwt_mad(coef[s]) == mad(plane), s
        assert alg.iuwt_max_abs(coef[s], 0.0, True) == max_abs(plane, 0.0, True), s


@pytest.mark.parametrize("sub_minor", [False, True])
def test_k5_facets_8x8_reduced_frame(rdl, oracle, torch_cuda, sub_minor):
    torch = torch_cuda
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built")
    n, c, grid = 1024, 1, (8, 8)
    res, psfs, weights, _ = hogbom_case(n, c, seed=41)
    kw = dict(max_iterations=400, minor_loop_gain=0.1, major_loop_gain=0.7, clean_border_ratio=0.0)
    ref = oracle.Reference("generic", n, n, c, weights=weights, use_sub_minor_optimization=sub_minor,
                           grid=grid, threads=1, **kw)
    r_res, r_model = res.copy(), np.zeros_like(res)
    alg = rdl.GenericClean(use_sub_minor_optimization=sub_minor)
    alg.apply(**kw)
    d_res, d_model = torch.from_numpy(res.copy()).cuda(), torch.zeros((c, n, n), device="cuda")
    d_psf = torch.from_numpy(psfs).cuda()
    cube = rdl.Cube(n, n, c, 1, weights)
    for call in range(2):
        w = ref.execute(r_res, r_model, psfs)
        g = alg.parallel_execute_major_iteration(grid[0], grid[1], cube, d_res, d_model, d_psf)
        assert g.another_iteration_required == w.another_iteration_required, call
        if sub_minor:
            assert g.final_peak_value == pytest.approx(w.final_peak_value, rel=1e-4), call
        else:
            assert np.float32(g.final_peak_value) == np.float32(w.final_peak_value), call
    assert len(alg.facets()) == 64
    assert np.count_nonzero(r_model) > 20
    peak = np.abs(res).max()
    g_res, g_model = d_res.cpu().numpy(), d_model.cpu().numpy()
    if not sub_minor:
        assert np.array_equal(g_res, r_res) and np.array_equal(g_model, r_model)
        return
    # Path A: CorrectResidualDirty goes through FFTs, so after the first call the two residuals
    # differ by FFT round-off (3e-7 of the peak). The dividers of the next call are minimum-cost
    # paths through THAT image: where two routes tie to within the round-off, a divider passes one
    # pixel further left or right, and that pixel belongs to the neighbouring facet, whose version
    # of it (its own components' sidelobes subtracted) differs. Seen here: ONE pixel of 2^20,
    # 1.7e-3 of the peak, same models; each facet alone agrees three ways (oracle, reference
    # sources, device) within 2e-7 (tests/tools/k5_debug2.py). Allow a few such pixels, everything else
    # within 1e-5.
    diff = np.abs(g_res - r_res)
    flipped = int(np.count_nonzero(diff > 1e-5 * peak))
    moved = int(np.count_nonzero(np.abs(g_model - r_model) > 1e-5 * peak))
    _record("k5_path_a_8x8.json", {"residual_pixels_beyond_1e-5": flipped,
                                   "model_pixels_differing": moved,
                                   "residual_max_rel": float(diff.max() / peak),
                                   "components": int(np.count_nonzero(r_model))})
    assert moved == 0
    assert flipped <= 8
    assert diff.max() <= 5e-3 * peak
