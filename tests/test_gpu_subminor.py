"""GPU parity of GenericClean path A (SubMinorLoop + CorrectResidualDirty,
generic_clean.cc:115-162, subminor_loop.cc) through the C ABI vs the CPU oracle.
Component positions, iteration counts and model images bit-exact (the loop works on gathered
raw pixels); the FFT-corrected residual within 1e-5 of the image peak."""
import numpy as np
import pytest

from helpers import hogbom_case, plus_psf_fixture
from test_gpu_hogbom import _run_both

pytestmark = pytest.mark.gpu
REL_TOL = 1e-5


def _check(o, oi, g, gi, alg, exact_model=True):
    assert g.iteration_number == o.iteration_number
    tr = alg.trace()
    assert len(tr) == len(o.trace)
    for k in ("iteration", "x", "y"):
        assert np.array_equal(tr[k], o.trace[k]), k
    assert np.array_equal(tr["value"].view(np.uint32), o.trace["value"].view(np.uint32))
    assert alg.last_selected_count == o.n_selected
    assert (g.starting_peak_value is None) == (o.starting_peak_value is None)
    if o.starting_peak_value is not None:
        assert g.starting_peak_value == o.starting_peak_value
    assert g.final_peak_value == o.final_peak_value
    assert g.another_iteration_required == o.another_iteration_required
    assert g.is_diverging == o.is_diverging
    if exact_model:
        assert np.array_equal(gi[1], oi[1])
    scale = max(np.abs(oi[0]).max(), np.abs(oi[1]).max(), 1e-30)
    assert np.abs(gi[0].astype(np.float64) - oi[0]).max() / scale <= REL_TOL


@pytest.mark.parametrize("shift", [(0, 0), (7, -11), (-9, 15), (-9, 23)])
def test_path_a_point_source_fixture(rdl, oracle, torch_cuda, shift):
    psf, res, src = plus_psf_fixture(*shift)
    st = dict(threshold=1e-8, max_iterations=1000, clean_border_ratio=0.0)
    o, oi, g, gi, alg = _run_both(rdl, oracle, res, psf, None, st, sub_minor=True)
    _check(o, oi, g, gi, alg)
    assert np.abs(gi[0]).max() < 2e-6
    assert gi[1].flat[src] == pytest.approx(2.5, rel=1e-6)


@pytest.mark.parametrize("n,c,iters,mgain,device_resident",
                         [(256, 1, 500, 0.8, False), (250, 2, 300, 0.9, True),
                          (512, 4, 2000, 0.95, True), (1024, 1, 5000, 0.9, True)])
def test_path_a_matches_oracle(rdl, oracle, torch_cuda, n, c, iters, mgain, device_resident):
    res, psfs, weights, freqs = hogbom_case(n, c)
    st = dict(threshold=0.0, max_iterations=iters, clean_border_ratio=0.0, minor_loop_gain=0.1,
              major_loop_gain=mgain)
    o, oi, g, gi, alg = _run_both(rdl, oracle, res, psfs, weights, st, sub_minor=True,
                                  device_resident=device_resident)
    assert o.iteration_number > 10
    _check(o, oi, g, gi, alg)


def test_path_a_variants(rdl, oracle, torch_cuda):
    res, psfs, weights, freqs = hogbom_case(256, 3)
    rng = np.random.default_rng(2)
    mask = (rng.uniform(size=(256, 256)) < 0.5).astype(np.uint8)
    for st in [dict(threshold=0.05, max_iterations=5000, clean_border_ratio=0.05, major_loop_gain=0.9),
               dict(max_iterations=300, major_loop_gain=0.8, clean_border_ratio=0.1),
               dict(max_iterations=300, allow_negative=False, clean_border_ratio=0.0, major_loop_gain=0.9),
               dict(max_iterations=300, stop_on_negative=True, clean_border_ratio=0.0, major_loop_gain=0.9),
               dict(max_iterations=200, clean_mask=mask, clean_border_ratio=0.0, major_loop_gain=0.9),
               dict(max_iterations=0, clean_border_ratio=0.0),
               dict(threshold=1e9, max_iterations=100, clean_border_ratio=0.0)]:  # nothing selected
        o, oi, g, gi, alg = _run_both(rdl, oracle, res, psfs, weights, st, sub_minor=True)
        _check(o, oi, g, gi, alg)
    st = dict(max_iterations=300, clean_border_ratio=0.0, major_loop_gain=0.9)
    o, oi, g, gi, alg = _run_both(rdl, oracle, res, psfs, weights, st, sub_minor=True,
                                  fit=(2, freqs, weights))
    _check(o, oi, g, gi, alg)


def test_path_a_divergence(rdl, oracle, torch_cuda):
    res, psfs, weights, _ = hogbom_case(128, 1)
    psfs = psfs * np.float32(-1.5)
    st = dict(max_iterations=1000, clean_border_ratio=0.0, minor_loop_gain=0.5, major_loop_gain=0.5)
    o, oi, g, gi, alg = _run_both(rdl, oracle, res, psfs, None, st, sub_minor=True)
    assert o.is_diverging
    _check(o, oi, g, gi, alg)
