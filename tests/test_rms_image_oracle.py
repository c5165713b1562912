"""cpp/math/rms_image.cc of the reference (compiled into oracle/_ref) against a numpy statement of
its window bounds: what the device kernels of csrc/rms_image.cu are written to. CPU only."""
import numpy as np
import pytest

import oracle

needs_ref = pytest.mark.skipif(oracle.ref_lib() is None, reason="oracle/_ref not built")


def sliding_minimum(a, window):
    h, w = a.shape
    w2 = window // 2
    t = np.empty_like(a)
    for x in range(w):
        t[:, x] = a[:, max(x, w2) - w2:min(x, w - w2) + w2].min(axis=1)
    o = np.empty_like(a)
    for y in range(h):
        o[y] = t[max(y, w2) - w2:min(y, h - w2) + w2].min(axis=0)
    return o


@needs_ref
@pytest.mark.parametrize("h,w,window", [(37, 53, 7), (40, 40, 2), (64, 96, 33), (30, 50, 60)])
def test_sliding_minimum_bounds(h, w, window):
    rng = np.random.default_rng(h * w)
    a = rng.standard_normal((h, w)).astype(np.float32)
    assert np.array_equal(oracle.ref_rms_sliding_minimum(a, window), sliding_minimum(a, window))
    assert np.array_equal(oracle.ref_rms_sliding_minimum(a, window, True),
                          -sliding_minimum(-a, window))


@needs_ref
def test_make_rms_factor_image():
    rng = np.random.default_rng(2)
    rms = rng.uniform(0.5, 2.0, (20, 30)).astype(np.float32)
    rms[3, 4] = 0.0
    f, s = oracle.ref_rms_make_factor_image(rms + 0, 1.0)
    nz = rms != 0
    assert s == float(rms.min()) == 0.0  # the zero pixel is the minimum: factors collapse to 0
    assert np.all(f[nz] == 0.0)
    rms[3, 4] = 0.7
    f, s = oracle.ref_rms_make_factor_image(rms, 1.0)
    assert s == float(rms.min())
    assert np.array_equal(f, (np.float64(s) / rms.astype(np.float64)).astype(np.float32))
    f0, _ = oracle.ref_rms_make_factor_image(rms, 0.0)
    assert np.all(f0 == 1.0)
