"""GPU parity of the IUWT decomposition kernels (csrc/iuwt.cu) through the C ABI against the
CPU oracle restatement (oracle/iuwt.cc, itself pinned to the reference's own
iuwt_decomposition.{h,cc}): BIT-EXACT coefficient planes, recomposed images and masked
recompositions; plus the committed golden vectors generated from the reference sources, and
size-independent properties at the K3 size (2048^2, 6 scales)."""
import os

import numpy as np
import pytest

from test_iuwt_oracle import GOLDEN, image

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    import torch
    import oracle
    import ska_sdp_func_radler_b200 as rdl
    return torch, oracle, rdl, rdl.DeconvolutionAlgorithm(rdl.IUWT)


def gpu_decompose(env, img, n_scales, include_largest=True):
    torch, _, _, alg = env
    d = torch.from_numpy(img).cuda()
    coef = torch.zeros((n_scales + (1 if include_largest else 0),) + img.shape, device="cuda")
    alg.iuwt_decompose(d, coef, n_scales, include_largest)
    torch.cuda.synchronize()
    return coef


@pytest.mark.parametrize("h,w,n_scales", [(64, 64, 2), (96, 80, 3), (70, 131, 4), (256, 256, 5),
                                          (30, 29, 3), (300, 517, 6)])
@pytest.mark.parametrize("include_largest", [True, False])
def test_decompose_bit_exact(env, h, w, n_scales, include_largest):
    _, oracle, _, _ = env
    img = image(h, w, h * 1000 + w)
    expect = oracle.iuwt_decompose(img, n_scales, include_largest)
    got = gpu_decompose(env, img, n_scales, include_largest).cpu().numpy()
    assert np.array_equal(got, expect)


@pytest.mark.parametrize("h,w,n_scales", [(64, 64, 2), (96, 80, 3), (70, 131, 4), (256, 256, 5)])
@pytest.mark.parametrize("masked", [False, True])
@pytest.mark.parametrize("include_largest", [True, False])
def test_recompose_bit_exact(env, h, w, n_scales, masked, include_largest):
    torch, oracle, _, alg = env
    img = image(h, w, 7 + h)
    coef = oracle.iuwt_decompose(img, n_scales, True)
    masks = (np.random.default_rng(3).random((n_scales, h, w)) > 0.4).astype(np.uint8) if masked else None
    expect = oracle.iuwt_recompose(coef, n_scales, include_largest, masks)
    out = torch.empty((h, w), device="cuda")
    alg.iuwt_recompose(torch.from_numpy(coef).cuda(), out, n_scales, include_largest,
                       None if masks is None else torch.from_numpy(masks).cuda())
    torch.cuda.synchronize()  # kernel-level entry points are asynchronous on the handle's stream
    assert np.array_equal(out.cpu().numpy(), expect)


@pytest.mark.parametrize("simple", [False, True])
@pytest.mark.parametrize("scale", [1, 2, 4])
def test_smooth_bit_exact(env, simple, scale):
    torch, oracle, _, alg = env
    img = image(150, 200, scale)
    out = torch.empty((150, 200), device="cuda")
    alg.iuwt_smooth(torch.from_numpy(img).cuda(), out, scale, simple)
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), oracle.iuwt_smooth(img, scale, mt=not simple))


def test_golden_vectors_from_reference_sources(env):
    torch, _, _, alg = env
    g = np.load(GOLDEN)
    n_scales = int(g["n_scales"])
    got = gpu_decompose(env, g["image"], n_scales, True).cpu().numpy()
    assert np.array_equal(got, g["coefficients"])
    out = torch.empty(g["image"].shape, device="cuda")
    alg.iuwt_recompose(torch.from_numpy(g["coefficients"]).cuda(), out, n_scales, False,
                       torch.from_numpy(g["masks"]).cuda())
    torch.cuda.synchronize()
    assert np.array_equal(out.cpu().numpy(), g["recomposed_masked"])


def test_reference_sources_directly(env):
    _, oracle, _, _ = env
    if oracle.ref_lib() is None:
        pytest.skip("oracle/_ref not built")
    img = image(200, 264, 99)
    assert np.array_equal(gpu_decompose(env, img, 4, True).cpu().numpy(),
                          oracle.ref_iuwt_decompose(img, 4, True))


def test_scale_too_large_is_an_error(env):
    torch, _, rdl, alg = env
    with pytest.raises(rdl.RadlerError):
        gpu_decompose(env, image(32, 32, 1), 5)


def test_k3_size_properties(env):
    """2048^2, 6 scales (BASELINE config 3): Recompose(Decompose(x)) == x to round-off and the
    transform is linear: D(a x + b y) == a D(x) + b D(y) within fp32 tolerance."""
    torch, _, _, alg = env
    n, s = 2048, 6
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn((n, n), device="cuda", generator=g)
    y = torch.randn((n, n), device="cuda", generator=g)
    cx, cy, cz = (torch.empty((s + 1, n, n), device="cuda") for _ in range(3))
    alg.iuwt_decompose(x, cx, s, True)
    alg.iuwt_decompose(y, cy, s, True)
    z = (2.0 * x - 0.5 * y).contiguous()
    alg.iuwt_decompose(z, cz, s, True)
    torch.cuda.synchronize()
    assert float((cz - (2.0 * cx - 0.5 * cy)).abs().max()) < 2e-5
    rec = torch.empty((n, n), device="cuda")
    alg.iuwt_recompose(cx, rec, s, True)
    torch.cuda.synchronize()
    assert float((rec - x).abs().max()) < 2e-5
