"""rdl_iuwt_decompose_static (csrc/static_scratch.cu): the IUWT decomposition with its scratch
planes in a static device array of the library, for callers that may not call cudaMalloc. No
counterpart in the reference. Checked through the C ABI against rdl_iuwt_decompose, BIT-EXACT
(the tolerance of the IUWT kernel tests), on calls in a row that share the static array.

Collected last on purpose."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_PIXELS = 1 << 20


def _both(torch, alg, img, n_scales, include_largest):
    d = torch.from_numpy(img).cuda()
    planes = n_scales + (1 if include_largest else 0)
    expect = torch.full((planes,) + img.shape, 7.0, device="cuda")
    got = torch.full((planes,) + img.shape, -3.0, device="cuda")
    alg.iuwt_decompose(d, expect, n_scales, include_largest)
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    alg.iuwt_decompose_static(d, got, n_scales, include_largest)
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    return expect.cpu().numpy(), got.cpu().numpy(), free_before, free_after


def test_static_scratch_decompose_matches_allocating_entry_point(rdl, torch_cuda):
    torch = torch_cuda
    alg = rdl.DeconvolutionAlgorithm(rdl.IUWT)
    rng = np.random.default_rng(5)
    # calls in a row on one device: every one finds the previous one's data in the array
    cases = [(300, 517, 5, True), (64, 96, 2, False), (1024, 1024, 6, True), (257, 131, 3, True),
             (300, 517, 5, True)]
    frees = []
    for h, w, n_scales, largest in cases:
        img = rng.standard_normal((h, w)).astype(np.float32)
        img[h // 3, w // 2] += 50.0
        expect, got, before, after = _both(torch, alg, img, n_scales, largest)
        assert np.array_equal(expect.view(np.uint32), got.view(np.uint32)), (h, w, n_scales)
        frees.append((before, after))
    # from the second call on nothing can be left to load lazily: device memory does not move
    for before, after in frees[1:]:
        assert after == before


def test_static_scratch_two_handles_same_array(rdl, torch_cuda):
    torch = torch_cuda
    a, b = rdl.DeconvolutionAlgorithm(rdl.IUWT), rdl.DeconvolutionAlgorithm(rdl.IUWT)
    rng = np.random.default_rng(6)
    for alg, shape in [(a, (200, 200)), (b, (128, 384)), (a, (128, 384))]:
        img = rng.standard_normal(shape).astype(np.float32)
        expect, got, _, _ = _both(torch, alg, img, 4, True)
        assert np.array_equal(expect.view(np.uint32), got.view(np.uint32))


def test_static_scratch_rejects_large_images(rdl, torch_cuda):
    torch = torch_cuda
    alg = rdl.DeconvolutionAlgorithm(rdl.IUWT)
    img = torch.zeros((1024, 1025), device="cuda")
    assert img.numel() > MAX_PIXELS
    coef = torch.zeros((3, 1024, 1025), device="cuda")
    with pytest.raises(Exception) as e:
        alg.iuwt_decompose_static(img, coef, 2, True)
    assert "static scratch" in str(e.value)
    # and the handle stays usable
    small = torch.ones((64, 64), device="cuda")
    out = torch.zeros((3, 64, 64), device="cuda")
    alg.iuwt_decompose_static(small, out, 2, True)
    torch.cuda.synchronize()
