"""Facet geometry of the product (csrc/facets.cc, host code inside libradler_b200.so; no GPU
needed) against (i) the known answer of the reference's own test
(cpp/math/test/test_dijkstra_splitter.cc:95-121, "vertical") and (ii) the reference's own
dijkstra_splitter.cc compiled in oracle/_ref, pixel for pixel on random and tie-heavy images."""
import ctypes as C

import numpy as np
import pytest

import oracle
import ska_sdp_func_radler_b200 as rdl

needs_ref = pytest.mark.skipif(oracle.ref_lib() is None, reason="oracle/_ref not built")


def ascii_image(width, text):
    a = np.array([0.1 if ch == "X" else 10.0 for ch in text], np.float32)
    return a.reshape(-1, width)


def path_str(img):
    return "\n".join("".join(" " if v == 0 else "X" for v in row) for row in img) + "\n"


def test_vertical_known_answer():
    img = ascii_image(10, "X         "
                          " X        "
                          "  X       "
                          "   XXX    "
                          "     X    "
                          "         X"
                          "   X      "
                          "    XXXX  "
                          "        X "
                          "      XX  ")
    out = rdl.facet_dividing_path(img, 0, 10, vertical=True)
    assert path_str(out) == ("X         \n"
                             " X        \n"
                             "  X       \n"
                             "   XX     \n"
                             "     X    \n"
                             "    X     \n"
                             "   X      \n"
                             "    XXXX  \n"
                             "        X \n"
                             "       X  \n")
    # the transposed problem gives the transposed path
    out_h = rdl.facet_dividing_path(np.ascontiguousarray(img.T), 0, 10, vertical=False)
    assert np.array_equal(out_h, out.T)


def images():
    rng = np.random.default_rng(4)
    yield rng.standard_normal((60, 84)).astype(np.float32)
    yield np.zeros((40, 52), np.float32)                      # every path ties
    yield (rng.integers(0, 3, (48, 64)) * 0.5).astype(np.float32)  # many exact ties
    big = rng.standard_normal((96, 128)).astype(np.float32)
    big[30:60, 40:90] += 5.0
    yield big


def ref_divide(img, lo, hi, vertical):
    L = oracle.ref_lib()
    out = np.zeros_like(img)
    fn = L.ref_divide_vertically if vertical else L.ref_divide_horizontally
    fn(img.ctypes.data, out.ctypes.data, img.shape[1], img.shape[0], lo, hi)
    return out


@needs_ref
def test_dividing_paths_match_reference_source():
    for img in images():
        h, w = img.shape
        for (lo, hi, vertical) in [(0, w, True), (w // 4, w // 2, True), (0, h, False),
                                   (h // 3, h // 3 + 7, False), (5, 6, True)]:
            got = rdl.facet_dividing_path(img, lo, hi, vertical)
            assert np.array_equal(got, ref_divide(img, lo, hi, vertical)), (img.shape, lo, hi, vertical)


@needs_ref
def test_long_unique_paths_match_reference_source():
    """Real-valued images have no exact ties, so the product takes its queue-free route
    (facets.cc UniqueDividingPath); the path must still be the reference's, pixel for pixel."""
    rng = np.random.default_rng(21)
    noise = rng.standard_normal((700, 900)).astype(np.float32)
    yy, xx = np.mgrid[0:512, 0:640]
    smooth = (np.sin(xx / 37.0) * np.cos(yy / 53.0)
              + 0.01 * rng.standard_normal((512, 640))).astype(np.float32)
    for img in (noise, smooth):
        h, w = img.shape
        for (lo, hi, vertical) in [(w // 4, 3 * w // 4, True), (h // 4, 3 * h // 4, False),
                                   (0, w, True), (10, 13, False)]:
            got = rdl.facet_dividing_path(img, lo, hi, vertical)
            assert np.array_equal(got, ref_divide(img, lo, hi, vertical)), (img.shape, lo, hi, vertical)
    # a strip with a few exact ties inside an otherwise real-valued image: queue route
    tied = noise.copy()
    tied[100:140, 300:500] = 0.0
    got = rdl.facet_dividing_path(tied, 250, 550, True)
    assert np.array_equal(got, ref_divide(tied, 250, 550, True))
    bad = noise.copy()
    bad[50, 400] = np.inf
    got = rdl.facet_dividing_path(bad, 250, 550, True)
    assert np.array_equal(got, ref_divide(bad, 250, 550, True))


@needs_ref
def test_flood_and_bounding_mask_match_reference_source():
    L = oracle.ref_lib()
    u64 = C.c_uint64
    for img in images():
        h, w = img.shape
        div_v = np.zeros_like(img)
        for d in (1, 2, 3):
            mid = w * d // 4
            div_v += rdl.facet_dividing_path(img, mid - w // 16, mid + w // 16, True)
        div_h = np.zeros_like(img)
        for d in (1, 2):
            mid = h * d // 3
            div_h += rdl.facet_dividing_path(img, mid - h // 12, mid + h // 12, False)
        for gx in range(4):
            inside = gx * w // 4 + w // 8
            vm, vx, vw = rdl.facet_flood_area(div_v, inside, True)
            rm = np.zeros((h, w), np.uint8)
            rx, rw = u64(), u64()
            L.ref_flood_vertical_area(div_v.ctypes.data, w, h, inside, rm.ctypes.data,
                                      C.byref(rx), C.byref(rw))
            assert (vx, vw) == (rx.value, rw.value) and np.array_equal(vm, rm)
            vcrop = np.ascontiguousarray(vm[:, vx:vx + vw])
            for gy in range(3):
                inside_y = gy * h // 3 + h // 6
                hm, hy, hh = rdl.facet_flood_area(div_h, inside_y, False)
                rm2 = np.zeros((h, w), np.uint8)
                ry, rh = u64(), u64()
                L.ref_flood_horizontal_area(div_h.ctypes.data, w, h, inside_y, rm2.ctypes.data,
                                            C.byref(ry), C.byref(rh))
                assert (hy, hh) == (ry.value, rh.value) and np.array_equal(hm, rm2)
                mask, box = rdl.facet_bounding_mask(vcrop, vx, hm)
                rmask = np.zeros((h, w), np.uint8)
                b = [u64() for _ in range(4)]
                L.ref_get_bounding_mask(vcrop.ctypes.data, w, h, vx, vw, hm.ctypes.data,
                                        rmask.ctypes.data, *[C.byref(v) for v in b])
                assert box == tuple(v.value for v in b)
                assert np.array_equal(mask, rmask)


def test_facet_grid_tiles_the_frame():
    rng = np.random.default_rng(8)
    img = rng.standard_normal((128, 160)).astype(np.float32)
    boxes = rdl.facet_grid(img, 4, 2)
    assert len(boxes) == 8
    cover = np.zeros(img.shape, np.int32)
    for (x, y, w, h) in boxes:
        assert w % 2 == 0 and h % 2 == 0 and x + w <= 160 and y + h <= 128
        cover[y:y + h, x:x + w] += 1
    assert cover.min() >= 1  # bounding boxes overlap, together they cover the frame


@pytest.mark.parametrize("shape,grid", [((128, 160), (4, 2)), ((96, 96), (3, 3)), ((131, 97), (2, 3)),
                                        ((200, 256), (5, 4))])
def test_facet_grid_equals_the_reference_steps(shape, grid):
    """MakeSubImages (parallel_deconvolution.cc:69-168) builds the facets from the splitter's
    steps; the product computes them from per-line runs instead of frame-sized masks. Boxes must
    equal the composition of the individual steps, each of which is pinned to the reference's
    source above (even / odd frame sizes: the even-size padding of the boxes)."""
    rng = np.random.default_rng(shape[0] + grid[0])
    img = rng.standard_normal(shape).astype(np.float32)
    h, w = shape
    gw, gh = grid
    avg_w, avg_h = w // gw, h // gh
    div_v = np.zeros_like(img)
    for d in range(1, gw):
        mid = w * d // gw
        div_v += rdl.facet_dividing_path(img, mid - avg_w // 4, mid + avg_w // 4, True)
    div_h = np.zeros_like(img)
    for d in range(1, gh):
        mid = h * d // gh
        div_h += rdl.facet_dividing_path(img, mid - avg_h // 4, mid + avg_h // 4, False)
    want = []
    for gy in range(gh):
        hm, _, _ = rdl.facet_flood_area(div_h, gy * h // gh + avg_h // 2, False)
        for gx in range(gw):
            vm, vx, vw = rdl.facet_flood_area(div_v, gx * w // gw + avg_w // 2, True)
            _, box = rdl.facet_bounding_mask(np.ascontiguousarray(vm[:, vx:vx + vw]), vx, hm)
            want.append(tuple(int(v) for v in box))
    got = [tuple(int(v) for v in b) for b in rdl.facet_grid(img, gw, gh)]
    assert got == want
