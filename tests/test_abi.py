"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol
include/radler_b200.h declares, and refuses to compute without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "radler_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(rdl_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_are_exported(rdl):
    declared = _declared_symbols()
    assert len(declared) >= 35
    lib = C.CDLL(rdl.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in include/radler_b200.h but not exported"
    assert sorted(rdl.EXPORTED_SYMBOLS) == declared


def test_no_torch_or_oracle_in_product():
    # the product never links or imports the checker
    import subprocess
    out = subprocess.run(["ldd", os.path.join(ROOT, "ska_sdp_func_radler_b200", "libradler_b200.so")],
                         capture_output=True, text=True).stdout
    assert "oracle" not in out and "torch" not in out
    for dirpath, _, files in os.walk(os.path.join(ROOT, "ska_sdp_func_radler_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cc", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert "import oracle" not in src and "liboracle" not in src, f


def test_fft_size_helpers(rdl):
    # cpp/utils/test/test_fft_size_calculations.cc:9-50 through the C ABI (host-only maths)
    for n in range(1154, 1177):
        assert rdl.good_fft_size(n) == 1176
    assert rdl.convolution_size(10, 1010, 1.1) == 1134
    assert rdl.convolution_size(0, 1010, 1.1) == 1120
    assert rdl.convolution_size(0, 1154, 1.0) == 1176


def test_fails_loudly_without_gpu(rdl):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    assert rdl.device_count() == 0
    with pytest.raises(rdl.RadlerError) as e:
        rdl.GenericClean()
    assert e.value.status == rdl.RDL_ERR_NO_DEVICE
    assert "no CPU fallback" in str(e.value)


def test_plane_pointer_tables():
    """Cubes are passed as one contiguous [n, H, W] array or as a list of separately allocated
    planes (what an ImageSet of aocommon::Images is); both give the same pointer table."""
    import ctypes

    import numpy as np

    import ska_sdp_func_radler_b200 as rdl
    cube = np.arange(3 * 4 * 5, dtype=np.float32).reshape(3, 4, 5)
    table, space = rdl._plane_pointers(cube, 3, 20)
    assert space == rdl.MEM_HOST
    assert [table[i] for i in range(3)] == [cube.ctypes.data + i * 80 for i in range(3)]
    planes = [np.ascontiguousarray(cube[i]) for i in range(3)]
    table2, space2 = rdl._plane_pointers(planes, 3, 20)
    assert space2 == rdl.MEM_HOST
    assert [table2[i] for i in range(3)] == [p.ctypes.data for p in planes]
    for i in range(3):  # the table really addresses the planes
        first = ctypes.cast(table2[i], ctypes.POINTER(ctypes.c_float))[0]
        assert first == planes[i][0, 0]
    with pytest.raises(ValueError):
        rdl._plane_pointers(planes[:2], 3, 20)
    with pytest.raises(ValueError):
        rdl._plane_pointers([p.astype(np.float64) for p in planes], 3, 20)
