"""The register-radix FFT kernels (csrc/fft_fast.cuh) run on the CPU: their pass bodies are
__host__ __device__ functions of (thread, thread count), and tests/fft_emulation.cu executes them
thread by thread and phase by phase with the launch geometry of csrc/fft.cu. The result is compared
with numpy's float64 FFT here. Catches index / pairing / twiddle errors without a GPU; the GPU run
itself is covered by tests/test_gpu_fft.py."""
import os
import subprocess
import tempfile

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "_fft_emulation")
SRC = os.path.join(ROOT, "tests", "fft_emulation.cu")
DEPS = [SRC] + [os.path.join(ROOT, "ska_sdp_func_radler_b200", "csrc", f)
                for f in ("fft_fast.cuh", "fft_fast_plan.h")]


@pytest.fixture(scope="module")
def emulator():
    if not os.path.exists(EXE) or any(os.path.getmtime(d) > os.path.getmtime(EXE) for d in DEPS):
        subprocess.run(["nvcc", "-std=c++17", "-O1", "--expt-relaxed-constexpr",
                        "-Wno-deprecated-gpu-targets", "-x", "cu", SRC, "-o", EXE], check=True)
    return EXE


CASES = [
    # W, H, source, destination, epilogue
    (2, 1, None, None, 0), (4, 3, None, None, 0), (16, 16, (13, 14), (12, 11), 1),
    (42, 30, (39, 28), (38, 25), 1), (64, 64, (60, 60), (60, 60), 2), (128, 18, None, None, 0),
    (36, 50, (33, 48), (32, 45), 1), (250, 14, None, None, 0), (96, 240, (92, 236), (92, 236), 2),
    (360, 336, (357, 334), (356, 331), 1), (512, 28, None, None, 0), (28, 1024, None, None, 0),
    (70, 1296, (66, 1292), (66, 1292), 2),
    # the row and column lengths of the BASELINE frames (other axis kept short)
    (4096, 2, None, None, 0), (4536, 2, None, None, 0), (4608, 2, None, None, 0),
    (4704, 2, None, None, 0), (4860, 2, None, None, 0), (8192, 2, None, None, 0),
    (9072, 2, None, None, 0), (9216, 2, (8192, 2), (8192, 2), 1), (9408, 2, None, None, 0),
    (16384, 2, None, None, 0), (32768, 2, None, None, 0),
    (32, 4096, None, None, 0), (32, 4536, None, None, 0), (32, 4608, None, None, 0),
    (32, 4704, None, None, 0), (32, 4860, None, None, 0), (32, 8192, None, None, 0),
    (32, 9072, None, None, 0), (32, 9216, (32, 8192), (32, 8192), 2), (32, 9408, None, None, 0),
    (32, 16384, None, None, 0), (32, 32768, None, None, 0),
]


# the whole-column kernel (H = ra * rb * rc in shared memory, 2 or 4 columns per tile): every
# column height of the BASELINE frames and their paddings, partial source / destination rows
WHOLE_CASES = [
    (32, 2048, None, None, 0), (32, 2304, (32, 2048), (32, 2048), 1), (32, 4096, None, None, 0),
    (32, 4536, None, None, 0), (32, 4608, (30, 4096), (30, 4096), 2), (32, 4704, None, None, 0),
    (32, 4860, (32, 4096), (32, 4000), 1), (32, 8192, None, None, 0), (32, 9072, None, None, 0),
    (32, 9216, (32, 8192), (32, 8192), 1), (32, 9408, (32, 8192), (32, 8192), 2),
    (32, 9720, None, None, 0),
]


@pytest.mark.parametrize("w,h,src,dst,mode", CASES)
def test_emulated_kernels_match_float64_fft(emulator, w, h, src, dst, mode):
    _run_case(emulator, w, h, src, dst, mode, False)


@pytest.mark.parametrize("w,h,src,dst,mode", WHOLE_CASES)
def test_emulated_whole_column_kernel_matches_float64_fft(emulator, w, h, src, dst, mode):
    _run_case(emulator, w, h, src, dst, mode, True)


def _run_case(emulator, w, h, src, dst, mode, whole):
    sw, sh = src or (w, h)
    dw, dh = dst or (w, h)
    with tempfile.TemporaryDirectory() as tmp:
        path = os.path.join(tmp, "dump.bin")
        r = subprocess.run([emulator] + (["whole"] if whole else []) +
                           ["dump", str(w), str(h), str(sw), str(sh), str(dw), str(dh),
                            str(mode), str(w * 7 + h), path], capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr
        data = np.fromfile(path, np.float32)
    a, b, c = sw * sh, sw * sh + w * h, sw * sh + w * h + dw * dh
    s, kernel = data[:a].reshape(sh, sw), data[a:b].reshape(h, w)
    dst0, got = data[b:c].reshape(dh, dw), data[c:].reshape(dh, dw)
    frame = np.zeros((h, w))
    ox, oy = (w - sw) // 2, (h - sh) // 2
    frame[oy:oy + sh, ox:ox + sw] = s
    conv = np.fft.irfft2(np.fft.rfft2(frame) * np.fft.rfft2(kernel.astype(np.float64)), s=(h, w))
    ox, oy = (w - dw) // 2, (h - dh) // 2
    want = conv[oy:oy + dh, ox:ox + dw]
    expect = want if mode == 0 else (dst0 - want if mode == 1 else dst0 + want)
    err = np.abs(got - expect).max() / np.abs(want).max()
    assert err < 2e-6, (r.stdout.strip(), err)
