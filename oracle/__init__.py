"""ORACLE — TEST INFRASTRUCTURE ONLY.

ctypes bindings of ``oracle/liboracle.so`` (the CPU restatement of Radler's minor
loops; see ``oracle/oracle.h``). Importable only from ``tests/``,
``__graft_entry__.smoke()`` and ``bench.py``'s cpu_baseline / ``--impl reference``
leg. The product package never imports this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass, field

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")


_REF_PATH = os.path.join(_HERE, "_ref", "libradler_ref.so")
_REFERENCE_ROOT = "/root/reference"


def build(force: bool = False) -> str:
    """Compile liboracle.so (and oracle/_ref when /root/reference exists)."""
    if force or not os.path.exists(_LIB_PATH):
        subprocess.run(["make", "-C", _HERE, "-B" if force else "-s", "liboracle.so"], check=True)
    if os.path.isdir(os.path.join(_REFERENCE_ROOT, "cpp")) and (force or not os.path.exists(_REF_PATH)):
        subprocess.run(["make", "-C", _HERE, "-s", "ref"], check=True)
    return _LIB_PATH


_ref = None


def ref_lib():
    """oracle/_ref/libradler_ref.so: the reference's own peak_finder.cc, simple_clean.cc and
    dijkstra_splitter.cc compiled where they lie. None when it was never built (it is
    prebuilt in the dev container and travels to the GPU box as a built file)."""
    global _ref
    if _ref is None:
        if not os.path.exists(_REF_PATH):
            if not os.path.isdir(os.path.join(_REFERENCE_ROOT, "cpp")):
                return None
            build()
        L = C.CDLL(_REF_PATH)
        u64, vp = C.c_uint64, C.c_void_p
        L.ref_find_peak.restype = C.c_int32
        L.ref_find_peak.argtypes = [C.c_int32, vp, u64, u64, C.c_int32, u64, u64, u64, u64, vp,
                                    C.POINTER(u64), C.POINTER(u64), C.POINTER(C.c_float)]
        L.ref_subtract_image.argtypes = [vp, vp, u64, u64, u64, u64, C.c_float]
        L.ref_partial_subtract_image.argtypes = [vp, u64, u64, vp, u64, u64, u64, u64, C.c_float,
                                                 u64, u64]
        L.ref_good_fft_size.restype = u64
        L.ref_good_fft_size.argtypes = [u64]
        L.ref_convolution_size.restype = u64
        L.ref_convolution_size.argtypes = [C.c_double, u64, C.c_double]
        for n in ["ref_divide_vertically", "ref_divide_horizontally"]:
            getattr(L, n).argtypes = [vp, vp, u64, u64, u64, u64]
        for n in ["ref_add_vertical_divider", "ref_add_horizontal_divider"]:
            getattr(L, n).argtypes = [vp, vp, vp, u64, u64, u64, u64]
        for n in ["ref_flood_vertical_area", "ref_flood_horizontal_area"]:
            getattr(L, n).argtypes = [vp, u64, u64, u64, vp, C.POINTER(u64), C.POINTER(u64)]
        L.ref_get_bounding_mask.argtypes = [vp, u64, u64, u64, u64, vp, vp] + [C.POINTER(u64)] * 4
        L.ref_iuwt_decompose.argtypes = [vp, u64, u64, C.c_int32, C.c_int32, vp]
        L.ref_iuwt_recompose.argtypes = [vp, u64, u64, C.c_int32, C.c_int32, vp, vp]
        L.ref_iuwt_smooth.argtypes = [vp, u64, u64, C.c_int32]
        i32, f32 = C.c_int32, C.c_float
        for pre in ["ref_"]:
            f = getattr(L, pre + "iuwt_select_structures")
            f.restype = u64
            f.argtypes = [vp, vp, u64, u64, i32, vp, i32, i32, f32, vp]
            f = getattr(L, pre + "iuwt_floodfill")
            f.restype = u64
            f.argtypes = [vp, vp, u64, u64, i32, vp, i32, i32, u64, u64, i32, f32, vp]
            f = getattr(L, pre + "iuwt_floodfill_2d")
            f.restype = u64
            f.argtypes = [vp, vp, u64, u64, f32, i32, u64, u64]
            f = getattr(L, pre + "iuwt_max_abs")
            f.restype = f32
            f.argtypes = [vp, u64, u64, f32, i32, vp, C.POINTER(u64), C.POINTER(u64)]
        L.ref_iuwt_mad.restype = f32
        L.ref_iuwt_mad.argtypes = [vp, u64, u64]
        _ref = L
    return _ref


def ref_find_peak(image, variant=1, allow_negative=True, start_y=0, end_y=None, h_border=0,
                  v_border=0, mask=None):
    """Reference peak_finder: variant 0 Simple, 1 Avx, 2 FindWithMask."""
    image = np.ascontiguousarray(image, np.float32)
    h, w = image.shape
    m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
    x, y, v = C.c_uint64(), C.c_uint64(), C.c_float()
    found = ref_lib().ref_find_peak(variant, image.ctypes.data, w, h, int(allow_negative), start_y,
                                    h if end_y is None else end_y, h_border, v_border,
                                    None if m is None else m.ctypes.data, C.byref(x), C.byref(y),
                                    C.byref(v))
    return (v.value if found else None), x.value, y.value


def ref_partial_subtract_image(image, psf, x, y, factor, start_y, end_y):
    assert image.dtype == np.float32 and image.flags.c_contiguous
    psf = np.ascontiguousarray(psf, np.float32)
    ref_lib().ref_partial_subtract_image(image.ctypes.data, image.shape[1], image.shape[0],
                                         psf.ctypes.data, psf.shape[1], psf.shape[0], x, y, factor,
                                         start_y, end_y)


def ref_subtract_image(image, psf, x, y, factor):
    assert image.dtype == np.float32 and image.flags.c_contiguous
    psf = np.ascontiguousarray(psf, np.float32)
    ref_lib().ref_subtract_image(image.ctypes.data, psf.ctypes.data, image.shape[1],
                                 image.shape[0], x, y, factor)


class _Cube(C.Structure):
    _fields_ = [("width", C.c_uint64), ("height", C.c_uint64), ("n_channels", C.c_uint64),
                ("n_pol", C.c_uint64), ("weights", C.POINTER(C.c_float)),
                ("pol_linked", C.POINTER(C.c_uint8)), ("pol_norm", C.c_float),
                ("squared_joins", C.c_int32)]


class _Settings(C.Structure):
    _fields_ = [("threshold", C.c_float), ("major_iteration_threshold", C.c_float),
                ("minor_loop_gain", C.c_float), ("major_loop_gain", C.c_float),
                ("clean_border_ratio", C.c_float), ("divergence_limit", C.c_float),
                ("max_iterations", C.c_uint64), ("iteration_number", C.c_uint64),
                ("allow_negative", C.c_int32), ("stop_on_negative", C.c_int32),
                ("clean_mask", C.POINTER(C.c_uint8)), ("rms_factor", C.POINTER(C.c_float)),
                ("spectral_mode", C.c_int32), ("n_terms", C.c_uint64),
                ("frequencies", C.POINTER(C.c_double)), ("fit_weights", C.POINTER(C.c_float)),
                ("record_trace", C.c_int32)]


class _Result(C.Structure):
    _fields_ = [("has_starting_peak", C.c_int32), ("starting_peak_value", C.c_float),
                ("final_peak_value", C.c_float), ("another_iteration_required", C.c_int32),
                ("is_diverging", C.c_int32), ("iteration_number", C.c_uint64),
                ("n_trace", C.c_uint64), ("n_selected", C.c_uint64)]


class _MsConfig(C.Structure):
    _fields_ = [("fast_sub_minor_loop", C.c_int32), ("sub_minor_loop_gain", C.c_double),
                ("scale_bias", C.c_double), ("max_scales", C.c_uint64),
                ("convolution_padding", C.c_double), ("scale_list", C.POINTER(C.c_double)),
                ("n_scale_list", C.c_uint64), ("shape", C.c_int32), ("beam_size", C.c_double),
                ("pixel_scale_x", C.c_double), ("pixel_scale_y", C.c_double)]


class _ScaleInfo(C.Structure):
    _fields_ = [("scale", C.c_float), ("psf_peak", C.c_float), ("kernel_peak", C.c_float),
                ("bias_factor", C.c_float), ("gain", C.c_float),
                ("max_normalized_image_value", C.c_float),
                ("max_unnormalized_image_value", C.c_float), ("rms", C.c_float),
                ("max_x", C.c_uint64), ("max_y", C.c_uint64), ("is_active", C.c_int32),
                ("n_components_cleaned", C.c_uint64), ("total_flux_cleaned", C.c_float)]


TRACE_DTYPE = np.dtype([("iteration", "<u4"), ("scale", "<u4"), ("x", "<u4"), ("y", "<u4"),
                        ("value", "<f4")])

_lib = None


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_LIB_PATH)
        L.orc_create_generic.restype = C.c_void_p
        L.orc_create_generic.argtypes = [C.c_int32]
        L.orc_create_multiscale.restype = C.c_void_p
        L.orc_create_multiscale.argtypes = [C.POINTER(_MsConfig)]
        L.orc_destroy.argtypes = [C.c_void_p]
        L.orc_last_error.restype = C.c_char_p
        L.orc_last_error.argtypes = [C.c_void_p]
        L.orc_execute.restype = C.c_int32
        L.orc_execute.argtypes = [C.c_void_p, C.POINTER(_Cube), C.POINTER(_Settings), C.c_void_p,
                                  C.c_void_p, C.c_void_p, C.POINTER(_Result)]
        L.orc_get_trace.restype = C.c_uint64
        L.orc_get_trace.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
        L.orc_ms_set_automask_mode.argtypes = [C.c_void_p, C.c_int32, C.c_int32]
        L.orc_ms_scale_count.restype = C.c_uint64
        L.orc_ms_scale_count.argtypes = [C.c_void_p]
        L.orc_ms_minor_iterations.restype = C.c_uint64
        L.orc_ms_minor_iterations.argtypes = [C.c_void_p]
        L.orc_ms_scale_info.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(_ScaleInfo)]
        L.orc_ms_scale_mask.restype = C.c_int32
        L.orc_ms_scale_mask.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p]
        L.orc_set_threads.argtypes = [C.c_uint64]
        L.orc_get_threads.restype = C.c_uint64
        L.orc_find_peak.restype = C.c_int32
        L.orc_find_peak.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_int32, C.c_uint64,
                                    C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p,
                                    C.POINTER(C.c_uint64), C.POINTER(C.c_uint64),
                                    C.POINTER(C.c_float)]
        L.orc_subtract_image.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64,
                                         C.c_uint64, C.c_uint64, C.c_float]
        L.orc_partial_subtract_image.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p,
                                                 C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64,
                                                 C.c_float, C.c_uint64, C.c_uint64]
        L.orc_integrate.argtypes = [C.POINTER(_Cube), C.c_void_p, C.c_int32, C.c_void_p]
        L.orc_integrated_psf.argtypes = [C.POINTER(_Cube), C.c_void_p, C.c_void_p]
        L.orc_spectral_fit.argtypes = [C.c_int32, C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p,
                                       C.c_void_p, C.c_void_p]
        L.orc_polynomial_projection.argtypes = [C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p,
                                                C.c_void_p]
        L.orc_convolve.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64]
        L.orc_prepare_convolution_kernel.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64,
                                                     C.c_uint64]
        L.orc_prepare_small_convolution_kernel.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64,
                                                           C.c_void_p, C.c_uint64]
        L.orc_fft1d.argtypes = [C.c_void_p, C.c_uint64, C.c_int32]
        L.orc_trim.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_uint64,
                               C.c_uint64]
        L.orc_untrim.argtypes = L.orc_trim.argtypes
        L.orc_good_fft_size.restype = C.c_uint64
        L.orc_good_fft_size.argtypes = [C.c_uint64]
        L.orc_convolution_size.restype = C.c_uint64
        L.orc_convolution_size.argtypes = [C.c_double, C.c_uint64, C.c_double]
        L.orc_make_shape_function.restype = C.c_uint64
        L.orc_make_shape_function.argtypes = [C.c_float, C.c_uint64, C.c_int32, C.c_void_p]
        L.orc_ms_transform.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_float,
                                       C.c_int32]
        L.orc_add_shape_component.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_float,
                                              C.c_uint64, C.c_uint64, C.c_float, C.c_int32]
        L.orc_iuwt_smooth.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_int32,
                                      C.c_int32]
        L.orc_iuwt_decompose.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_int32, C.c_int32,
                                         C.c_void_p]
        L.orc_iuwt_recompose.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_int32, C.c_int32,
                                         C.c_void_p, C.c_void_p]
        i32, f32 = C.c_int32, C.c_float
        for pre in ["orc_"]:
            f = getattr(L, pre + "iuwt_select_structures")
            f.restype = C.c_uint64
            f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, i32, C.c_void_p, i32, i32, f32, C.c_void_p]
            f = getattr(L, pre + "iuwt_floodfill")
            f.restype = C.c_uint64
            f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, i32, C.c_void_p, i32, i32, C.c_uint64, C.c_uint64, i32, f32, C.c_void_p]
            f = getattr(L, pre + "iuwt_floodfill_2d")
            f.restype = C.c_uint64
            f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, f32, i32, C.c_uint64, C.c_uint64]
            f = getattr(L, pre + "iuwt_max_abs")
            f.restype = f32
            f.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, f32, i32, C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        L.orc_iuwt_mad.restype = f32
        L.orc_iuwt_mad.argtypes = [C.c_void_p, C.c_uint64]
        _lib = L
    return _lib


def _f32(a):
    return np.ascontiguousarray(a, dtype=np.float32)


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


@dataclass
class Cube:
    """Layout metadata of an ImageSet: images ordered [channel][pol]."""
    width: int
    height: int
    n_channels: int = 1
    n_pol: int = 1
    weights: np.ndarray | None = None
    pol_linked: np.ndarray | None = None
    pol_norm: float = 1.0
    squared_joins: bool = False
    _keep: list = field(default_factory=list, repr=False)

    def c(self) -> _Cube:
        w = None if self.weights is None else _f32(self.weights)
        p = None if self.pol_linked is None else np.ascontiguousarray(self.pol_linked, np.uint8)
        self._keep = [w, p]
        return _Cube(self.width, self.height, self.n_channels, self.n_pol,
                     None if w is None else w.ctypes.data_as(C.POINTER(C.c_float)),
                     None if p is None else p.ctypes.data_as(C.POINTER(C.c_uint8)),
                     self.pol_norm, int(self.squared_joins))


@dataclass
class Settings:
    """Mirror of DeconvolutionAlgorithm's settings (deconvolution_algorithm.h:189-202)."""
    threshold: float = 0.0
    major_iteration_threshold: float = 0.0
    minor_loop_gain: float = 0.1
    major_loop_gain: float = 1.0
    clean_border_ratio: float = 0.05
    divergence_limit: float = 4.0
    max_iterations: int = 500
    iteration_number: int = 0
    allow_negative: bool = True
    stop_on_negative: bool = False
    clean_mask: np.ndarray | None = None
    rms_factor: np.ndarray | None = None
    spectral_mode: int = 0
    n_terms: int = 0
    frequencies: np.ndarray | None = None
    fit_weights: np.ndarray | None = None
    record_trace: bool = False


@dataclass
class Result:
    starting_peak_value: float | None
    final_peak_value: float
    another_iteration_required: bool
    is_diverging: bool
    iteration_number: int
    trace: np.ndarray
    n_selected: int = 0


class Algorithm:
    def __init__(self, handle):
        self._h = handle

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_destroy(self._h)
            self._h = None

    def execute(self, cube: Cube, s: Settings, residual, model, psfs) -> Result:
        """Runs ExecuteMajorIteration; residual/model ([C,H,W] float32) are updated in place."""
        assert residual.dtype == np.float32 and residual.flags.c_contiguous
        assert model.dtype == np.float32 and model.flags.c_contiguous
        psfs = _f32(psfs)
        mask = None if s.clean_mask is None else np.ascontiguousarray(s.clean_mask, np.uint8)
        rms = None if s.rms_factor is None else _f32(s.rms_factor)
        fr = None if s.frequencies is None else np.ascontiguousarray(s.frequencies, np.float64)
        fw = None if s.fit_weights is None else _f32(s.fit_weights)
        cs = _Settings(s.threshold, s.major_iteration_threshold, s.minor_loop_gain,
                       s.major_loop_gain, s.clean_border_ratio, s.divergence_limit,
                       s.max_iterations, s.iteration_number, int(s.allow_negative),
                       int(s.stop_on_negative),
                       None if mask is None else mask.ctypes.data_as(C.POINTER(C.c_uint8)),
                       None if rms is None else rms.ctypes.data_as(C.POINTER(C.c_float)),
                       s.spectral_mode, s.n_terms,
                       None if fr is None else fr.ctypes.data_as(C.POINTER(C.c_double)),
                       None if fw is None else fw.ctypes.data_as(C.POINTER(C.c_float)),
                       int(s.record_trace))
        cc = cube.c()
        out = _Result()
        rc = lib().orc_execute(self._h, C.byref(cc), C.byref(cs), _ptr(residual), _ptr(model),
                               _ptr(psfs), C.byref(out))
        if rc != 0:
            raise RuntimeError(lib().orc_last_error(self._h).decode())
        trace = np.zeros(out.n_trace, TRACE_DTYPE)
        if out.n_trace:
            lib().orc_get_trace(self._h, _ptr(trace), out.n_trace)
        return Result(out.starting_peak_value if out.has_starting_peak else None,
                      out.final_peak_value, bool(out.another_iteration_required),
                      bool(out.is_diverging), out.iteration_number, trace, out.n_selected)


class GenericClean(Algorithm):
    def __init__(self, use_sub_minor_optimization: bool = True):
        super().__init__(lib().orc_create_generic(int(use_sub_minor_optimization)))


class MultiScale(Algorithm):
    def __init__(self, beam_size=0.0, pixel_scale=1.0, fast_sub_minor_loop=True,
                 sub_minor_loop_gain=0.2, scale_bias=0.6, max_scales=0, convolution_padding=1.1,
                 scale_list=None, shape=0):
        sl = None if not scale_list else np.ascontiguousarray(scale_list, np.float64)
        cfg = _MsConfig(int(fast_sub_minor_loop), sub_minor_loop_gain, scale_bias, max_scales,
                        convolution_padding,
                        None if sl is None else sl.ctypes.data_as(C.POINTER(C.c_double)),
                        0 if sl is None else len(sl), shape, beam_size, pixel_scale, pixel_scale)
        super().__init__(lib().orc_create_multiscale(C.byref(cfg)))

    def set_automask_mode(self, track: bool, use: bool):
        lib().orc_ms_set_automask_mode(self._h, int(track), int(use))

    def scale_infos(self):
        out = []
        for i in range(lib().orc_ms_scale_count(self._h)):
            s = _ScaleInfo()
            lib().orc_ms_scale_info(self._h, i, C.byref(s))
            out.append({k: getattr(s, k) for k, _ in _ScaleInfo._fields_})
        return out

    def minor_iterations(self) -> int:
        return lib().orc_ms_minor_iterations(self._h)

    def scale_mask(self, i, width, height):
        m = np.zeros((height, width), np.uint8)
        return m if lib().orc_ms_scale_mask(self._h, i, _ptr(m)) else None


# ---- leaf functions ----------------------------------------------------------------
def set_threads(n: int):
    lib().orc_set_threads(n)


def find_peak(image, allow_negative=True, start_y=0, end_y=None, h_border=0, v_border=0,
              mask=None):
    image = _f32(image)
    h, w = image.shape
    if end_y is None:
        end_y = h
    m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
    x, y, v = C.c_uint64(), C.c_uint64(), C.c_float()
    found = lib().orc_find_peak(_ptr(image), w, h, int(allow_negative), start_y, end_y, h_border,
                                v_border, _ptr(m), C.byref(x), C.byref(y), C.byref(v))
    return (v.value if found else None), x.value, y.value


def subtract_image(image, psf, x, y, factor):
    assert image.dtype == np.float32 and image.flags.c_contiguous
    psf = _f32(psf)
    h, w = image.shape
    lib().orc_subtract_image(_ptr(image), _ptr(psf), w, h, x, y, factor)


def subtract_image_whole(image, psf, x, y, factor):
    """simple_clean::SubtractImage (simple_clean.cc:17-59), incl. the dropped odd column."""
    assert image.dtype == np.float32 and image.flags.c_contiguous
    psf = _f32(psf)
    h, w = image.shape
    f = lib().orc_subtract_image_whole
    f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64, C.c_float]
    f(_ptr(image), _ptr(psf), w, h, x, y, factor)


def partial_subtract_image(image, psf, x, y, factor, start_y, end_y):
    assert image.dtype == np.float32 and image.flags.c_contiguous
    psf = _f32(psf)
    lib().orc_partial_subtract_image(_ptr(image), image.shape[1], image.shape[0], _ptr(psf),
                                     psf.shape[1], psf.shape[0], x, y, factor, start_y, end_y)


def integrate(cube: Cube, images, square=False):
    images = _f32(images)
    dest = np.zeros((cube.height, cube.width), np.float32)
    cc = cube.c()
    lib().orc_integrate(C.byref(cc), _ptr(images), int(square), _ptr(dest))
    return dest


def integrated_psf(cube: Cube, psfs):
    psfs = _f32(psfs)
    dest = np.zeros((cube.height, cube.width), np.float32)
    cc = cube.c()
    lib().orc_integrated_psf(C.byref(cc), _ptr(psfs), _ptr(dest))
    return dest


def spectral_fit(values, freqs, weights, n_terms, mode=1, n_pol=1):
    values = _f32(values).copy()
    freqs = np.ascontiguousarray(freqs, np.float64)
    weights = _f32(weights)
    lib().orc_spectral_fit(mode, n_terms, len(freqs), n_pol, _ptr(freqs), _ptr(weights),
                           _ptr(values))
    return values


def polynomial_projection(freqs, weights, n_terms):
    freqs = np.ascontiguousarray(freqs, np.float64)
    weights = _f32(weights)
    out = np.zeros((len(freqs), len(freqs)), np.float64)
    lib().orc_polynomial_projection(n_terms, len(freqs), _ptr(freqs), _ptr(weights), _ptr(out))
    return out


def convolve(image, kernel):
    image = _f32(image).copy()
    kernel = _f32(kernel)
    h, w = image.shape
    lib().orc_convolve(_ptr(image), _ptr(kernel), w, h)
    return image


def prepare_convolution_kernel(src):
    src = _f32(src)
    dst = np.zeros_like(src)
    lib().orc_prepare_convolution_kernel(_ptr(dst), _ptr(src), src.shape[1], src.shape[0])
    return dst


def prepare_small_convolution_kernel(kernel, w, h):
    kernel = _f32(kernel)
    dst = np.zeros((h, w), np.float32)
    lib().orc_prepare_small_convolution_kernel(_ptr(dst), w, h, _ptr(kernel), kernel.shape[0])
    return dst


def fft1d(z, inverse=False):
    z = np.ascontiguousarray(z, np.complex64).copy()
    lib().orc_fft1d(_ptr(z), len(z), int(inverse))
    return z


def trim(src, dw, dh):
    src = _f32(src)
    dst = np.zeros((dh, dw), np.float32)
    lib().orc_trim(_ptr(dst), dw, dh, _ptr(src), src.shape[1], src.shape[0])
    return dst


def untrim(src, dw, dh):
    src = _f32(src)
    dst = np.zeros((dh, dw), np.float32)
    lib().orc_untrim(_ptr(dst), dw, dh, _ptr(src), src.shape[1], src.shape[0])
    return dst


def iuwt_smooth(image, scale, mt=True):
    """One B3-spline a-trous smoothing (iuwt_decomposition.cc:55-67 convolveMT when mt, else the
    "simple" convolve of iuwt_decomposition.h:249-267)."""
    a = _f32(image)
    out = np.empty_like(a)
    lib().orc_iuwt_smooth(_ptr(out), _ptr(a), a.shape[1], a.shape[0], scale, int(mt))
    return out


def iuwt_decompose(image, n_scales, include_largest=True):
    """IuwtDecomposition::DecomposeMt (iuwt_decomposition.cc:9-53) -> [n_scales(+1), H, W]."""
    a = _f32(image)
    out = np.zeros((n_scales + (1 if include_largest else 0),) + a.shape, np.float32)
    lib().orc_iuwt_decompose(_ptr(a), a.shape[1], a.shape[0], n_scales, int(include_largest),
                             _ptr(out))
    return out


def iuwt_recompose(coefficients, n_scales, include_largest=True, masks=None):
    """ApplyMask (when masks) + IuwtDecomposition::Recompose (iuwt_decomposition.h:121-169)."""
    c = _f32(coefficients)
    out = np.empty(c.shape[1:], np.float32)
    m = None if masks is None else np.ascontiguousarray(masks, np.uint8)
    lib().orc_iuwt_recompose(_ptr(c), c.shape[2], c.shape[1], n_scales, int(include_largest),
                             _ptr(m), _ptr(out))
    return out


def ref_iuwt_decompose(image, n_scales, include_largest=True):
    a = _f32(image)
    out = np.zeros((n_scales + (1 if include_largest else 0),) + a.shape, np.float32)
    ref_lib().ref_iuwt_decompose(_ptr(a), a.shape[1], a.shape[0], n_scales, int(include_largest),
                                 _ptr(out))
    return out


def ref_iuwt_decompose_aliased(image, n_scales):
    """IuwtDecomposition::Decompose(x, x, false) of the reference's own source: input aliased to
    the scratch image, as IuwtDeconvolutionAlgorithm calls it."""
    a = _f32(image).copy()
    out = np.zeros((n_scales,) + a.shape, np.float32)
    f = ref_lib().ref_iuwt_decompose_aliased
    f.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_int32, C.c_void_p]
    f(_ptr(a), a.shape[1], a.shape[0], n_scales, _ptr(out))
    return out


def ref_iuwt_recompose(coefficients, n_scales, include_largest=True, masks=None):
    c = _f32(coefficients)
    if c.shape[0] == n_scales:  # the wrapper reads n_scales + 1 planes
        c = np.concatenate([c, np.zeros((1,) + c.shape[1:], np.float32)])
    out = np.empty(c.shape[1:], np.float32)
    m = None if masks is None else np.ascontiguousarray(masks, np.uint8)
    ref_lib().ref_iuwt_recompose(_ptr(c), c.shape[2], c.shape[1], n_scales, int(include_largest),
                                 _ptr(m), _ptr(out))
    return out


def ref_iuwt_smooth(image, to_scale):
    a = _f32(image).copy()
    ref_lib().ref_iuwt_smooth(_ptr(a), a.shape[1], a.shape[0], to_scale)
    return a


def _iuwt_analysis(prefix):
    """Structure selection / flood fills / plane measurements, from liboracle.so (prefix "orc_")
    or from the reference's own sources in oracle/_ref (prefix "ref_"). Masks are uint8 arrays
    [n_scales, H, W] (2-D for the 2-D fill); every call returns (new masks, area) or a scalar."""
    L = lib() if prefix == "orc_" else ref_lib()

    def select_structures(coefficients, masks, thresholds, min_scale, end_scale, clean_border=0.0,
                          prior_mask=None):
        c = _f32(coefficients)
        m = np.ascontiguousarray(masks, np.uint8).copy()
        t = _f32(thresholds)
        p = None if prior_mask is None else np.ascontiguousarray(prior_mask, np.uint8)
        area = getattr(L, prefix + "iuwt_select_structures")(
            _ptr(c), _ptr(m), c.shape[2], c.shape[1], c.shape[0], _ptr(t), min_scale, end_scale,
            clean_border, _ptr(p))
        return m, area

    def floodfill(coefficients, masks, thresholds, min_scale, end_scale, seed, clean_border=0.0,
                  prior_mask=None):
        c = _f32(coefficients)
        m = np.ascontiguousarray(masks, np.uint8).copy()
        t = _f32(thresholds)
        p = None if prior_mask is None else np.ascontiguousarray(prior_mask, np.uint8)
        x, y, scale = seed
        area = getattr(L, prefix + "iuwt_floodfill")(
            _ptr(c), _ptr(m), c.shape[2], c.shape[1], c.shape[0], _ptr(t), min_scale, end_scale,
            x, y, scale, clean_border, _ptr(p))
        return m, area

    def floodfill_2d(image, mask, threshold, seed, use_abs=False):
        a = _f32(image)
        m = np.ascontiguousarray(mask, np.uint8).copy()
        area = getattr(L, prefix + "iuwt_floodfill_2d")(_ptr(a), _ptr(m), a.shape[1], a.shape[0],
                                                        threshold, int(use_abs), seed[0], seed[1])
        return m, area

    def mad(image):
        a = _f32(image)
        if prefix == "orc_":
            return float(L.orc_iuwt_mad(_ptr(a), a.size))
        return float(L.ref_iuwt_mad(_ptr(a), a.shape[1], a.shape[0]))

    def max_abs(image, clean_border=0.0, allow_negative=True, mask=None):
        a = _f32(image)
        m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
        x, y = C.c_uint64(), C.c_uint64()
        v = getattr(L, prefix + "iuwt_max_abs")(_ptr(a), a.shape[1], a.shape[0], clean_border,
                                                int(allow_negative), _ptr(m), C.byref(x),
                                                C.byref(y))
        return float(v), x.value, y.value

    return select_structures, floodfill, floodfill_2d, mad, max_abs


def iuwt_analysis():
    """(select_structures, floodfill, floodfill_2d, mad, max_abs) of the oracle restatement
    (oracle/iuwt_analysis.cc)."""
    return _iuwt_analysis("orc_")


def ref_rms_sliding_minimum(image, window, maximum=False):
    """math::rms_image::SlidingMinimum / SlidingMaximum of the reference's own rms_image.cc."""
    a = _f32(image)
    out = np.empty_like(a)
    f = ref_lib().ref_rms_sliding_minimum
    f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_int]
    f(_ptr(a), _ptr(out), a.shape[1], a.shape[0], int(window), int(maximum))
    return out


def ref_rms_make_factor_image(image, strength):
    """math::rms_image::MakeRmsFactorImage -> (factor image, stddev)."""
    a = _f32(image).copy()
    f = ref_lib().ref_rms_make_factor_image
    f.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_double]
    f.restype = C.c_double
    stddev = f(_ptr(a), a.shape[1], a.shape[0], float(strength))
    return a, stddev


def ref_load_and_average(inputs, weights, n_pol, n_deconvolution):
    """ImageSet::LoadAndAverage(true) of the reference's own image_set.cc: `inputs` is
    [original channel][pol] planes (shape (n_original * n_pol, h, w)), `weights` one per plane;
    returns (n_deconvolution * n_pol, h, w)."""
    a = _f32(inputs)
    n_in, h, w = a.shape
    n_original = n_in // n_pol
    wts = np.ascontiguousarray(weights, np.float64)
    out = np.zeros((min(n_deconvolution, n_original) * n_pol, h, w), np.float32)
    in_ptrs = (C.c_void_p * n_in)(*[a[i].ctypes.data for i in range(n_in)])
    out_ptrs = (C.c_void_p * len(out))(*[out[i].ctypes.data for i in range(len(out))])
    f = ref_lib().refa_load_and_average
    f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64,
                  C.c_uint64, C.c_void_p]
    if f(in_ptrs, wts.ctypes.data, n_original, n_pol, n_deconvolution, w, h, out_ptrs) != 0:
        raise RuntimeError("refa_load_and_average failed")
    return out


def ref_load_and_average_psfs(psfs, weights, n_deconvolution):
    """ImageSet::LoadAndAveragePsfs of the reference's own image_set.cc: (n_original, h, w), one
    weight per original channel -> (n_deconvolution, h, w)."""
    a = _f32(psfs)
    n, h, w = a.shape
    wts = np.ascontiguousarray(weights, np.float64)
    out = np.zeros((min(n_deconvolution, n) if n_deconvolution else n, h, w), np.float32)
    in_ptrs = (C.c_void_p * n)(*[a[i].ctypes.data for i in range(n)])
    out_ptrs = (C.c_void_p * len(out))(*[out[i].ctypes.data for i in range(len(out))])
    f = ref_lib().refa_load_and_average_psfs
    f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint64,
                  C.c_void_p]
    if f(in_ptrs, wts.ctypes.data, n, n_deconvolution, w, h, out_ptrs) != 0:
        raise RuntimeError("refa_load_and_average_psfs failed")
    return out


def ref_interpolate_model(models, n_pol, fit_frequencies, fit_weights, n_terms,
                          original_frequencies):
    """ImageSet::InterpolateAndStoreModel of the reference's own image_set.cc through the
    polynomial SpectralFitter stand-in: `models` (n_deconvolution * n_pol, h, w) ->
    (n_original * n_pol, h, w)."""
    a = _f32(models)
    n_in, h, w = a.shape
    n_dec = n_in // n_pol
    ff = np.ascontiguousarray(fit_frequencies, np.float64)
    fw = np.ascontiguousarray(fit_weights, np.float32)
    of = np.ascontiguousarray(original_frequencies, np.float64)
    out = np.zeros((len(of) * n_pol, h, w), np.float32)
    in_ptrs = (C.c_void_p * n_in)(*[a[i].ctypes.data for i in range(n_in)])
    out_ptrs = (C.c_void_p * len(out))(*[out[i].ctypes.data for i in range(len(out))])
    f = ref_lib().refa_interpolate_model
    f.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_uint64,
                  C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p]
    if f(in_ptrs, n_dec, n_pol, ff.ctypes.data, fw.ctypes.data, n_terms, of.ctypes.data, len(of),
         w, h, out_ptrs) != 0:
        raise RuntimeError("refa_interpolate_model failed")
    return out


def ref_radler_perform(residuals, models, psfs, frequencies, weights, n_deconvolution, *,
                       auto_threshold_sigma, minor_loop_gain, major_loop_gain,
                       minor_iteration_count, fit_terms, auto_mask_sigma=0.0,
                       n_major_iterations=1, use_sub_minor_optimization=False,
                       multiscale=False, beam_size=0.0, pixel_scale=1.0, iuwt=False):
    """radler::Radler::Perform of the reference's own radler.cc (major iterations of
    GenericClean, one polarisation). `residuals` / `models`
    (n_original, h, w) are overwritten the way wsclean's load-and-store accessors are.
    -> (another_iteration_required, iteration_number), or a list of them, one per major
    iteration, when n_major_iterations > 1 (auto_mask_sigma > 0: two-phase auto-masking)."""
    for a in (residuals, models, psfs):
        assert a.dtype == np.float32 and a.flags.c_contiguous
    n, h, w = residuals.shape
    fr = np.ascontiguousarray(frequencies, np.float64)
    wt = np.ascontiguousarray(weights, np.float64)
    tables = [(C.c_void_p * n)(*[a[i].ctypes.data for i in range(n)])
              for a in (residuals, models, psfs)]
    another = (C.c_int32 * n_major_iterations)()
    iterations = (C.c_uint64 * n_major_iterations)()
    f = ref_lib().refa_radler_perform
    f.argtypes = [C.c_uint64] * 4 + [C.c_void_p] * 5 + [C.c_double] * 3 + [C.c_uint64] * 2 + \
        [C.c_double, C.c_uint64, C.c_int32, C.c_int32, C.c_double, C.c_double] + \
        [C.c_void_p] * 2
    if f(w, h, n, n_deconvolution, fr.ctypes.data, wt.ctypes.data, tables[0], tables[1],
         tables[2], float(auto_threshold_sigma), float(minor_loop_gain), float(major_loop_gain),
         int(minor_iteration_count), int(fit_terms), float(auto_mask_sigma),
         int(n_major_iterations), int(use_sub_minor_optimization),
         2 if iuwt else int(multiscale),
         float(beam_size), float(pixel_scale), another, iterations) != 0:
        err = ref_lib().refa_last_error
        err.restype = C.c_char_p
        err.argtypes = [C.c_void_p]
        raise RuntimeError("Radler::Perform: " + err(None).decode())
    per_major = [(bool(a), int(i)) for a, i in zip(another, iterations)]
    return per_major[0] if n_major_iterations == 1 else per_major


def ref_median_and_stddev_from_mad(image):
    """aocommon::Image::MedianAndStdDevFromMAD, the stand-in of oracle/ref_shim (aocommon is not
    in the reference tree: restated, unpinned) -> (median, stddev) as float32."""
    a = _f32(image).ravel()
    m, s = C.c_float(), C.c_float()
    f = ref_lib().refa_median_and_stddev_from_mad
    f.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p]
    f.restype = None
    f(_ptr(a), a.size, C.byref(m), C.byref(s))
    return np.float32(m.value), np.float32(s.value)


def ref_iuwt_analysis():
    """The same five functions from the reference's own image_analysis.cc /
    iuwt_deconvolution_algorithm.cc compiled in oracle/_ref."""
    return _iuwt_analysis("ref_")


def good_fft_size(n):
    return lib().orc_good_fft_size(n)


def convolution_size(scale, n, padding):
    return lib().orc_convolution_size(scale, n, padding)


def make_shape_function(scale, max_n, shape=0):
    n = lib().orc_make_shape_function(scale, max_n, shape, None)
    out = np.zeros((n, n), np.float32)
    lib().orc_make_shape_function(scale, max_n, shape, _ptr(out))
    return out


def ms_transform(images, scale, shape=0):
    images = _f32(images).copy()
    n, h, w = images.shape
    lib().orc_ms_transform(_ptr(images), n, w, h, scale, shape)
    return images


def add_shape_component(image, scale, x, y, gain, shape=0):
    assert image.dtype == np.float32 and image.flags.c_contiguous
    lib().orc_add_shape_component(_ptr(image), image.shape[1], image.shape[0], scale, x, y, gain,
                                  shape)


# ---- oracle/_ref: the reference's OWN algorithm sources (oracle/ref_algo.cc) ------------
class _RefConfig(C.Structure):
    _fields_ = [("algorithm", C.c_int32), ("use_sub_minor_optimization", C.c_int32),
                ("threshold", C.c_float), ("minor_loop_gain", C.c_float),
                ("major_loop_gain", C.c_float), ("clean_border_ratio", C.c_float),
                ("divergence_limit", C.c_float), ("max_iterations", C.c_uint64),
                ("allow_negative", C.c_int32), ("stop_on_negative", C.c_int32),
                ("ms_fast_sub_minor_loop", C.c_int32), ("ms_sub_minor_loop_gain", C.c_double),
                ("ms_scale_bias", C.c_double), ("ms_max_scales", C.c_uint64),
                ("ms_convolution_padding", C.c_double), ("ms_scale_list", C.POINTER(C.c_double)),
                ("ms_n_scale_list", C.c_uint64), ("ms_shape", C.c_int32),
                ("beam_size", C.c_double), ("pixel_scale", C.c_double),
                ("track_components", C.c_int32), ("fit_mode", C.c_int32),
                ("fit_terms", C.c_uint64), ("grid_width", C.c_uint64),
                ("grid_height", C.c_uint64), ("thread_count", C.c_uint64),
                ("squared_joins", C.c_int32)]


class _RefResult(C.Structure):
    _fields_ = [("has_start_peak", C.c_int32), ("start_peak", C.c_float), ("end_peak", C.c_float),
                ("another_iteration_required", C.c_int32), ("iteration_number", C.c_uint64)]


_refa_ready = False


_ADAPTER_PATH = os.path.join(_HERE, "_ref", "libradler_ref_adapter.so")
_adapter = None


def adapter_lib():
    """oracle/_ref/libradler_ref_adapter.so: the reference's own sources again, plus
    include/cuda_deconvolution.h (the adapter a Radler maintainer adds) so that the algorithm
    behind the reference's ParallelDeconvolution can be the CUDA library. Built by `make -C oracle
    adapter` where /root/reference exists; None when it was never built."""
    global _adapter
    if _adapter is None:
        if not os.path.exists(_ADAPTER_PATH):
            if not os.path.isdir(os.path.join(_REFERENCE_ROOT, "cpp")):
                return None
            subprocess.run(["make", "-C", _HERE, "-s", "adapter"], check=True)
        _adapter = C.CDLL(_ADAPTER_PATH)
        _bind_refa(_adapter)
    return _adapter


def _refa():
    global _refa_ready
    L = ref_lib()
    if L is None:
        return None
    if not _refa_ready:
        _bind_refa(L)
        _refa_ready = True
    return L


def _bind_refa(L):
    if True:
        vp, u64 = C.c_void_p, C.c_uint64
        L.refa_last_error.restype = C.c_char_p
        L.refa_last_error.argtypes = [vp]
        L.refa_create.restype = vp
        L.refa_create.argtypes = [C.POINTER(_RefConfig), u64, u64, u64, u64, vp, vp, vp]
        L.refa_destroy.argtypes = [vp]
        L.refa_set_threshold.argtypes = [vp, C.c_double]
        L.refa_set_minor_loop_gain.argtypes = [vp, C.c_double]
        L.refa_set_auto_mask_mode.argtypes = [vp, C.c_int32, C.c_int32]
        L.refa_set_clean_mask.argtypes = [vp, vp]
        L.refa_set_rms_factor_image.argtypes = [vp, vp]
        L.refa_set_iteration_number.argtypes = [vp, u64]
        L.refa_set_max_iterations.argtypes = [vp, u64]
        L.refa_set_major_iteration_threshold.argtypes = [vp, C.c_float]
        L.refa_execute.argtypes = [vp, vp, vp, vp, u64, u64, u64, vp, vp, C.c_double,
                                   C.POINTER(_RefResult)]
        L.refa_iteration_number.restype = u64
        L.refa_iteration_number.argtypes = [vp]
        L.refa_scale_count.restype = u64
        L.refa_scale_count.argtypes = [vp]
        L.refa_scale_size.restype = C.c_float
        L.refa_scale_size.argtypes = [vp, u64]
        L.refa_scale_mask_count.restype = u64
        L.refa_scale_mask_count.argtypes = [vp]
        L.refa_scale_mask.argtypes = [vp, u64, vp]
        L.refa_component_count.restype = u64
        L.refa_component_count.argtypes = [vp, u64]
        L.refa_component_scales.restype = u64
        L.refa_component_scales.argtypes = [vp]
        L.refa_component_get.argtypes = [vp, u64, u64, C.POINTER(u64), C.POINTER(u64), vp]
        L.refa_parallel_scale_mask_count.restype = u64
        L.refa_parallel_scale_mask_count.argtypes = [vp]
        L.refa_parallel_scale_mask.argtypes = [vp, u64, u64, u64, vp]


@dataclass
class RefResult:
    starting_peak_value: float | None
    final_peak_value: float
    another_iteration_required: bool
    iteration_number: int


# aocommon::PolarizationEnum order of oracle/ref_shim/aocommon/polarization.h
POL = {"I": 0, "Q": 1, "U": 2, "V": 3, "RR": 4, "RL": 5, "LR": 6, "LL": 7, "XX": 8, "XY": 9,
       "YX": 10, "YY": 11}


def ref_convolve_stats(reset=True):
    """(sizes -> count of schaapcommon::math::Convolve calls, seconds spent in them) since the last
    reset, for the reference's sources in oracle/_ref."""
    L = _refa()
    L.refa_convolve_stats.restype = C.c_uint64
    L.refa_convolve_stats.argtypes = [C.c_void_p, C.c_uint64, C.POINTER(C.c_double), C.c_int32]
    buf = np.zeros(3 * 64, np.uint64)
    sec = C.c_double()
    n = L.refa_convolve_stats(_ptr(buf), 64, C.byref(sec), int(reset))
    return {(int(buf[3 * i]), int(buf[3 * i + 1])): int(buf[3 * i + 2]) for i in range(min(n, 64))}, sec.value


class Reference:
    """The reference's own GenericClean / MultiScaleAlgorithm / IuwtDeconvolution behind its
    ParallelDeconvolution, driven like Radler::InitializeDeconvolutionAlgorithm + Perform do
    (oracle/ref_algo.cc). algorithm: "generic", "multiscale" or "iuwt"."""

    def __init__(self, algorithm, width, height, n_channels=1, pols=("I",), weights=None,
                 frequencies=None, *, use_sub_minor_optimization=True, threshold=0.0,
                 minor_loop_gain=0.1, major_loop_gain=1.0, clean_border_ratio=0.0,
                 divergence_limit=4.0, max_iterations=500, allow_negative=True,
                 stop_on_negative=False, fast_sub_minor_loop=True, sub_minor_loop_gain=0.2,
                 scale_bias=0.6, max_scales=0, convolution_padding=1.1, scale_list=None, shape=0,
                 beam_size=0.0, pixel_scale=1.0, track_components=False, fit_terms=0,
                 grid=(1, 1), threads=0, squared_joins=False, adapter=False):
        """adapter=True: the same driver with radler::algorithms::CudaDeconvolution
        (include/cuda_deconvolution.h) as the algorithm behind the reference's
        ParallelDeconvolution instead of the reference's CPU algorithm."""
        L = adapter_lib() if adapter else _refa()
        if L is None:
            raise RuntimeError("oracle/_ref is not built")
        self._L = L
        sl = np.ascontiguousarray(scale_list if scale_list else [], np.float64)
        cfg = _RefConfig(
            {"generic": 0, "multiscale": 1, "iuwt": 2}[algorithm] + (10 if adapter else 0),
            int(use_sub_minor_optimization),
            threshold, minor_loop_gain, major_loop_gain, clean_border_ratio, divergence_limit,
            max_iterations, int(allow_negative), int(stop_on_negative), int(fast_sub_minor_loop),
            sub_minor_loop_gain, scale_bias, max_scales, convolution_padding,
            sl.ctypes.data_as(C.POINTER(C.c_double)), len(sl), shape, beam_size, pixel_scale,
            int(track_components), 1 if fit_terms else 0, fit_terms, grid[0], grid[1],
            threads or (os.cpu_count() or 1), int(squared_joins))
        p = np.ascontiguousarray([POL[x] for x in pols], np.int32)
        w = None if weights is None else np.ascontiguousarray(weights, np.float64)
        f = None if frequencies is None else np.ascontiguousarray(frequencies, np.float64)
        self.width, self.height, self.n_channels, self.n_pol = width, height, n_channels, len(pols)
        self._major_loop_gain = major_loop_gain
        self._h = L.refa_create(C.byref(cfg), width, height, n_channels, len(pols), _ptr(p),
                                _ptr(w), _ptr(f))
        if not self._h:
            raise RuntimeError(self._L.refa_last_error(None).decode())

    def __del__(self):
        if getattr(self, "_h", None) and getattr(self, "_L", None) is not None:
            self._L.refa_destroy(self._h)
            self._h = None

    def _check(self, rc):
        if rc != 0:
            raise RuntimeError(self._L.refa_last_error(self._h).decode())

    def set_threshold(self, v): self._check(self._L.refa_set_threshold(self._h, v))
    def set_minor_loop_gain(self, v): self._check(self._L.refa_set_minor_loop_gain(self._h, v))
    def set_major_loop_gain(self, v): self._major_loop_gain = v
    def set_auto_mask_mode(self, track, use):
        self._check(self._L.refa_set_auto_mask_mode(self._h, int(track), int(use)))
    def set_iteration_number(self, n): self._check(self._L.refa_set_iteration_number(self._h, n))
    def set_max_iterations(self, n): self._check(self._L.refa_set_max_iterations(self._h, n))
    def set_major_iteration_threshold(self, v):
        self._check(self._L.refa_set_major_iteration_threshold(self._h, v))

    def set_clean_mask(self, mask):
        m = None if mask is None else np.ascontiguousarray(mask, np.uint8)
        self._check(self._L.refa_set_clean_mask(self._h, _ptr(m)))

    def set_rms_factor_image(self, image):
        a = None if image is None else _f32(image)
        self._check(self._L.refa_set_rms_factor_image(self._h, _ptr(a)))

    def execute(self, residual, model, psfs, psf_offsets=None) -> RefResult:
        """residual/model [n_images,H,W] updated in place; psfs [n_channels,h,w] or
        [n_dd,n_channels,h,w] with psf_offsets [(x,y)] for direction-dependent PSFs."""
        assert residual.dtype == np.float32 and residual.flags.c_contiguous
        assert model.dtype == np.float32 and model.flags.c_contiguous
        psfs = _f32(psfs)
        if psfs.ndim == 3:
            psfs = psfs[None]
        n_dd, _, ph, pw = psfs.shape
        ox = oy = None
        if psf_offsets is not None:
            ox = np.ascontiguousarray([o[0] for o in psf_offsets], np.int64)
            oy = np.ascontiguousarray([o[1] for o in psf_offsets], np.int64)
        out = _RefResult()
        self._check(self._L.refa_execute(self._h, _ptr(residual), _ptr(model), _ptr(psfs), n_dd, pw,
                                      ph, _ptr(ox), _ptr(oy), self._major_loop_gain,
                                      C.byref(out)))
        return RefResult(out.start_peak if out.has_start_peak else None, out.end_peak,
                         bool(out.another_iteration_required), out.iteration_number)

    @property
    def iteration_number(self): return self._L.refa_iteration_number(self._h)

    def scale_sizes(self):
        return [self._L.refa_scale_size(self._h, i) for i in range(self._L.refa_scale_count(self._h))]

    def scale_mask(self, i):
        m = np.zeros((self.height, self.width), np.uint8)
        self._check(self._L.refa_scale_mask(self._h, i, _ptr(m)))
        return m

    def component_list(self, n_images):
        """MultiScaleAlgorithm::GetComponentList() after MergeDuplicates: per scale
        (x[n], y[n], values[n, n_images])."""
        out = []
        for scale in range(self._L.refa_component_scales(self._h)):
            n = self._L.refa_component_count(self._h, scale)
            x, y = np.zeros(n, np.uint64), np.zeros(n, np.uint64)
            v = np.zeros((n, n_images), np.float32)
            cx, cy = C.c_uint64(), C.c_uint64()
            for i in range(n):
                self._check(self._L.refa_component_get(self._h, scale, i, C.byref(cx), C.byref(cy),
                                                    _ptr(v[i])))
                x[i], y[i] = cx.value, cy.value
            out.append((x, y, v))
        return out

    def parallel_scale_masks(self):
        """ParallelDeconvolution's frame-sized per-scale masks (facet runs)."""
        out = []
        for i in range(self._L.refa_parallel_scale_mask_count(self._h)):
            m = np.zeros((self.height, self.width), np.uint8)
            self._check(self._L.refa_parallel_scale_mask(self._h, i, self.width, self.height, _ptr(m)))
            out.append(m)
        return out
