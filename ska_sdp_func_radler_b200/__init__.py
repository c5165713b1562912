"""radler_b200 — host-side Python view of ``libradler_b200.so`` (include/radler_b200.h).

The product is the C++/CUDA library; this module is a thin ctypes layer that mirrors the
reference's algorithm interface (``DeconvolutionAlgorithm::ExecuteMajorIteration`` and
its setters, cpp/algorithms/deconvolution_algorithm.h:60-210) so tests and ``bench.py``
call the hot path exactly as a host program would: plain pointers and sizes, no torch
types in any signature. torch is used only to own device memory when the caller wants
HBM-resident cubes (``RDL_MEM_DEVICE``).

There is NO CPU fallback and nothing here imports ``oracle/``: if the CUDA library is
missing, or no sm_100 device is present, construction fails loudly.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from dataclasses import dataclass, field

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libradler_b200.so")

RDL_OK, RDL_ERR_INVALID_ARGUMENT, RDL_ERR_NO_DEVICE, RDL_ERR_CUDA, RDL_ERR_UNSUPPORTED, \
    RDL_ERR_OUT_OF_MEMORY = range(6)
GENERIC_CLEAN, MULTISCALE, IUWT = 0, 1, 2
SHAPE_TAPERED_QUADRATIC, SHAPE_GAUSSIAN = 0, 1
FIT_NONE, FIT_POLYNOMIAL = 0, 1
MEM_HOST, MEM_DEVICE = 0, 1


class RadlerError(RuntimeError):
    """std::runtime_error of the reference, carried over the C ABI as status + message."""

    def __init__(self, status: int, message: str):
        super().__init__(f"[rdl status {status}] {message}")
        self.status = status


class _Config(C.Structure):
    _fields_ = [("algorithm", C.c_int32), ("device", C.c_int32),
                ("use_sub_minor_optimization", C.c_int32), ("ms_fast_sub_minor_loop", C.c_int32),
                ("ms_sub_minor_loop_gain", C.c_double), ("ms_scale_bias", C.c_double),
                ("ms_max_scales", C.c_uint64), ("ms_convolution_padding", C.c_double),
                ("ms_scale_list", C.POINTER(C.c_double)), ("ms_n_scale_list", C.c_uint64),
                ("ms_shape", C.c_int32), ("beam_size", C.c_double),
                ("pixel_scale_x", C.c_double), ("pixel_scale_y", C.c_double),
                ("track_components", C.c_int32)]


class _CubeDesc(C.Structure):
    _fields_ = [("width", C.c_uint64), ("height", C.c_uint64), ("n_channels", C.c_uint64),
                ("n_pol", C.c_uint64), ("channel_weights", C.POINTER(C.c_float)),
                ("pol_linked", C.POINTER(C.c_uint8)), ("pol_norm", C.c_float),
                ("squared_joins", C.c_int32), ("memory_space", C.c_int32)]


class _Result(C.Structure):
    _fields_ = [("has_starting_peak", C.c_int32), ("starting_peak_value", C.c_float),
                ("final_peak_value", C.c_float), ("another_iteration_required", C.c_int32),
                ("is_diverging", C.c_int32), ("iteration_number", C.c_uint64)]


class _ScaleInfo(C.Structure):
    _fields_ = [("scale", C.c_float), ("psf_peak", C.c_float), ("kernel_peak", C.c_float),
                ("bias_factor", C.c_float), ("gain", C.c_float),
                ("max_normalized_image_value", C.c_float),
                ("max_unnormalized_image_value", C.c_float), ("rms", C.c_float),
                ("max_x", C.c_uint64), ("max_y", C.c_uint64), ("is_active", C.c_int32),
                ("n_components_cleaned", C.c_uint64), ("total_flux_cleaned", C.c_float)]


class _FacetInfo(C.Structure):
    _fields_ = [("x", C.c_uint64), ("y", C.c_uint64), ("width", C.c_uint64),
                ("height", C.c_uint64), ("peak", C.c_double),
                ("reached_major_threshold", C.c_int32), ("iteration_number", C.c_uint64)]


TRACE_DTYPE = np.dtype([("iteration", "<u4"), ("scale", "<u4"), ("x", "<u4"), ("y", "<u4"),
                        ("value", "<f4")])

# every symbol include/radler_b200.h declares (checked by tests/test_abi.py)
EXPORTED_SYMBOLS = [
    "rdl_config_default", "rdl_create", "rdl_clone", "rdl_destroy", "rdl_last_error",
    "rdl_set_max_iterations", "rdl_set_threshold", "rdl_set_major_iteration_threshold",
    "rdl_set_minor_loop_gain", "rdl_set_major_loop_gain", "rdl_set_allow_negative_components",
    "rdl_set_stop_on_negative_components", "rdl_set_clean_border_ratio",
    "rdl_set_divergence_limit", "rdl_set_iteration_number", "rdl_set_clean_mask",
    "rdl_set_rms_factor_image", "rdl_set_spectral_fitter", "rdl_set_auto_mask_mode",
    "rdl_set_psfs_constant", "rdl_set_pin_host_buffers",
    "rdl_set_trace_capacity", "rdl_iteration_number", "rdl_scale_count", "rdl_scale_info_get",
    "rdl_scale_mask_get", "rdl_trace_size", "rdl_trace_get", "rdl_last_selected_count",
    "rdl_minor_iteration_count", "rdl_kernel_launch_count", "rdl_reference_convolutions", "rdl_execute_major_iteration",
    "rdl_find_peak", "rdl_subtract_image", "rdl_integrate", "rdl_convolve",
    "rdl_padded_convolution", "rdl_multiscale_transform", "rdl_iuwt_smooth", "rdl_iuwt_decompose", "rdl_iuwt_decompose_static", "rdl_iuwt_decompose_aliased", "rdl_iuwt_recompose",
    "rdl_iuwt_select_structures", "rdl_iuwt_floodfill", "rdl_iuwt_floodfill_2d", "rdl_iuwt_mad",
    "rdl_iuwt_max_abs", "rdl_rms_sliding_minimum", "rdl_rms_make_factor_image", "rdl_rms_make",
    "rdl_median_and_stddev_from_mad", "rdl_load_and_average", "rdl_load_and_average_psfs", "rdl_auto_mask_from_model",
    "rdl_interpolate_model",
    "rdl_parallel_execute_major_iteration", "rdl_parallel_execute_major_iteration_dd",
    "rdl_parallel_facet_count", "rdl_parallel_facet_get",
    "rdl_parallel_timing",
    "rdl_facet_dividing_path", "rdl_facet_flood_area", "rdl_facet_bounding_mask", "rdl_facet_grid",
    "rdl_dist_unique_id", "rdl_dist_init", "rdl_dist_set_row_band", "rdl_dist_set_row_blocks",
    "rdl_dist_finalize",
    "rdl_release_cached_memory", "rdl_component_scale_count", "rdl_component_count",
    "rdl_component_get", "rdl_component_list_clear",
    "rdl_good_fft_size", "rdl_convolution_size", "rdl_version",
    "rdl_device_count",
]

_lib = None


def build(force: bool = False) -> str:
    """Compiles libradler_b200.so for sm_100a (nvcc cross-compiles without a GPU)."""
    if force or not os.path.exists(LIB_PATH):
        subprocess.run(["make", "-C", os.path.join(_HERE, "csrc"), "-j8"] +
                       (["-B"] if force else []), check=True)
    return LIB_PATH


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(f"{LIB_PATH} is missing: build it with "
                          "`make -C ska_sdp_func_radler_b200/csrc` (there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    vp, u64, i32, f32 = C.c_void_p, C.c_uint64, C.c_int32, C.c_float
    L.rdl_config_default.argtypes = [C.POINTER(_Config)]
    L.rdl_create.argtypes = [C.POINTER(_Config), C.POINTER(vp)]
    L.rdl_clone.argtypes = [vp, C.POINTER(vp)]
    L.rdl_destroy.argtypes = [vp]
    L.rdl_destroy.restype = None
    L.rdl_last_error.argtypes = [vp]
    L.rdl_last_error.restype = C.c_char_p
    for name, t in [("rdl_set_max_iterations", u64), ("rdl_set_threshold", f32),
                    ("rdl_set_major_iteration_threshold", f32), ("rdl_set_minor_loop_gain", f32),
                    ("rdl_set_major_loop_gain", f32), ("rdl_set_allow_negative_components", i32),
                    ("rdl_set_stop_on_negative_components", i32),
                    ("rdl_set_clean_border_ratio", f32), ("rdl_set_divergence_limit", f32),
                    ("rdl_set_iteration_number", u64), ("rdl_set_trace_capacity", u64)]:
        getattr(L, name).argtypes = [vp, t]
    L.rdl_set_clean_mask.argtypes = [vp, vp, u64, u64]
    L.rdl_set_rms_factor_image.argtypes = [vp, vp, u64, u64]
    L.rdl_set_spectral_fitter.argtypes = [vp, i32, u64, vp, vp, u64, u64]
    L.rdl_set_auto_mask_mode.argtypes = [vp, i32, i32]
    for name in ["rdl_iteration_number", "rdl_scale_count", "rdl_trace_size",
                 "rdl_last_selected_count", "rdl_minor_iteration_count",
                 "rdl_kernel_launch_count"]:
        getattr(L, name).argtypes = [vp]
        getattr(L, name).restype = u64
    L.rdl_set_psfs_constant.argtypes = [vp, i32]
    L.rdl_set_pin_host_buffers.argtypes = [vp, i32]
    L.rdl_scale_info_get.argtypes = [vp, u64, C.POINTER(_ScaleInfo)]
    L.rdl_scale_mask_get.argtypes = [vp, u64, vp]
    L.rdl_trace_get.argtypes = [vp, vp, u64]
    L.rdl_trace_get.restype = u64
    L.rdl_execute_major_iteration.argtypes = [vp, C.POINTER(_CubeDesc), vp, vp, vp, vp,
                                              C.POINTER(_Result)]
    L.rdl_find_peak.argtypes = [vp, C.POINTER(_CubeDesc), vp, i32, i32, u64, u64, u64, u64, vp,
                                vp, vp, C.POINTER(i32), C.POINTER(u64), C.POINTER(u64),
                                C.POINTER(f32)]
    L.rdl_subtract_image.argtypes = [vp, C.POINTER(_CubeDesc), vp, vp, u64, u64, vp, vp]
    L.rdl_integrate.argtypes = [vp, C.POINTER(_CubeDesc), vp, i32, vp, vp]
    L.rdl_convolve.argtypes = [vp, vp, u64, vp, u64, u64, vp]
    L.rdl_multiscale_transform.argtypes = [vp, vp, u64, u64, u64, f32, i32, vp]
    L.rdl_padded_convolution.argtypes = [vp, vp, vp, vp, u64, u64, u64, u64, vp]
    L.rdl_good_fft_size.argtypes = [u64]
    L.rdl_good_fft_size.restype = u64
    L.rdl_parallel_execute_major_iteration.argtypes = [vp, u64, u64, C.POINTER(_CubeDesc), vp, vp,
                                                       vp, vp, C.POINTER(_Result)]
    L.rdl_parallel_facet_count.argtypes = [vp]
    L.rdl_parallel_facet_count.restype = u64
    L.rdl_parallel_facet_get.argtypes = [vp, u64, C.POINTER(_FacetInfo), vp]
    L.rdl_component_scale_count.argtypes = [vp]
    L.rdl_component_scale_count.restype = u64
    L.rdl_component_count.argtypes = [vp, u64, C.POINTER(u64)]
    L.rdl_component_get.argtypes = [vp, u64, u64, u64, vp, vp, vp]
    L.rdl_component_list_clear.argtypes = [vp]
    L.rdl_component_list_clear.restype = None
    L.rdl_parallel_timing.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_double)]
    L.rdl_facet_dividing_path.argtypes = [vp, vp, u64, u64, u64, u64, i32]
    L.rdl_facet_flood_area.argtypes = [vp, u64, u64, u64, i32, vp, C.POINTER(u64), C.POINTER(u64)]
    L.rdl_facet_bounding_mask.argtypes = [vp, u64, u64, vp, vp, u64, u64, vp]
    L.rdl_facet_grid.argtypes = [vp, u64, u64, vp, u64, u64, vp]
    L.rdl_dist_unique_id.argtypes = [vp]
    L.rdl_dist_init.argtypes = [vp, vp, i32, i32]
    L.rdl_dist_set_row_band.argtypes = [vp, u64, u64]
    L.rdl_dist_set_row_blocks.argtypes = [vp, u64, u64]
    L.rdl_dist_finalize.argtypes = [vp]
    L.rdl_iuwt_smooth.argtypes = [vp, vp, vp, u64, u64, C.c_int32, C.c_int32, vp]
    L.rdl_iuwt_decompose.argtypes = [vp, vp, u64, u64, C.c_int32, C.c_int32, vp, vp]
    L.rdl_iuwt_decompose_static.argtypes = [vp, vp, u64, u64, C.c_int32, C.c_int32, vp, vp]
    L.rdl_iuwt_recompose.argtypes = [vp, vp, u64, u64, C.c_int32, C.c_int32, vp, vp, vp]
    i32, f32, pu64 = C.c_int32, C.c_float, C.POINTER(u64)
    L.rdl_iuwt_select_structures.argtypes = [vp, vp, vp, u64, u64, i32, C.POINTER(f32), i32, i32,
                                             f32, vp, pu64, vp]
    L.rdl_iuwt_floodfill.argtypes = [vp, vp, vp, u64, u64, i32, C.POINTER(f32), i32, i32, u64, u64,
                                     i32, f32, vp, pu64, vp]
    L.rdl_iuwt_floodfill_2d.argtypes = [vp, vp, vp, u64, u64, f32, i32, u64, u64, pu64, vp]
    L.rdl_iuwt_mad.argtypes = [vp, vp, u64, C.POINTER(f32), vp]
    L.rdl_iuwt_max_abs.argtypes = [vp, vp, u64, u64, f32, i32, vp, C.POINTER(f32), pu64, pu64, vp]
    L.rdl_convolution_size.argtypes = [C.c_double, u64, C.c_double]
    L.rdl_convolution_size.restype = u64
    L.rdl_version.restype = C.c_char_p
    L.rdl_device_count.restype = i32
    _lib = L
    return L


@dataclass
class Cube:
    """ImageSet layout (cpp/image_set.h:19-381): images ordered [channel][polarization]."""
    width: int
    height: int
    n_channels: int = 1
    n_pol: int = 1
    weights: np.ndarray | None = None
    pol_linked: np.ndarray | None = None
    pol_norm: float = 1.0
    squared_joins: bool = False
    _keep: list = field(default_factory=list, repr=False)

    @property
    def n_images(self) -> int:
        return self.n_channels * self.n_pol

    def desc(self, memory_space: int) -> _CubeDesc:
        w = None if self.weights is None else np.ascontiguousarray(self.weights, np.float32)
        p = None if self.pol_linked is None else np.ascontiguousarray(self.pol_linked, np.uint8)
        self._keep = [w, p]
        return _CubeDesc(self.width, self.height, self.n_channels, self.n_pol,
                         None if w is None else w.ctypes.data_as(C.POINTER(C.c_float)),
                         None if p is None else p.ctypes.data_as(C.POINTER(C.c_uint8)),
                         self.pol_norm, int(self.squared_joins), memory_space)


@dataclass
class Result:
    """DeconvolutionResult (deconvolution_algorithm.h:31-58)."""
    starting_peak_value: float | None
    final_peak_value: float
    another_iteration_required: bool
    is_diverging: bool
    iteration_number: int


def _is_torch(a) -> bool:
    return type(a).__module__.startswith("torch")


def _plane_pointers(a, n: int, plane_elems: int):
    """Pointer table (void*[n]) of the n image planes of `a` + its memory space. `a` is one
    contiguous [n, H, W] array / tensor, or a list of n separately allocated planes (what an
    ImageSet of aocommon::Images is)."""
    if isinstance(a, (list, tuple)):
        if len(a) != n:
            raise ValueError(f"expected {n} planes")
        ptrs, spaces = [], set()
        for plane in a:
            t, space = _plane_pointers(plane, 1, plane_elems)
            ptrs.append(t[0])
            spaces.add(space)
        if len(spaces) != 1:
            raise ValueError("planes of one cube must live in one memory space")
        return (C.c_void_p * n)(*ptrs), spaces.pop()
    if _is_torch(a):
        if not a.is_cuda:
            raise ValueError("torch tensors passed to radler_b200 must live on a CUDA device")
        if not a.is_contiguous() or a.dtype.itemsize != 4 or not a.dtype.is_floating_point:
            raise ValueError("expected a contiguous float32 tensor")
        if a.numel() != n * plane_elems:
            raise ValueError(f"expected {n} planes of {plane_elems} pixels")
        base, space = a.data_ptr(), MEM_DEVICE
    else:
        if a.dtype != np.float32 or not a.flags.c_contiguous:
            raise ValueError("expected a C-contiguous float32 array")
        if a.size != n * plane_elems:
            raise ValueError(f"expected {n} planes of {plane_elems} pixels")
        base, space = a.ctypes.data, MEM_HOST
    table = (C.c_void_p * n)(*[base + i * plane_elems * 4 for i in range(n)])
    return table, space


class DeconvolutionAlgorithm:
    """Mirror of radler::algorithms::DeconvolutionAlgorithm over the C ABI."""

    def __init__(self, algorithm: int = GENERIC_CLEAN, *, device: int = 0,
                 use_sub_minor_optimization: bool = True, fast_sub_minor_loop: bool = True,
                 sub_minor_loop_gain: float = 0.2, scale_bias: float = 0.6, max_scales: int = 0,
                 convolution_padding: float = 1.1, scale_list=None, shape: int = 0,
                 beam_size: float = 0.0, pixel_scale: float = 1.0, track_components: bool = False,
                 _handle=None):
        L = lib()
        self._h = None
        if _handle is not None:
            self._h = _handle
            return
        cfg = _Config()
        L.rdl_config_default(C.byref(cfg))
        cfg.algorithm, cfg.device = algorithm, device
        cfg.use_sub_minor_optimization = int(use_sub_minor_optimization)
        cfg.ms_fast_sub_minor_loop = int(fast_sub_minor_loop)
        cfg.ms_sub_minor_loop_gain = sub_minor_loop_gain
        cfg.ms_scale_bias = scale_bias
        cfg.ms_max_scales = max_scales
        cfg.ms_convolution_padding = convolution_padding
        sl = None if not scale_list else np.ascontiguousarray(scale_list, np.float64)
        if sl is not None:
            cfg.ms_scale_list = sl.ctypes.data_as(C.POINTER(C.c_double))
            cfg.ms_n_scale_list = len(sl)
        cfg.ms_shape = shape
        cfg.beam_size = beam_size
        cfg.pixel_scale_x = cfg.pixel_scale_y = pixel_scale
        cfg.track_components = int(track_components)
        h = C.c_void_p()
        st = L.rdl_create(C.byref(cfg), C.byref(h))
        if st != RDL_OK:
            raise RadlerError(st, L.rdl_last_error(None).decode())
        self._h = h

    def __del__(self):
        # at interpreter shutdown module globals may already be gone
        h, L = getattr(self, "_h", None), _lib
        if h and L is not None:
            L.rdl_destroy(h)
            self._h = None

    def _check(self, st: int):
        if st != RDL_OK:
            raise RadlerError(st, lib().rdl_last_error(self._h).decode())

    def clone(self) -> "DeconvolutionAlgorithm":
        h = C.c_void_p()
        self._check(lib().rdl_clone(self._h, C.byref(h)))
        return DeconvolutionAlgorithm(_handle=h)

    # -- setters (deconvolution_algorithm.h:75-166) ---------------------------------
    def set_max_iterations(self, n): self._check(lib().rdl_set_max_iterations(self._h, n))
    def set_threshold(self, v): self._check(lib().rdl_set_threshold(self._h, v))
    def set_major_iteration_threshold(self, v):
        self._check(lib().rdl_set_major_iteration_threshold(self._h, v))
    def set_minor_loop_gain(self, v): self._check(lib().rdl_set_minor_loop_gain(self._h, v))
    def set_major_loop_gain(self, v): self._check(lib().rdl_set_major_loop_gain(self._h, v))
    def set_allow_negative_components(self, v):
        self._check(lib().rdl_set_allow_negative_components(self._h, int(v)))
    def set_stop_on_negative_components(self, v):
        self._check(lib().rdl_set_stop_on_negative_components(self._h, int(v)))
    def set_clean_border_ratio(self, v): self._check(lib().rdl_set_clean_border_ratio(self._h, v))
    def set_divergence_limit(self, v): self._check(lib().rdl_set_divergence_limit(self._h, v))
    def set_iteration_number(self, n): self._check(lib().rdl_set_iteration_number(self._h, n))
    def set_trace_capacity(self, n): self._check(lib().rdl_set_trace_capacity(self._h, n))

    def set_clean_mask(self, mask):
        if mask is None:
            self._check(lib().rdl_set_clean_mask(self._h, None, 0, 0))
            return
        if hasattr(mask, "data_ptr"):  # a uint8 device plane, e.g. from auto_mask_from_model
            self._check(lib().rdl_set_clean_mask(self._h, mask.data_ptr(), mask.shape[1],
                                                 mask.shape[0]))
            return
        m = np.ascontiguousarray(mask, np.uint8)
        self._check(lib().rdl_set_clean_mask(self._h, m.ctypes.data, m.shape[1], m.shape[0]))

    def set_rms_factor_image(self, image):
        if image is None:
            self._check(lib().rdl_set_rms_factor_image(self._h, None, 0, 0))
            return
        if hasattr(image, "data_ptr"):  # a float32 device plane, e.g. from rms_make_factor_image
            self._check(lib().rdl_set_rms_factor_image(self._h, image.data_ptr(), image.shape[1],
                                                       image.shape[0]))
            return
        a = np.ascontiguousarray(image, np.float32)
        self._check(lib().rdl_set_rms_factor_image(self._h, a.ctypes.data, a.shape[1], a.shape[0]))

    def set_spectral_fitter(self, mode, n_terms=0, frequencies=None, weights=None, n_pol=1):
        fr = None if frequencies is None else np.ascontiguousarray(frequencies, np.float64)
        w = None if weights is None else np.ascontiguousarray(weights, np.float32)
        self._check(lib().rdl_set_spectral_fitter(
            self._h, mode, n_terms, None if fr is None else fr.ctypes.data,
            None if w is None else w.ctypes.data, 0 if fr is None else len(fr), n_pol))

    def set_psfs_constant(self, constant: bool):
        """Host buffers: PSF planes passed again under the same pointers are not uploaded again."""
        self._check(lib().rdl_set_psfs_constant(self._h, int(constant)))

    def set_pin_host_buffers(self, enable: bool):
        """Host buffers: register the caller's planes with cudaHostRegister on first use."""
        self._check(lib().rdl_set_pin_host_buffers(self._h, int(enable)))

    def set_auto_mask_mode(self, track: bool, use: bool):
        self._check(lib().rdl_set_auto_mask_mode(self._h, int(track), int(use)))

    def apply(self, **settings):
        """Convenience: apply(threshold=..., minor_loop_gain=...) -> the setters above."""
        for k, v in settings.items():
            getattr(self, "set_" + k)(v)
        return self

    # -- getters ----------------------------------------------------------------------
    @property
    def iteration_number(self) -> int: return lib().rdl_iteration_number(self._h)
    @property
    def last_selected_count(self) -> int: return lib().rdl_last_selected_count(self._h)
    @property
    def minor_iteration_count(self) -> int: return lib().rdl_minor_iteration_count(self._h)
    @property
    def kernel_launch_count(self) -> int: return lib().rdl_kernel_launch_count(self._h)
    def reference_convolutions(self) -> dict:
        """{(width, height): count} of the Convolve calls the reference makes for the last call."""
        f = lib().rdl_reference_convolutions
        f.restype = C.c_uint64
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64]
        buf = np.zeros(3 * 64, np.uint64)
        n = f(self._h, buf.ctypes.data, 64)
        return {(int(buf[3 * i]), int(buf[3 * i + 1])): int(buf[3 * i + 2]) for i in range(min(n, 64))}

    @property
    def scale_count(self) -> int: return lib().rdl_scale_count(self._h)

    def scale_infos(self):
        out = []
        for i in range(self.scale_count):
            s = _ScaleInfo()
            self._check(lib().rdl_scale_info_get(self._h, i, C.byref(s)))
            out.append({k: getattr(s, k) for k, _ in _ScaleInfo._fields_})
        return out

    def component_list(self, n_images):
        """MultiScaleAlgorithm::GetComponentList() after MergeDuplicates: per scale a tuple
        (x[n], y[n], values[n, n_images])."""
        out = []
        for scale in range(lib().rdl_component_scale_count(self._h)):
            n = C.c_uint64()
            self._check(lib().rdl_component_count(self._h, scale, C.byref(n)))
            x = np.zeros(n.value, np.uint64)
            y = np.zeros(n.value, np.uint64)
            v = np.zeros((n.value, n_images), np.float32)
            if n.value:
                self._check(lib().rdl_component_get(self._h, scale, 0, n.value, x.ctypes.data,
                                                    y.ctypes.data, v.ctypes.data))
            out.append((x, y, v))
        return out

    def clear_component_list(self):
        lib().rdl_component_list_clear(self._h)

    def scale_mask(self, i, width, height):
        m = np.zeros((height, width), np.uint8)
        self._check(lib().rdl_scale_mask_get(self._h, i, m.ctypes.data))
        return m

    def trace(self) -> np.ndarray:
        n = lib().rdl_trace_size(self._h)
        t = np.zeros(n, TRACE_DTYPE)
        if n:
            lib().rdl_trace_get(self._h, t.ctypes.data, n)
        return t

    # -- the hot path -------------------------------------------------------------------
    def execute_major_iteration(self, cube: Cube, residual, model, psfs, stream=None) -> Result:
        """ExecuteMajorIteration (deconvolution_algorithm.h:69-71). residual/model
        ([n_images,H,W]) are updated in place; psfs is [n_channels,H,W]. All three are
        either numpy arrays (host memory, staged inside the call) or CUDA torch tensors."""
        n = cube.width * cube.height
        rt, s0 = _plane_pointers(residual, cube.n_images, n)
        mt, s1 = _plane_pointers(model, cube.n_images, n)
        # row-sharded cubes (dist_set_row_band) carry full-frame PSFs
        psf_n = cube.width * self._frame_height if getattr(self, "_frame_height", 0) else n
        pt, s2 = _plane_pointers(psfs, cube.n_channels, psf_n)
        if not (s0 == s1 == s2):
            raise ValueError("residual, model and psfs must live in the same memory space")
        d = cube.desc(s0)
        out = _Result()
        self._check(lib().rdl_execute_major_iteration(self._h, C.byref(d), rt, mt, pt, stream,
                                                      C.byref(out)))
        return Result(out.starting_peak_value if out.has_starting_peak else None,
                      out.final_peak_value, bool(out.another_iteration_required),
                      bool(out.is_diverging), out.iteration_number)

    # -- facets: ParallelDeconvolution (parallel_deconvolution.cc:494-654) ----------------
    def parallel_execute_major_iteration(self, grid_width: int, grid_height: int, cube: Cube,
                                         residual, model, psfs, stream=None,
                                         psf_offsets=None) -> Result:
        """psf_offsets [(x, y)]: direction-dependent PSFs, psfs = [n_dd, n_channels, ph, pw]."""
        n = cube.width * cube.height
        rt, s0 = _plane_pointers(residual, cube.n_images, n)
        mt, s1 = _plane_pointers(model, cube.n_images, n)
        out = _Result()
        if psf_offsets is not None:
            n_dd, n_c, ph, pw = psfs.shape
            assert n_c == cube.n_channels and n_dd == len(psf_offsets)
            flat = psfs.reshape(n_dd * n_c, ph, pw)
            pt, s2 = _plane_pointers(flat, n_dd * n_c, ph * pw)
            if not (s0 == s1 == s2):
                raise ValueError("residual, model and psfs must live in the same memory space")
            ox = np.ascontiguousarray([o[0] for o in psf_offsets], np.int64)
            oy = np.ascontiguousarray([o[1] for o in psf_offsets], np.int64)
            d = cube.desc(s0)
            f = lib().rdl_parallel_execute_major_iteration_dd
            f.argtypes = [C.c_void_p, C.c_uint64, C.c_uint64, C.POINTER(_CubeDesc), C.c_void_p,
                          C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64, C.c_void_p,
                          C.c_void_p, C.c_void_p, C.POINTER(_Result)]
            self._check(f(self._h, grid_width, grid_height, C.byref(d), rt, mt, pt, n_dd, pw, ph,
                          ox.ctypes.data, oy.ctypes.data, stream, C.byref(out)))
            return Result(out.starting_peak_value if out.has_starting_peak else None,
                          out.final_peak_value, bool(out.another_iteration_required),
                          bool(out.is_diverging), out.iteration_number)
        pt, s2 = _plane_pointers(psfs, cube.n_channels, n)
        if not (s0 == s1 == s2):
            raise ValueError("residual, model and psfs must live in the same memory space")
        d = cube.desc(s0)
        self._check(lib().rdl_parallel_execute_major_iteration(
            self._h, grid_width, grid_height, C.byref(d), rt, mt, pt, stream, C.byref(out)))
        return Result(out.starting_peak_value if out.has_starting_peak else None,
                      out.final_peak_value, bool(out.another_iteration_required),
                      bool(out.is_diverging), out.iteration_number)

    def parallel_timing(self):
        """(host divider search ms, whole call ms) of the last facet run."""
        g, t = C.c_double(), C.c_double()
        self._check(lib().rdl_parallel_timing(self._h, C.byref(g), C.byref(t)))
        return g.value, t.value

    def facets(self, with_masks: bool = False):
        """SubImage records of the last facet run."""
        out = []
        for i in range(lib().rdl_parallel_facet_count(self._h)):
            f = _FacetInfo()
            self._check(lib().rdl_parallel_facet_get(self._h, i, C.byref(f), None))
            rec = {k: getattr(f, k) for k, _ in _FacetInfo._fields_}
            if with_masks:
                m = np.zeros((f.height, f.width), np.uint8)
                self._check(lib().rdl_parallel_facet_get(self._h, i, C.byref(f), m.ctypes.data))
                rec["boundary_mask"] = m
            out.append(rec)
        return out

    # -- kernel-level entry points (device tensors) -----------------------------------
    def find_peak(self, cube: Cube, images, *, square=False, allow_negative=True, start_y=0,
                  end_y=None, h_border=0, v_border=0, mask=None, rms_factor=None, stream=None):
        t, space = _plane_pointers(images, cube.n_images, cube.width * cube.height)
        assert space == MEM_DEVICE
        d = cube.desc(space)
        found, x, y, v = C.c_int32(), C.c_uint64(), C.c_uint64(), C.c_float()
        self._check(lib().rdl_find_peak(
            self._h, C.byref(d), t, int(square), int(allow_negative), start_y,
            cube.height if end_y is None else end_y, h_border, v_border,
            None if mask is None else mask.data_ptr(),
            None if rms_factor is None else rms_factor.data_ptr(), stream, C.byref(found),
            C.byref(x), C.byref(y), C.byref(v)))
        return (v.value if found.value else None), x.value, y.value

    def subtract_image(self, cube: Cube, images, psfs, x, y, factors, stream=None):
        t, space = _plane_pointers(images, cube.n_images, cube.width * cube.height)
        p, _ = _plane_pointers(psfs, cube.n_channels, cube.width * cube.height)
        assert space == MEM_DEVICE
        d = cube.desc(space)
        f = np.ascontiguousarray(factors, np.float32)
        assert f.size == cube.n_images
        self._check(lib().rdl_subtract_image(self._h, C.byref(d), t, p, x, y, f.ctypes.data,
                                             stream))

    def integrate(self, cube: Cube, images, dest, *, square=False, stream=None):
        t, space = _plane_pointers(images, cube.n_images, cube.width * cube.height)
        assert space == MEM_DEVICE
        d = cube.desc(space)
        self._check(lib().rdl_integrate(self._h, C.byref(d), t, int(square), dest.data_ptr(),
                                        stream))

    def convolve(self, images, kernel, stream=None):
        """schaapcommon::math::Convolve on [n,H,W] device images with a prepared kernel."""
        n, h, w = images.shape
        t, space = _plane_pointers(images, n, h * w)
        assert space == MEM_DEVICE
        self._check(lib().rdl_convolve(self._h, t, n, kernel.data_ptr(), w, h, stream))

    def padded_psf_convolve(self, image, psf, padded_width, padded_height, stream=None):
        """Trim(Convolve(Untrim(image), PrepareConvolutionKernel(Untrim(psf)))) on [H,W] device
        images (subminor_loop.cc:195-218); returns a new device image."""
        h, w = image.shape
        out = image.clone()
        self._check(lib().rdl_padded_convolution(self._h, image.data_ptr(), psf.data_ptr(),
                                                 out.data_ptr(), w, h, padded_width, padded_height,
                                                 stream))
        return out

    def multiscale_transform(self, images, scale, shape=0, stream=None):
        n, h, w = images.shape
        t, space = _plane_pointers(images, n, h * w)
        assert space == MEM_DEVICE
        self._check(lib().rdl_multiscale_transform(self._h, t, n, w, h, scale, shape, stream))


    # ---- multi-GPU: rows of the cube sharded over one process per GPU ---------------------------
    def dist_init(self, unique_id: bytes, rank: int, world: int):
        buf = (C.c_uint8 * 128).from_buffer_copy(unique_id)
        self._check(lib().rdl_dist_init(self._h, buf, rank, world))

    def dist_set_row_band(self, y_offset: int, global_height: int):
        self._check(lib().rdl_dist_set_row_band(self._h, y_offset, global_height))
        self._frame_height = global_height

    def dist_set_row_blocks(self, block_rows: int, global_height: int):
        self._check(lib().rdl_dist_set_row_blocks(self._h, block_rows, global_height))
        self._frame_height = global_height

    def dist_finalize(self):
        self._check(lib().rdl_dist_finalize(self._h))

    # ---- IUWT decomposition (iuwt_decomposition.{h,cc}); torch device tensors ------------------
    def iuwt_smooth(self, image, out, scale, simple=False, stream=None):
        h, w = image.shape
        self._check(lib().rdl_iuwt_smooth(self._h, image.data_ptr(), out.data_ptr(), w, h, scale,
                                          int(simple), stream))

    def iuwt_decompose(self, image, coefficients, n_scales, include_largest=True, stream=None):
        """coefficients: [n_scales (+1), H, W] device tensor, filled in place."""
        h, w = image.shape
        assert coefficients.shape[0] >= n_scales + (1 if include_largest else 0)
        self._check(lib().rdl_iuwt_decompose(self._h, image.data_ptr(), w, h, n_scales,
                                             int(include_largest), coefficients.data_ptr(), stream))

    def iuwt_decompose_static(self, image, coefficients, n_scales, include_largest=True,
                              stream=None):
        """iuwt_decompose without device allocations: the scratch planes are a static device
        array of the library (width * height <= 2**20; calls must not overlap on the GPU)."""
        h, w = image.shape
        assert coefficients.shape[0] >= n_scales + (1 if include_largest else 0)
        self._check(lib().rdl_iuwt_decompose_static(self._h, image.data_ptr(), w, h, n_scales,
                                                    int(include_largest), coefficients.data_ptr(),
                                                    stream))

    def iuwt_decompose_aliased(self, x, coefficients, n_scales, stream=None):
        """IuwtDecomposition::Decompose(x, x, false) (input aliased to the scratch image)."""
        h, w = x.shape
        f = lib().rdl_iuwt_decompose_aliased
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_int32, C.c_void_p,
                      C.c_void_p]
        self._check(f(self._h, x.data_ptr(), w, h, n_scales, coefficients.data_ptr(), stream))

    def iuwt_recompose(self, coefficients, out, n_scales, include_largest=True, masks=None,
                       stream=None):
        h, w = out.shape
        self._check(lib().rdl_iuwt_recompose(self._h, coefficients.data_ptr(), w, h, n_scales,
                                             int(include_largest),
                                             None if masks is None else masks.data_ptr(),
                                             out.data_ptr(), stream))

    # IUWT structure selection (cpp/algorithms/iuwt/image_analysis.cc) and plane measurements
    # (iuwt_deconvolution_algorithm.cc:104-167). coefficients: float32 CUDA tensor [S, H, W];
    # masks: uint8 CUDA tensor [S, H, W], updated in place; thresholds: S host floats.
    def iuwt_select_structures(self, coefficients, masks, thresholds, min_scale, end_scale,
                               clean_border=0.0, prior_mask=None, stream=None):
        """image_analysis::SelectStructures (image_analysis.cc:208-239) -> voxels added."""
        s, h, w = coefficients.shape
        thr = (C.c_float * s)(*[float(t) for t in thresholds])
        area = C.c_uint64()
        self._check(lib().rdl_iuwt_select_structures(
            self._h, coefficients.data_ptr(), masks.data_ptr(), w, h, s, thr, min_scale, end_scale,
            clean_border, None if prior_mask is None else prior_mask.data_ptr(), C.byref(area),
            stream))
        return area.value

    def iuwt_floodfill(self, coefficients, masks, thresholds, min_scale, end_scale, seed,
                       clean_border=0.0, prior_mask=None, stream=None):
        """image_analysis::Floodfill / MaskedFloodfill (image_analysis.cc:86-206); seed =
        (x, y, scale) -> voxels filled."""
        s, h, w = coefficients.shape
        thr = (C.c_float * s)(*[float(t) for t in thresholds])
        area = C.c_uint64()
        self._check(lib().rdl_iuwt_floodfill(
            self._h, coefficients.data_ptr(), masks.data_ptr(), w, h, s, thr, min_scale, end_scale,
            seed[0], seed[1], seed[2], clean_border,
            None if prior_mask is None else prior_mask.data_ptr(), C.byref(area), stream))
        return area.value

    def iuwt_floodfill_2d(self, image, mask, threshold, seed, use_abs=False, stream=None):
        """image_analysis::FloodFill2D (image_analysis.cc:241-333) -> pixels filled."""
        h, w = image.shape
        area = C.c_uint64()
        self._check(lib().rdl_iuwt_floodfill_2d(self._h, image.data_ptr(), mask.data_ptr(), w, h,
                                                threshold, int(use_abs), seed[0], seed[1],
                                                C.byref(area), stream))
        return area.value

    # -- local-RMS maps (cpp/math/rms_image.cc) -----------------------------------------
    def rms_sliding_minimum(self, image, out, window_size, maximum=False, stream=None):
        h, w = image.shape
        f = lib().rdl_rms_sliding_minimum
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.c_uint64,
                      C.c_int32, C.c_void_p]
        self._check(f(self._h, image.data_ptr(), out.data_ptr(), w, h, int(window_size),
                      int(maximum), stream))

    def rms_make_factor_image(self, image, local_rms_strength=1.0, stream=None) -> float:
        """In place; returns the lowest RMS (MakeRmsFactorImage's return value)."""
        f = lib().rdl_rms_make_factor_image
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_double, C.POINTER(C.c_double),
                      C.c_void_p]
        s = C.c_double()
        self._check(f(self._h, image.data_ptr(), image.numel(), float(local_rms_strength),
                      C.byref(s), stream))
        return s.value

    def rms_make(self, image, out, window_size, beam_major, beam_minor, beam_pa, pixel_scale_l,
                 pixel_scale_m, with_negativity_limit=False, stream=None):
        h, w = image.shape
        f = lib().rdl_rms_make
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64] + \
            [C.c_double] * 6 + [C.c_int32, C.c_void_p]
        self._check(f(self._h, image.data_ptr(), out.data_ptr(), w, h, float(window_size),
                      float(beam_major), float(beam_minor), float(beam_pa), float(pixel_scale_l),
                      float(pixel_scale_m), int(with_negativity_limit), stream))

    def median_and_stddev_from_mad(self, image, stream=None):
        """integrated.MedianAndStdDevFromMAD() of Radler::Perform (radler.cc:162-169) on a device
        plane -> (median, stddev, n_finite)."""
        f = lib().rdl_median_and_stddev_from_mad
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p,
                      C.c_void_p]
        m, s, n = C.c_float(), C.c_float(), C.c_uint64()
        self._check(f(self._h, image.data_ptr() if image.numel() else None, image.numel(),
                      C.byref(m), C.byref(s), C.byref(n), stream))
        return m.value, s.value, n.value

    def load_and_average(self, inputs, weights, output_index, outputs, stream=None, psfs=False):
        """ImageSet::LoadAndAverage (image_set.cc:105-140) with the planes already on the device:
        `inputs` / `outputs` are sequences of equally sized device planes. psfs=True:
        ImageSet::LoadAndAveragePsfs (:142-205), whose products stay double."""
        n_in, n_out = len(inputs), len(outputs)
        in_ptrs = (C.c_void_p * max(n_in, 1))(*[t.data_ptr() for t in inputs])
        out_ptrs = (C.c_void_p * max(n_out, 1))(*[t.data_ptr() for t in outputs])
        w = (C.c_double * max(n_in, 1))(*[float(v) for v in weights])
        idx = (C.c_uint32 * max(n_in, 1))(*[int(v) for v in output_index])
        f = lib().rdl_load_and_average_psfs if psfs else lib().rdl_load_and_average
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p,
                      C.c_uint64, C.c_uint64, C.c_void_p]
        n_pixels = outputs[0].numel() if n_out else 0
        self._check(f(self._h, in_ptrs, w, idx, n_in, out_ptrs, n_out, n_pixels, stream))

    def auto_mask_from_model(self, models, mask, stream=None):
        """The auto-mask Radler::Perform derives from the model planes (radler.cc:254-266):
        `mask` (uint8 device plane) = 1 where any of `models` is finite and non-zero."""
        n = len(models)
        ptrs = (C.c_void_p * max(n, 1))(*[t.data_ptr() for t in models])
        f = lib().rdl_auto_mask_from_model
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p]
        self._check(f(self._h, ptrs, n, mask.data_ptr(), mask.numel(), stream))

    def interpolate_model(self, models, output_frequencies, outputs, stream=None):
        """ImageSet::InterpolateAndStoreModel (image_set.cc:208-283) for one polarisation: the
        polynomial fitter of set_spectral_fitter fitted per pixel to `models` (device planes, one
        per deconvolution channel) and evaluated at `output_frequencies` into `outputs`."""
        n, n_out = len(models), len(outputs)
        m_ptrs = (C.c_void_p * max(n, 1))(*[t.data_ptr() for t in models])
        o_ptrs = (C.c_void_p * max(n_out, 1))(*[t.data_ptr() for t in outputs])
        freqs = (C.c_double * max(n_out, 1))(*[float(f) for f in output_frequencies])
        f = lib().rdl_interpolate_model
        f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_void_p,
                      C.c_uint64, C.c_void_p]
        self._check(f(self._h, m_ptrs, n, freqs, n_out, o_ptrs,
                      models[0].numel() if n else 0, stream))

    def iuwt_mad(self, image, stream=None):
        """IuwtDeconvolutionAlgorithm::Mad (iuwt_deconvolution_algorithm.cc:104-110)."""
        v = C.c_float()
        self._check(lib().rdl_iuwt_mad(self._h, image.data_ptr(), image.numel(), C.byref(v),
                                       stream))
        return v.value

    def iuwt_max_abs(self, image, clean_border=0.0, allow_negative=True, mask=None, stream=None):
        """GetMaxAbsWithoutMask / WithMask (iuwt_deconvolution_algorithm.cc:112-167) ->
        (value, x, y)."""
        h, w = image.shape
        v, x, y = C.c_float(), C.c_uint64(), C.c_uint64()
        self._check(lib().rdl_iuwt_max_abs(self._h, image.data_ptr(), w, h, clean_border,
                                           int(allow_negative),
                                           None if mask is None else mask.data_ptr(), C.byref(v),
                                           C.byref(x), C.byref(y), stream))
        return v.value, x.value, y.value


def GenericClean(use_sub_minor_optimization: bool = True, **kw) -> DeconvolutionAlgorithm:
    """GenericClean (cpp/algorithms/generic_clean.h)."""
    return DeconvolutionAlgorithm(GENERIC_CLEAN,
                                  use_sub_minor_optimization=use_sub_minor_optimization, **kw)


def MultiScaleAlgorithm(beam_size: float, pixel_scale: float = 1.0, **kw) -> DeconvolutionAlgorithm:
    """MultiScaleAlgorithm (cpp/algorithms/multiscale_algorithm.h)."""
    return DeconvolutionAlgorithm(MULTISCALE, beam_size=beam_size, pixel_scale=pixel_scale, **kw)


# ---- facet geometry on the host (math/dijkstra_splitter.h; no GPU needed) ---------------------
def _host_status(st):
    if st != RDL_OK:
        raise RadlerError(st, lib().rdl_last_error(None).decode())


def facet_dividing_path(image, lo, hi, vertical=True):
    img = np.ascontiguousarray(image, np.float32)
    out = np.zeros_like(img)
    _host_status(lib().rdl_facet_dividing_path(img.ctypes.data, out.ctypes.data, img.shape[1],
                                               img.shape[0], lo, hi, int(vertical)))
    return out


def facet_flood_area(dividers, inside, vertical=True):
    d = np.ascontiguousarray(dividers, np.float32)
    mask = np.zeros(d.shape, np.uint8)
    start, extent = C.c_uint64(), C.c_uint64()
    _host_status(lib().rdl_facet_flood_area(d.ctypes.data, d.shape[1], d.shape[0], inside,
                                            int(vertical), mask.ctypes.data, C.byref(start),
                                            C.byref(extent)))
    return mask, start.value, extent.value


def facet_bounding_mask(vertical_mask, vertical_mask_x, horizontal_mask, mask=None):
    v = np.ascontiguousarray(vertical_mask, np.uint8)
    hm = np.ascontiguousarray(horizontal_mask, np.uint8)
    mask = np.zeros(hm.shape, np.uint8) if mask is None else mask
    box = (C.c_uint64 * 4)()
    _host_status(lib().rdl_facet_bounding_mask(v.ctypes.data, vertical_mask_x, v.shape[1],
                                               hm.ctypes.data, mask.ctypes.data, hm.shape[1],
                                               hm.shape[0], box))
    return mask, tuple(box)


def facet_grid(image, grid_width, grid_height, user_mask=None):
    img = np.ascontiguousarray(image, np.float32)
    m = None if user_mask is None else np.ascontiguousarray(user_mask, np.uint8)
    out = (_FacetInfo * (grid_width * grid_height))()
    _host_status(lib().rdl_facet_grid(img.ctypes.data, img.shape[1], img.shape[0],
                                      None if m is None else m.ctypes.data, grid_width,
                                      grid_height, out))
    return [(f.x, f.y, f.width, f.height) for f in out]


def dist_unique_id() -> bytes:
    """ncclGetUniqueId (128 bytes) for rdl_dist_init; create on rank 0 and broadcast."""
    buf = (C.c_uint8 * 128)()
    st = lib().rdl_dist_unique_id(buf)
    if st != RDL_OK:
        raise RadlerError(st, "ncclGetUniqueId failed (is libnccl.so.2 loadable?)")
    return bytes(buf)


def good_fft_size(n: int) -> int:
    return lib().rdl_good_fft_size(n)


def convolution_size(scale: float, n: int, padding: float) -> int:
    return lib().rdl_convolution_size(scale, n, padding)


def version() -> str:
    return lib().rdl_version().decode()


def device_count() -> int:
    return lib().rdl_device_count()
