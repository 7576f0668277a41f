"""ctypes binding of the C ABI in include/mof.h (mrs_optic_flow_b200/libmof_b200.so).

The shared library is built in-tree by `make -C mrs_optic_flow_b200/csrc` (see __graft_entry__.build).
There is no fallback: if the library is missing, or no CUDA device is usable, the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# MOF_B200_LIB: an alternative build of the same library (A/B kernel experiments, tools/)
LIB_PATH = os.environ.get("MOF_B200_LIB") or os.path.join(_HERE, "libmof_b200.so")

MOF_OK, MOF_ERR_INVALID, MOF_ERR_CUDA, MOF_ERR_BUSY, MOF_ERR_NOMEM = 0, 1, 2, 3, 4

# every symbol include/mof.h declares (tests check that the library exports all of them)
EXPORTS = [
    "mof_default_config", "mof_create", "mof_destroy", "mof_last_error", "mof_get_geometry",
    "mof_fft_pairs", "mof_fft_pairs_device", "mof_fft_stream_push", "mof_fft_stream_push_device",
    "mof_fft_stream_set_prev", "mof_fft_stream_reset", "mof_bm_geometry", "mof_bm_pairs",
    "mof_bm_pairs_device", "mof_scalerot_pairs", "mof_scalerot_pairs_device", "mof_scalerot_push", "mof_scalerot_reset", "mof_logpolar",
    "mof_sync", "mof_kernel_launches", "mof_fft_kernel_time", "mof_enable_kernel_timing",
    "mof_version", "mof_fft_pairs_noalloc_device",
    "mof_frontend", "mof_frontend_device", "mof_fft_stream_push_raw", "mof_fft_stream_push_raw_device",
    "mof_longrange_geometry", "mof_fft_pairs_longrange", "mof_fft_pairs_longrange_device",
    "mof_point_pairs", "mof_point_pairs_device", "mof_bm_refine", "mof_bm_refine_device",
    "mof_fft_stream_push_many", "mof_fft_stream_push_many_device", "mof_fft_stream_push_many_raw",
    "mof_fft_stream_push_many_raw_device",
    "mof_multi_create", "mof_multi_destroy", "mof_multi_device_count", "mof_multi_last_error",
    "mof_multi_get_geometry", "mof_multi_fft_pairs", "mof_multi_fft_stream_push",
    "mof_multi_fft_stream_push_many", "mof_multi_fft_stream_push_many_raw", "mof_multi_kernel_launches",
]


class MofConfig(C.Structure):
    _fields_ = [
        ("device", C.c_int), ("frame_size", C.c_int), ("sample_point_size", C.c_int),
        ("max_px_speed", C.c_double), ("hann", C.c_int), ("max_batch", C.c_int),
        ("n_streams", C.c_int), ("bm_block", C.c_int), ("bm_step", C.c_int),
        ("bm_radius", C.c_int), ("sr_resolution", C.c_int), ("sr_m", C.c_double),
        ("bm_no_gate", C.c_int), ("ocl_compat", C.c_int), ("reserved", C.c_int * 6),
    ]


class MofFrontendDesc(C.Structure):
    _fields_ = [
        ("width", C.c_int), ("height", C.c_int), ("channels", C.c_int), ("row_stride", C.c_int),
        ("image_stride", C.c_size_t), ("scale_factor", C.c_int), ("cx", C.c_int),
    ]


class MofError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"mof error {code}: {msg}")
        self.code = code


_lib = None


def load() -> C.CDLL:
    """Load libmof_b200.so; raises (loudly) when the extension has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build the CUDA backend first "
            "(python -c 'import __graft_entry__ as g; g.build()' or make -C mrs_optic_flow_b200/csrc). "
            "There is no CPU fallback.")
    L = C.CDLL(LIB_PATH)
    vp, u8p, f64p = C.c_void_p, C.c_void_p, C.c_void_p
    L.mof_default_config.argtypes = [C.POINTER(MofConfig)]
    L.mof_default_config.restype = None
    L.mof_create.argtypes = [C.POINTER(vp), C.POINTER(MofConfig)]
    L.mof_destroy.argtypes = [vp]
    L.mof_destroy.restype = None
    L.mof_last_error.argtypes = [vp]
    L.mof_last_error.restype = C.c_char_p
    L.mof_get_geometry.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.mof_fft_pairs.argtypes = [vp, u8p, u8p, C.c_size_t, C.c_int, C.c_int, f64p]
    L.mof_fft_pairs_device.argtypes = [vp, u8p, u8p, C.c_size_t, C.c_int, C.c_int, f64p, vp]
    L.mof_fft_pairs_noalloc_device.argtypes = [C.c_int, u8p, u8p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_double,
                                               C.c_int, f64p, vp, C.c_char_p, C.c_size_t]
    L.mof_fft_pairs_noalloc_device.restype = C.c_int
    L.mof_fft_stream_push.argtypes = [vp, C.c_int, C.POINTER(C.c_int), u8p, C.c_size_t, C.c_int, f64p]
    L.mof_fft_stream_push_device.argtypes = [vp, C.c_int, C.POINTER(C.c_int), u8p, C.c_size_t,
                                             C.c_int, f64p, vp]
    L.mof_fft_stream_set_prev.argtypes = [vp, C.c_int, u8p, C.c_int]
    L.mof_fft_stream_reset.argtypes = [vp, C.c_int]
    L.mof_bm_geometry.argtypes = [vp, C.c_int, C.c_int, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.mof_bm_pairs.argtypes = [vp, u8p, u8p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int, vp, vp]
    L.mof_bm_pairs_device.argtypes = [vp, u8p, u8p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int,
                                      vp, vp, vp]
    L.mof_scalerot_pairs.argtypes = [vp, u8p, u8p, C.c_size_t, C.c_int, C.c_int, f64p]
    L.mof_scalerot_pairs_device.argtypes = [vp, u8p, u8p, C.c_size_t, C.c_int, C.c_int, f64p, vp]
    L.mof_scalerot_push.argtypes = [vp, u8p, C.c_int, f64p]
    L.mof_scalerot_reset.argtypes = [vp]
    L.mof_logpolar.argtypes = [vp, u8p, C.c_int, C.c_int, u8p, C.c_int, C.c_int, C.c_size_t,
                               C.c_size_t]
    L.mof_sync.argtypes = [vp]
    L.mof_kernel_launches.argtypes = [vp]
    L.mof_kernel_launches.restype = C.c_longlong
    L.mof_fft_kernel_time.argtypes = [vp, C.POINTER(C.c_double), C.POINTER(C.c_longlong), C.c_int]
    L.mof_enable_kernel_timing.argtypes = [vp, C.c_int]
    L.mof_version.argtypes = []
    L.mof_version.restype = C.c_char_p
    fdp = C.POINTER(MofFrontendDesc)
    L.mof_frontend.argtypes = [vp, fdp, u8p, C.c_int, u8p]
    L.mof_frontend_device.argtypes = [vp, fdp, u8p, C.c_int, u8p, C.c_size_t, C.c_int, vp]
    L.mof_fft_stream_push_raw.argtypes = [vp, C.c_int, C.POINTER(C.c_int), fdp, u8p, f64p]
    L.mof_fft_stream_push_raw_device.argtypes = [vp, C.c_int, C.POINTER(C.c_int), fdp, u8p, f64p, vp]
    L.mof_longrange_geometry.argtypes = [vp, C.POINTER(C.c_int), C.POINTER(C.c_int)]
    L.mof_fft_pairs_longrange.argtypes = [vp, u8p, u8p, C.c_size_t, C.c_int, C.c_int, f64p]
    L.mof_fft_pairs_longrange_device.argtypes = [vp, u8p, u8p, C.c_size_t, C.c_int, C.c_int, f64p, vp]
    L.mof_point_pairs.argtypes = [vp, f64p, C.c_int, C.c_int, C.c_int, f64p, f64p, vp]
    L.mof_point_pairs_device.argtypes = [vp, f64p, C.c_int, C.c_int, C.c_int, f64p, f64p, vp, vp]
    L.mof_bm_refine.argtypes = [vp, u8p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_int, f64p]
    L.mof_bm_refine_device.argtypes = [vp, u8p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.c_int,
                                       f64p, vp]
    L.mof_bm_refine.restype = C.c_int
    L.mof_bm_refine_device.restype = C.c_int
    L.mof_point_pairs.restype = C.c_int
    L.mof_point_pairs_device.restype = C.c_int
    for name in ("mof_frontend", "mof_frontend_device", "mof_fft_stream_push_raw",
                 "mof_fft_stream_push_raw_device", "mof_longrange_geometry", "mof_fft_pairs_longrange",
                 "mof_fft_pairs_longrange_device"):
        getattr(L, name).restype = C.c_int
    for name in ("mof_create", "mof_get_geometry", "mof_fft_pairs", "mof_fft_pairs_device",
                 "mof_fft_stream_push", "mof_fft_stream_push_device", "mof_fft_stream_set_prev",
                 "mof_fft_stream_reset", "mof_bm_geometry", "mof_bm_pairs", "mof_bm_pairs_device",
                 "mof_scalerot_pairs", "mof_scalerot_pairs_device", "mof_scalerot_push", "mof_scalerot_reset", "mof_logpolar", "mof_sync",
                 "mof_fft_kernel_time", "mof_enable_kernel_timing"):
        getattr(L, name).restype = C.c_int
    ip = C.POINTER(C.c_int)
    L.mof_fft_stream_push_many.argtypes = [vp, C.c_int, ip, C.c_int, u8p, C.c_size_t, C.c_size_t, C.c_int, f64p]
    L.mof_fft_stream_push_many_device.argtypes = [vp, C.c_int, ip, C.c_int, u8p, C.c_size_t, C.c_size_t, C.c_int,
                                                  f64p, vp]
    L.mof_fft_stream_push_many_raw.argtypes = [vp, C.c_int, ip, C.c_int, fdp, u8p, f64p]
    L.mof_fft_stream_push_many_raw_device.argtypes = [vp, C.c_int, ip, C.c_int, fdp, u8p, f64p, vp]
    L.mof_multi_create.argtypes = [C.POINTER(vp), C.POINTER(MofConfig), ip, C.c_int]
    L.mof_multi_destroy.argtypes = [vp]
    L.mof_multi_destroy.restype = None
    L.mof_multi_device_count.argtypes = [vp]
    L.mof_multi_last_error.argtypes = [vp]
    L.mof_multi_last_error.restype = C.c_char_p
    L.mof_multi_get_geometry.argtypes = [vp, ip, ip, ip]
    L.mof_multi_fft_pairs.argtypes = [vp, u8p, u8p, C.c_size_t, C.c_int, C.c_int, f64p]
    L.mof_multi_fft_stream_push.argtypes = [vp, C.c_int, ip, u8p, C.c_size_t, C.c_int, f64p]
    L.mof_multi_fft_stream_push_many.argtypes = [vp, C.c_int, ip, C.c_int, u8p, C.c_size_t, C.c_size_t, C.c_int, f64p]
    L.mof_multi_fft_stream_push_many_raw.argtypes = [vp, C.c_int, ip, C.c_int, fdp, u8p, f64p]
    L.mof_multi_kernel_launches.argtypes = [vp]
    L.mof_multi_kernel_launches.restype = C.c_longlong
    for name in ("mof_fft_stream_push_many", "mof_fft_stream_push_many_device", "mof_fft_stream_push_many_raw",
                 "mof_fft_stream_push_many_raw_device", "mof_multi_create", "mof_multi_device_count",
                 "mof_multi_get_geometry", "mof_multi_fft_pairs", "mof_multi_fft_stream_push",
                 "mof_multi_fft_stream_push_many", "mof_multi_fft_stream_push_many_raw"):
        getattr(L, name).restype = C.c_int
    _lib = L
    return L
