"""Host-side mirror of the reference's operator interface for the hot path, on top of the C ABI.

`Backend` wraps one `mof_ctx` (one CUDA device).  `FftMethod`, `FastSpacedBMMethod` and
`scaleRotationEstimator` keep the reference's constructor / processImage signatures
(include/FftMethod.h:434-441, include/FastSpacedBMMethod_OCL.h:39-42,
include/scaleRotationEstimator.h:15-16) so that callers and tests read like the reference's.
Images are numpy uint8 arrays (the cv::Mat CV_8UC1 of the reference) or torch tensors; CUDA torch
tensors take the device-pointer entry points and stay on the GPU.
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np

from . import _lib
from ._lib import MofConfig, MofError, MofFrontendDesc

try:  # torch is plumbing only (device memory / streams); the backend itself does not need it
    import torch
except Exception:  # pragma: no cover
    torch = None


def _stream_handle(device) -> int:
    """cudaStream_t of torch's current stream; the legacy default stream (handle 0) is passed as
    cudaStreamLegacy (0x1) because NULL means "the context's own stream" in the C ABI."""
    h = torch.cuda.current_stream(device).cuda_stream
    return h if h else 1


def _is_torch(x) -> bool:
    return torch is not None and isinstance(x, torch.Tensor)


def _host_u8(x, ndim: int):
    """-> (keepalive, pointer, shape, strides) of a uint8 host array with contiguous pixels."""
    if _is_torch(x):
        assert x.dtype == torch.uint8 and not x.is_cuda
        if x.stride(-1) != 1:
            x = x.contiguous()
        return x, x.data_ptr(), tuple(x.shape), tuple(s for s in x.stride())
    a = np.asarray(x)
    assert a.dtype == np.uint8, "images must be uint8 (CV_8UC1)"
    if a.strides[-1] != 1:
        a = np.ascontiguousarray(a)
    assert a.ndim == ndim
    return a, a.ctypes.data, a.shape, a.strides


class Backend:
    """One CUDA context of the B200 backend (RAII over mof_create / mof_destroy)."""

    def __init__(self, frame_size: int = 480, sample_point_size: int = 120,
                 max_px_speed: float = 80.0, *, hann: bool = False, max_batch: int = 64,
                 n_streams: int = 0, bm_block: int = 120, bm_step: int = 24, bm_radius: int = 21,
                 sr_resolution: int = 0, sr_m: float = 49.9, device: int = 0, bm_no_gate: bool = False,
                 ocl_compat: bool = False):
        self._L = _lib.load()
        cfg = MofConfig()
        self._L.mof_default_config(C.byref(cfg))
        cfg.device = device
        cfg.ocl_compat = 1 if ocl_compat else 0
        cfg.frame_size = frame_size
        cfg.sample_point_size = sample_point_size
        cfg.max_px_speed = float(max_px_speed)
        cfg.hann = 1 if hann else 0
        cfg.max_batch = max_batch
        cfg.n_streams = n_streams
        cfg.bm_block, cfg.bm_step, cfg.bm_radius = bm_block, bm_step, bm_radius
        cfg.sr_resolution, cfg.sr_m = sr_resolution, float(sr_m)
        cfg.bm_no_gate = 1 if bm_no_gate else 0
        self._ctx = C.c_void_p()
        rc = self._L.mof_create(C.byref(self._ctx), C.byref(cfg))
        if rc != 0:
            raise MofError(rc, self._L.mof_last_error(None).decode())
        F, N, S = C.c_int(), C.c_int(), C.c_int()
        self._L.mof_get_geometry(self._ctx, C.byref(F), C.byref(N), C.byref(S))
        self.F, self.N, self.S = F.value, N.value, S.value
        self.device = device
        self.cfg = cfg

    # -- plumbing ---------------------------------------------------------------------------
    def close(self):
        if getattr(self, "_ctx", None) is not None and self._ctx.value:
            self._L.mof_destroy(self._ctx)
            self._ctx = C.c_void_p()

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    def _chk(self, rc: int):
        if rc != 0:
            raise MofError(rc, self._L.mof_last_error(self._ctx).decode())

    def sync(self):
        self._chk(self._L.mof_sync(self._ctx))

    @property
    def kernel_launches(self) -> int:
        return int(self._L.mof_kernel_launches(self._ctx))

    def enable_kernel_timing(self, on: bool = True):
        self._chk(self._L.mof_enable_kernel_timing(self._ctx, 1 if on else 0))

    def fft_kernel_time(self, reset: bool = True):
        """-> (average device ms per phase-correlation launch, launches measured)."""
        ms, n = C.c_double(), C.c_longlong()
        self._chk(self._L.mof_fft_kernel_time(self._ctx, C.byref(ms), C.byref(n), 1 if reset else 0))
        return ms.value, n.value

    # -- FftMethod path -----------------------------------------------------------------------
    def fft_pairs(self, curr, prev, out=None):
        """shift = -cv::phaseCorrelate(curr_roi, prev_roi) for every section of every pair.

        curr/prev: [n, H, W] uint8 (H, W >= F; the top-left F x F is used).  Host arrays go through
        mof_fft_pairs (copies included), CUDA tensors through mof_fft_pairs_device (async on the
        current torch stream).  Returns [n, S, 2] float64 (numpy or CUDA tensor)."""
        if _is_torch(curr) and curr.is_cuda:
            assert prev.is_cuda and curr.dtype == torch.uint8 and prev.dtype == torch.uint8
            assert curr.stride(-1) == 1 and prev.stride() == curr.stride() and curr.dim() == 3
            n = curr.shape[0]
            if out is None:
                out = torch.empty((n, self.S, 2), dtype=torch.float64, device=curr.device)
            st = _stream_handle(curr.device)
            self._chk(self._L.mof_fft_pairs_device(self._ctx, curr.data_ptr(), prev.data_ptr(),
                                                   curr.stride(0), curr.stride(1), n,
                                                   out.data_ptr(), st))
            return out
        kc, pc, shp, strd = _host_u8(curr, 3)
        kp, pp, shp2, strd2 = _host_u8(prev, 3)
        assert shp == shp2 and strd == strd2 and shp[1] >= self.F and shp[2] >= self.F
        n = shp[0]
        if out is None:
            out = np.empty((n, self.S, 2), np.float64)
        optr = out.data_ptr() if _is_torch(out) else out.ctypes.data
        self._chk(self._L.mof_fft_pairs(self._ctx, pc, pp, strd[0], strd[1], n, optr))
        return out

    def fft_stream_push(self, stream_ids, frames, out=None):
        """One frame for each of the given streams (stateful imPrev semantics) -> [n, S, 2]."""
        ids = np.ascontiguousarray(stream_ids, dtype=np.int32)
        n = ids.shape[0]
        idp = ids.ctypes.data_as(C.POINTER(C.c_int))
        if _is_torch(frames) and frames.is_cuda:
            assert frames.dtype == torch.uint8 and frames.dim() == 3 and frames.stride(-1) == 1
            if out is None:
                out = torch.empty((n, self.S, 2), dtype=torch.float64, device=frames.device)
            st = _stream_handle(frames.device)
            self._chk(self._L.mof_fft_stream_push_device(self._ctx, n, idp, frames.data_ptr(),
                                                         frames.stride(0), frames.stride(1),
                                                         out.data_ptr(), st))
            return out
        kf, pf, shp, strd = _host_u8(frames, 3)
        assert shp[0] == n
        if out is None:
            out = np.empty((n, self.S, 2), np.float64)
        optr = out.data_ptr() if _is_torch(out) else out.ctypes.data
        self._chk(self._L.mof_fft_stream_push(self._ctx, n, idp, pf, strd[0], strd[1], optr))
        return out

    def fft_stream_push_many(self, stream_ids, frames, out=None):
        """T consecutive frames for each of the given streams in one call / one launch.
        frames [n, T, H, W] uint8 -> [n, T, S, 2]: what T successive fft_stream_push calls would return."""
        ids = np.ascontiguousarray(stream_ids, dtype=np.int32)
        n = ids.shape[0]
        idp = ids.ctypes.data_as(C.POINTER(C.c_int))
        if _is_torch(frames) and frames.is_cuda:
            assert frames.dtype == torch.uint8 and frames.dim() == 4 and frames.stride(-1) == 1 and frames.shape[0] == n
            T = frames.shape[1]
            if out is None:
                out = torch.empty((n, T, self.S, 2), dtype=torch.float64, device=frames.device)
            self._chk(self._L.mof_fft_stream_push_many_device(self._ctx, n, idp, T, frames.data_ptr(), frames.stride(0),
                                                              frames.stride(1), frames.stride(2), out.data_ptr(),
                                                              _stream_handle(frames.device)))
            return out
        kf, pf, shp, strd = _host_u8(frames, 4)
        assert shp[0] == n
        T = shp[1]
        if out is None:
            out = np.empty((n, T, self.S, 2), np.float64)
        optr = out.data_ptr() if _is_torch(out) else out.ctypes.data
        self._chk(self._L.mof_fft_stream_push_many(self._ctx, n, idp, T, pf, strd[0], strd[1], strd[2], optr))
        return out

    def fft_stream_push_many_raw(self, stream_ids, raw, scale_factor: int = 1, cx=None, out=None):
        """Raw camera frames [n, T, H, W(, 3)] (contiguous) -> front end -> push_many -> [n, T, S, 2]."""
        ids = np.ascontiguousarray(stream_ids, dtype=np.int32)
        n = ids.shape[0]
        idp = ids.ctypes.data_as(C.POINTER(C.c_int))
        T = raw.shape[1]
        flat = raw.reshape((n * T,) + tuple(raw.shape[2:]))
        if _is_torch(raw) and raw.is_cuda:
            assert raw.is_contiguous()
            d = self._fe_desc(flat, scale_factor, cx)
            if out is None:
                out = torch.empty((n, T, self.S, 2), dtype=torch.float64, device=raw.device)
            self._chk(self._L.mof_fft_stream_push_many_raw_device(self._ctx, n, idp, T, C.byref(d), raw.data_ptr(),
                                                                  out.data_ptr(), _stream_handle(raw.device)))
            return out
        if _is_torch(flat):            # host tensor (e.g. pinned): no copy
            assert flat.is_contiguous() and flat.dtype == torch.uint8
            a, aptr = flat, flat.data_ptr()
        else:
            a = np.ascontiguousarray(flat)
            aptr = a.ctypes.data
        d = self._fe_desc(a, scale_factor, cx)
        if out is None:
            out = np.empty((n, T, self.S, 2), np.float64)
        optr = out.data_ptr() if _is_torch(out) else out.ctypes.data
        self._chk(self._L.mof_fft_stream_push_many_raw(self._ctx, n, idp, T, C.byref(d), aptr, optr))
        return out

    def fft_stream_set_prev(self, stream_id: int, frame):
        kf, pf, shp, strd = _host_u8(frame, 2)
        self._chk(self._L.mof_fft_stream_set_prev(self._ctx, stream_id, pf, strd[0]))

    def fft_stream_reset(self, stream_id: int):
        self._chk(self._L.mof_fft_stream_reset(self._ctx, stream_id))

    # -- camera front end (src/optic_flow.cpp:1603-1622) ---------------------------------------------
    @staticmethod
    def _fe_desc(raw, scale_factor: int, cx):
        """raw: [n, H, W] (MONO8) or [n, H, W, 3] (BGR8), pixels contiguous."""
        shp = tuple(raw.shape)
        strd = tuple(raw.stride()) if _is_torch(raw) else tuple(raw.strides)
        ch = 1 if len(shp) == 3 else shp[3]
        assert len(shp) in (3, 4) and ch in (1, 3)
        if len(shp) == 4:
            assert strd[3] == 1 and strd[2] == ch
        else:
            assert strd[2] == 1
        d = MofFrontendDesc()
        d.height, d.width, d.channels = shp[1], shp[2], ch
        d.row_stride, d.image_stride = strd[1], strd[0]
        d.scale_factor = int(scale_factor)
        d.cx = int(cx) if cx is not None else (shp[2] // max(1, int(scale_factor))) // 2
        return d

    def frontend(self, raw, scale_factor: int = 1, cx=None, out=None):
        """cv::resize(1/scale_factor) -> crop F x F at (cx - F/2, h/2 - F/2) -> RGB2GRAY.
        -> [n, F, F] uint8 (numpy for host input, CUDA tensor for CUDA input)."""
        if _is_torch(raw) and raw.is_cuda:
            d = self._fe_desc(raw, scale_factor, cx)
            n = raw.shape[0]
            if out is None:
                out = torch.empty((n, self.F, self.F), dtype=torch.uint8, device=raw.device)
            st = _stream_handle(raw.device)
            self._chk(self._L.mof_frontend_device(self._ctx, C.byref(d), raw.data_ptr(), n, out.data_ptr(),
                                                  out.stride(0), out.stride(1), st))
            return out
        a = np.asarray(raw)
        assert a.dtype == np.uint8
        if a.strides[-1] != 1 or (a.ndim == 4 and a.strides[2] != a.shape[3]):
            a = np.ascontiguousarray(a)
        d = self._fe_desc(a, scale_factor, cx)
        g = np.empty((a.shape[0], self.F, self.F), np.uint8)
        self._chk(self._L.mof_frontend(self._ctx, C.byref(d), a.ctypes.data, a.shape[0], g.ctypes.data))
        return g

    def fft_stream_push_raw(self, stream_ids, raw, scale_factor: int = 1, cx=None, out=None):
        """Raw camera frames -> front end -> stateful FftMethod::processImage, all on the device."""
        ids = np.ascontiguousarray(stream_ids, dtype=np.int32)
        n = ids.shape[0]
        idp = ids.ctypes.data_as(C.POINTER(C.c_int))
        if _is_torch(raw) and raw.is_cuda:
            d = self._fe_desc(raw, scale_factor, cx)
            if out is None:
                out = torch.empty((n, self.S, 2), dtype=torch.float64, device=raw.device)
            st = _stream_handle(raw.device)
            self._chk(self._L.mof_fft_stream_push_raw_device(self._ctx, n, idp, C.byref(d), raw.data_ptr(),
                                                             out.data_ptr(), st))
            return out
        a = np.asarray(raw)
        assert a.dtype == np.uint8 and a.shape[0] == n
        if a.strides[-1] != 1 or (a.ndim == 4 and a.strides[2] != a.shape[3]):
            a = np.ascontiguousarray(a)
        d = self._fe_desc(a, scale_factor, cx)
        if out is None:
            out = np.empty((n, self.S, 2), np.float64)
        self._chk(self._L.mof_fft_stream_push_raw(self._ctx, n, idp, C.byref(d), a.ctypes.data, out.ctypes.data))
        return out

    # -- long-range mode (src/FftMethod.cpp:1905-1990) -----------------------------------------------
    def longrange_geometry(self):
        """-> (F_lr, S_lr): down-scaled frame side and sections per pair."""
        f, s_ = C.c_int(), C.c_int()
        self._chk(self._L.mof_longrange_geometry(self._ctx, C.byref(f), C.byref(s_)))
        return f.value, s_.value

    def fft_pairs_longrange(self, curr, prev, out=None):
        _, s_lr = self.longrange_geometry()
        if _is_torch(curr) and curr.is_cuda:
            assert prev.is_cuda and curr.stride(-1) == 1 and prev.stride() == curr.stride() and curr.dim() == 3
            n = curr.shape[0]
            if out is None:
                out = torch.empty((n, s_lr, 2), dtype=torch.float64, device=curr.device)
            st = _stream_handle(curr.device)
            self._chk(self._L.mof_fft_pairs_longrange_device(self._ctx, curr.data_ptr(), prev.data_ptr(),
                                                             curr.stride(0), curr.stride(1), n, out.data_ptr(), st))
            return out
        kc, pc, shp, strd = _host_u8(curr, 3)
        kp, pp, shp2, strd2 = _host_u8(prev, 3)
        assert shp == shp2 and strd == strd2 and shp[1] >= self.F and shp[2] >= self.F
        if out is None:
            out = np.empty((shp[0], s_lr, 2), np.float64)
        self._chk(self._L.mof_fft_pairs_longrange(self._ctx, pc, pp, strd[0], strd[1], shp[0], out.ctypes.data))
        return out

    # -- point pairs for getRT / get2DT (src/optic_flow.cpp:527-550) -------------------------------
    def point_pairs(self, shifts, sq=None, section_size=None):
        """shifts [n, S, 2] -> (initial [n, S, 2], shifted [n, S, 2], counts [n]): the finite sections in
        section order, NaN past counts[f]."""
        sq = int(round(math.sqrt(shifts.shape[1]))) if sq is None else sq
        N = self.N if section_size is None else section_size
        n = shifts.shape[0]
        if _is_torch(shifts) and shifts.is_cuda:
            assert shifts.dtype == torch.float64 and shifts.is_contiguous()
            a, b = torch.empty_like(shifts), torch.empty_like(shifts)
            cnt = torch.empty((n,), dtype=torch.int32, device=shifts.device)
            self._chk(self._L.mof_point_pairs_device(self._ctx, shifts.data_ptr(), n, sq, N, a.data_ptr(),
                                                     b.data_ptr(), cnt.data_ptr(), _stream_handle(shifts.device)))
            return a, b, cnt
        sh = np.ascontiguousarray(shifts, dtype=np.float64)
        a, b = np.empty_like(sh), np.empty_like(sh)
        cnt = np.empty((n,), np.int32)
        self._chk(self._L.mof_point_pairs(self._ctx, sh.ctypes.data, n, sq, N, a.ctypes.data, b.ctypes.data,
                                          cnt.ctypes.data))
        return a, b, cnt

    def bm_refine(self, curr, fullpix, passes: int = 2):
        """BlockMethod::Refine (src/BlockMethod.cpp:96-147): curr [n, h, w] uint8, fullpix [n, 2] int32
        (x, y) -> refined [n, 2] float64."""
        if _is_torch(curr) and curr.is_cuda:
            n, H, W = curr.shape
            assert curr.dtype == torch.uint8 and curr.stride(2) == 1
            fp = fullpix.to(device=curr.device, dtype=torch.int32).contiguous()
            out = torch.empty((n, 2), dtype=torch.float64, device=curr.device)
            self._chk(self._L.mof_bm_refine_device(self._ctx, curr.data_ptr(), curr.stride(0), curr.stride(1), W, H,
                                                   n, fp.data_ptr(), passes, out.data_ptr(),
                                                   _stream_handle(curr.device)))
            return out
        kc, pc, shp, strd = _host_u8(curr, 3)
        n, H, W = shp
        fp = np.ascontiguousarray(np.asarray(fullpix).reshape(n, 2), dtype=np.int32)
        out = np.empty((n, 2), np.float64)
        self._chk(self._L.mof_bm_refine(self._ctx, pc, strd[0], strd[1], W, H, n, fp.ctypes.data, passes,
                                        out.ctypes.data))
        return out

    # -- FastSpacedBM path ----------------------------------------------------------------------
    def bm_geometry(self, width: int, height: int):
        gx, gy = C.c_int(), C.c_int()
        self._chk(self._L.mof_bm_geometry(self._ctx, width, height, C.byref(gx), C.byref(gy)))
        return gx.value, gy.value

    def bm_pairs(self, curr, prev):
        """-> (flow int8 [n, 2, gy, gx] (x plane, y plane), voted float32 [n, 2])."""
        if _is_torch(curr) and curr.is_cuda:
            n, H, W = curr.shape
            gx, gy = self.bm_geometry(W, H)
            flow = torch.empty((n, 2, gy, gx), dtype=torch.int8, device=curr.device)
            voted = torch.empty((n, 2), dtype=torch.float32, device=curr.device)
            st = _stream_handle(curr.device)
            self._chk(self._L.mof_bm_pairs_device(self._ctx, curr.data_ptr(), prev.data_ptr(),
                                                  curr.stride(0), W, H, curr.stride(1), n,
                                                  flow.data_ptr(), voted.data_ptr(), st))
            return flow, voted
        kc, pc, shp, strd = _host_u8(curr, 3)
        kp, pp, shp2, strd2 = _host_u8(prev, 3)
        assert shp == shp2 and strd == strd2
        n, H, W = shp
        gx, gy = self.bm_geometry(W, H)
        flow = np.empty((n, 2, gy, gx), np.int8)
        voted = np.empty((n, 2), np.float32)
        self._chk(self._L.mof_bm_pairs(self._ctx, pc, pp, strd[0], W, H, strd[1], n,
                                       flow.ctypes.data, voted.ctypes.data))
        return flow, voted

    # -- scale / rotation path --------------------------------------------------------------------
    def scalerot_pairs(self, curr, prev):
        if _is_torch(curr) and curr.is_cuda:
            n = curr.shape[0]
            out = torch.empty((n, 2), dtype=torch.float64, device=curr.device)
            st = _stream_handle(curr.device)
            self._chk(self._L.mof_scalerot_pairs_device(self._ctx, curr.data_ptr(), prev.data_ptr(),
                                                        curr.stride(0), curr.stride(1), n,
                                                        out.data_ptr(), st))
            return out
        kc, pc, shp, strd = _host_u8(curr, 3)
        kp, pp, shp2, strd2 = _host_u8(prev, 3)
        assert shp == shp2 and strd == strd2
        out = np.empty((shp[0], 2), np.float64)
        self._chk(self._L.mof_scalerot_pairs(self._ctx, pc, pp, strd[0], strd[1], shp[0],
                                             out.ctypes.data))
        return out

    def scalerot_push(self, frame):
        kf, pf, shp, strd = _host_u8(frame, 2)
        out = np.empty(2, np.float64)
        self._chk(self._L.mof_scalerot_push(self._ctx, pf, strd[0], out.ctypes.data))
        return float(out[0]), float(out[1])

    def scalerot_reset(self):
        self._chk(self._L.mof_scalerot_reset(self._ctx))

    def logpolar(self, src, interp: int, dst=None):
        """cv::logPolar(src, dst, (res/2, res/2), M, interp) on [n, res, res] or [res, res] uint8."""
        a = np.ascontiguousarray(src, dtype=np.uint8)
        single = a.ndim == 2
        if single:
            a = a[None]
        n, res, _ = a.shape
        d = np.zeros_like(a) if dst is None else np.ascontiguousarray(dst, dtype=np.uint8).reshape(a.shape).copy()
        self._chk(self._L.mof_logpolar(self._ctx, a.ctypes.data, res, interp, d.ctypes.data, res, n,
                                       res * res, res * res))
        return d[0] if single else d


class MultiBackend:
    """Several GPUs behind one handle (mof_multi_*): one context + one host worker thread per device
    inside the library; pairs in contiguous slices, stream s on device s % G; host arrays only."""

    def __init__(self, devices, frame_size: int = 480, sample_point_size: int = 120, max_px_speed: float = 80.0, *,
                 max_batch: int = 64, n_streams: int = 0, hann: bool = False, ocl_compat: bool = False):
        self._L = _lib.load()
        cfg = MofConfig()
        self._L.mof_default_config(C.byref(cfg))
        cfg.frame_size, cfg.sample_point_size = frame_size, sample_point_size
        cfg.max_px_speed = float(max_px_speed)
        cfg.max_batch, cfg.n_streams = max_batch, n_streams
        cfg.hann = 1 if hann else 0
        cfg.ocl_compat = 1 if ocl_compat else 0
        cfg.sr_resolution = 0
        dev = (C.c_int * len(devices))(*devices)
        self._m = C.c_void_p()
        rc = self._L.mof_multi_create(C.byref(self._m), C.byref(cfg), dev, len(devices))
        if rc != 0:
            raise MofError(rc, self._L.mof_multi_last_error(None).decode())
        F, N, S = C.c_int(), C.c_int(), C.c_int()
        self._L.mof_multi_get_geometry(self._m, C.byref(F), C.byref(N), C.byref(S))
        self.F, self.N, self.S = F.value, N.value, S.value
        self.n_devices = len(devices)

    def close(self):
        if getattr(self, "_m", None) is not None and self._m.value:
            self._L.mof_multi_destroy(self._m)
            self._m = C.c_void_p()

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass

    def _chk(self, rc: int):
        if rc != 0:
            raise MofError(rc, self._L.mof_multi_last_error(self._m).decode())

    @property
    def kernel_launches(self) -> int:
        return int(self._L.mof_multi_kernel_launches(self._m))

    def fft_pairs(self, curr, prev, out=None):
        kc, pc, shp, strd = _host_u8(curr, 3)
        kp, pp, shp2, strd2 = _host_u8(prev, 3)
        assert shp == shp2 and strd == strd2
        if out is None:
            out = np.empty((shp[0], self.S, 2), np.float64)
        optr = out.data_ptr() if _is_torch(out) else out.ctypes.data
        self._chk(self._L.mof_multi_fft_pairs(self._m, pc, pp, strd[0], strd[1], shp[0], optr))
        return out

    def fft_stream_push(self, stream_ids, frames, out=None):
        ids = np.ascontiguousarray(stream_ids, dtype=np.int32)
        kf, pf, shp, strd = _host_u8(frames, 3)
        if out is None:
            out = np.empty((shp[0], self.S, 2), np.float64)
        self._chk(self._L.mof_multi_fft_stream_push(self._m, ids.shape[0], ids.ctypes.data_as(C.POINTER(C.c_int)), pf,
                                                    strd[0], strd[1], out.ctypes.data))
        return out

    def fft_stream_push_many(self, stream_ids, frames, out=None):
        ids = np.ascontiguousarray(stream_ids, dtype=np.int32)
        kf, pf, shp, strd = _host_u8(frames, 4)
        if out is None:
            out = np.empty((shp[0], shp[1], self.S, 2), np.float64)
        optr = out.data_ptr() if _is_torch(out) else out.ctypes.data
        self._chk(self._L.mof_multi_fft_stream_push_many(self._m, ids.shape[0], ids.ctypes.data_as(C.POINTER(C.c_int)),
                                                         shp[1], pf, strd[0], strd[1], strd[2], optr))
        return out

    def fft_stream_push_many_raw(self, stream_ids, raw, scale_factor: int = 1, cx=None, out=None):
        ids = np.ascontiguousarray(stream_ids, dtype=np.int32)
        n, T = raw.shape[0], raw.shape[1]
        if _is_torch(raw):
            assert raw.is_contiguous() and not raw.is_cuda
            flat = raw.reshape((n * T,) + tuple(raw.shape[2:]))
            ptr = raw.data_ptr()
        else:
            raw = np.ascontiguousarray(raw)
            flat = raw.reshape((n * T,) + raw.shape[2:])
            ptr = raw.ctypes.data
        d = Backend._fe_desc(flat, scale_factor, cx)
        if out is None:
            out = np.empty((n, T, self.S, 2), np.float64)
        optr = out.data_ptr() if _is_torch(out) else out.ctypes.data
        self._chk(self._L.mof_multi_fft_stream_push_many_raw(self._m, n, ids.ctypes.data_as(C.POINTER(C.c_int)), T,
                                                             C.byref(d), ptr, optr))
        return out


class FftMethod:
    """Drop-in for the reference's FftMethod (OpticFlowCalc subclass).

    Constructor and processImage keep the reference's argument lists (include/FftMethod.h:434-441);
    storeVideo / raw_enable / rot_corr_enable / tilt_corr_enable / videoPath / videoFPS /
    cl_file_name / useOCL are accepted and ignored, exactly the arguments the CPU branch never reads
    (GUI and video are out of scope on a headless box)."""

    def __init__(self, frameSize, samplePointSize, max_px_speed, storeVideo=False, raw_enable=False,
                 rot_corr_enable=False, tilt_corr_enable=False, videoPath=None, videoFPS=30,
                 cl_file_name="", useOCL=False, *, device=0, hann=False, ocl_compat=False):
        # useOCL selects the reference's OpenCL kernel, which cannot run here; ocl_compat=True reproduces
        # that branch's numerics instead (parity unpinned), the default is the CPU branch
        self._be = Backend(frameSize, samplePointSize, max_px_speed, n_streams=1, max_batch=1,
                           sr_resolution=0, device=device, hann=hann, ocl_compat=ocl_compat)
        self.frameSize, self.samplePointSize = self._be.F, self._be.N
        self.sqNum = self._be.F // self._be.N
        self.max_px_speed_sq = float(max_px_speed) ** 2
        self._ids = np.zeros(1, np.int32)
        self.running = False
        # one `first` flag and one imPrev for both modes (src/FftMethod.cpp:1761, :1791-1793, :1872,
        # :1900, :1920, :2004); _spec_valid: the device cache holds the spectrum of exactly self.imPrev
        self.first = True
        self.imPrev = None
        self._spec_valid = False

    def setImPrev(self, imPrev):  # include/OpticFlowCalc.h:14-16
        """Stores the frame; like the reference, the FIRST processImage call overwrites it with the
        current frame (src/FftMethod.cpp:1791-1793), later calls correlate against it."""
        F = self._be.F
        self.imPrev = np.array(np.asarray(imPrev)[:F, :F], dtype=np.uint8, copy=True)
        self._spec_valid = False

    def processImage(self, imCurr, gui=False, debug=False, midPoint_t=None, yaw_angle=0.0,
                     rot_center=None, raw_output=None, i_fx=300.0, i_fy=300.0):
        """-> list of (x, y) per section, index i + j*sqNum; (nan, nan) for invalid sections;
        [] when re-entered while running (src/FftMethod.cpp:1775-1776)."""
        if self.running:
            return []
        self.running = True
        try:
            F = self._be.F
            im = np.array(np.asarray(imCurr)[:F, :F], dtype=np.uint8, copy=True)
            if self.first:
                self._be.fft_stream_reset(0)          # prev := curr on the device
            elif not self._spec_valid:
                self._be.fft_stream_set_prev(0, self.imPrev)
            out = self._be.fft_stream_push(self._ids, im[None])
            self.imPrev, self._spec_valid, self.first = im, True, False   # :1872, :1900
            return [(float(x), float(y)) for x, y in out[0]]
        finally:
            self.running = False

    def processImageLongRange(self, imCurr, gui=False, debug=False, midPoint_t=None, yaw_angle=0.0,
                              rot_center=None, raw_output=None, i_fx=300.0, i_fy=300.0):
        """include/FftMethod.h:440-441, src/FftMethod.cpp:1905-1990: both frames down-scaled by 4,
        sqNum/4 x sqNum/4 sections of samplePointSize; shares `first` and imPrev with processImage."""
        if self.running:
            return []
        self.running = True
        try:
            F = self._be.F
            im = np.array(np.asarray(imCurr)[:F, :F], dtype=np.uint8, copy=True)
            if self.first or self.imPrev is None:
                self.imPrev = im                       # :1920-1922
            out = self._be.fft_pairs_longrange(im[None], self.imPrev[None])
            self.imPrev, self._spec_valid, self.first = im, False, False   # :1993, :2004
            return [(float(x), float(y)) for x, y in out[0]]
        finally:
            self.running = False


class FastSpacedBMMethod:
    """Drop-in for FastSpacedBMMethod (include/FastSpacedBMMethod_OCL.h:39-42): the camera
    intrinsics / distortion arguments are stored but, as in the reference, never used by the search."""

    def __init__(self, samplePointSize, scanRadius, stepSize, cx=0, cy=0, fx=0, fy=0, k1=0, k2=0,
                 k3=0, p1=0, p2=0, storeVideo=False, videoPath=None, *, frame_size=240, device=0):
        self._be = Backend(frame_size, frame_size, 1e9, bm_block=samplePointSize, bm_step=stepSize,
                           bm_radius=scanRadius, max_batch=1, device=device)
        self.imPrev = None

    def setImPrev(self, imPrev):
        self.imPrev = np.array(imPrev, dtype=np.uint8, copy=True)

    def processImage(self, imCurr, gui=False, debug=False, midPoint_t=None, yaw_angle=0.0,
                     tiltCorr=None):
        """-> [(x, y)]: the voted shift (src/FastSpacedBMMethod_OCL.cpp:172-175)."""
        im = np.ascontiguousarray(imCurr, dtype=np.uint8)
        if self.imPrev is None:
            self.imPrev = im.copy()
        flow, voted = self._be.bm_pairs(im[None], self.imPrev[None])
        self.flow = flow[0]
        self.imPrev = im.copy()  # :181
        return [(float(voted[0, 0]), float(voted[0, 1]))]


class BlockMethod:
    """Integer stage of the reference's BlockMethod (include/BlockMethod.h, src/BlockMethod.cpp:3-94):
    same constructor arguments; processImage returns the voted full-pixel shift passed through the
    sub-pixel `Refine` pass (src/BlockMethod.cpp:78, :96-147 -- which in the reference compares the
    up-scaled current frame with itself) as [(x, y)]; `voted` keeps the integer vote."""

    def __init__(self, frameSize, samplePointSize, scanRadius, scanDiameter=None, scanCount=None,
                 stepSize=0, *, device=0):
        self._be = Backend(frameSize, frameSize, 1e9, bm_block=samplePointSize, bm_step=0,
                           bm_radius=scanRadius, bm_no_gate=True, max_batch=1, device=device)
        self.frameSize = frameSize
        self.imPrev = np.zeros((frameSize, frameSize), np.uint8)   # src/BlockMethod.cpp:19-20

    def processImage(self, imCurr, gui=False, debug=False, midPoint_t=None, yaw_angle=0.0,
                     tiltCorr=None):
        im = np.ascontiguousarray(np.asarray(imCurr)[:self.frameSize, :self.frameSize], dtype=np.uint8)
        flow, voted = self._be.bm_pairs(im[None], self.imPrev[None])
        self.flow = flow[0]
        self.voted = (int(voted[0, 0]), int(voted[0, 1]))
        refined = self._be.bm_refine(im[None], np.array([self.voted], np.int32), 2)   # :78
        self.imPrev = im.copy()   # :86
        return [(float(refined[0, 0]), float(refined[0, 1]))]


class scaleRotationEstimator:  # noqa: N801  (the reference's class name)
    """Drop-in for scaleRotationEstimator (include/scaleRotationEstimator.h:15-16)."""

    def __init__(self, resolution, m, storeVideo=False, videoPath=None, videoFPS=30, *, device=0):
        self._be = Backend(resolution, resolution, 1e9, sr_resolution=resolution, sr_m=m,
                           max_batch=1, device=device)
        self.resolution = resolution

    def processImage(self, imCurr, gui=False, debug=False):
        """-> (scale, rotation[rad]); (1, 0) on the first call and when rejected."""
        return self._be.scalerot_push(np.ascontiguousarray(imCurr, dtype=np.uint8))


def expected_rot_from_shift(pt_y: float, res: int) -> float:
    """(pt.y / Ky) * pi/180 with Ky = res/360 (src/scaleRotationEstimator.cpp:26,124)."""
    return (pt_y / (res / 360.0)) * (math.pi / 180.0)
