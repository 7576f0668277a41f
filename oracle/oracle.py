"""ctypes binding of the CPU oracle (oracle/libmof_oracle.so).  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this module; the product package (mrs_optic_flow_b200) never does.  See mof_oracle.h for
the reference file:line each function restates.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libmof_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile the C restatement (gcc, a second or two).  Building the checker is not using it."""
    src = os.path.join(_HERE, "mof_oracle.c")
    if (force or not os.path.exists(_LIB_PATH)
            or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src)):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _LIB_PATH


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = C.CDLL(_LIB_PATH)
        u8p, f32p, f64p = C.POINTER(C.c_uint8), C.POINTER(C.c_float), C.POINTER(C.c_double)
        i8p, i32p = C.POINTER(C.c_int8), C.POINTER(C.c_int32)
        L.mof_oracle_phase_correlate_surface.argtypes = [
            f32p, C.c_int, f32p, C.c_int, C.c_int, f32p, C.c_int, f64p, f64p, f32p]
        L.mof_oracle_phase_correlate_surface.restype = C.c_int
        L.mof_oracle_fft_pairs.argtypes = [
            u8p, u8p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_double, C.c_int, C.c_int, f64p,
            C.c_int]
        L.mof_oracle_fft_pairs.restype = C.c_int
        L.mof_oracle_fft_pairs_ex.argtypes = L.mof_oracle_fft_pairs.argtypes
        L.mof_oracle_fft_pairs_ex.restype = C.c_int
        L.mof_oracle_hann.argtypes = [C.c_int, f32p]
        L.mof_oracle_hann.restype = None
        L.mof_oracle_bm.argtypes = [u8p, u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int,
                                    i8p, i8p, f32p, i32p]
        L.mof_oracle_bm.restype = C.c_int
        L.mof_oracle_bm_pairs.argtypes = [u8p, u8p, C.c_size_t, C.c_int, C.c_int, C.c_int, C.c_int,
                                          C.c_int, C.c_int, C.c_int, i8p, f32p, C.c_int]
        L.mof_oracle_bm_pairs.restype = C.c_int
        L.mof_oracle_logpolar.argtypes = [u8p, C.c_int, C.c_int, C.c_double, C.c_int, u8p, C.c_int]
        L.mof_oracle_logpolar.restype = C.c_int
        L.mof_oracle_scalerot_pairs.argtypes = [u8p, u8p, C.c_size_t, C.c_int, C.c_int, C.c_double,
                                                C.c_int, f64p, C.c_int]
        L.mof_oracle_scalerot_pairs.restype = C.c_int
        L.mof_oracle_interp_table.argtypes = [C.c_int]
        L.mof_oracle_blockmethod.argtypes = [u8p, u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p,
                                             C.c_void_p, C.c_void_p]
        L.mof_oracle_frontend.argtypes = [u8p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, u8p]
        L.mof_oracle_quarter.argtypes = [u8p, C.c_int, C.c_int, u8p]
        L.mof_oracle_interp_table.restype = C.POINTER(C.c_int16)
        _lib = L
    return _lib


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def phase_correlate(src1: np.ndarray, src2: np.ndarray, window=None, old_mag: bool = False,
                    want_surface: bool = False):
    """cv::phaseCorrelate(src1, src2[, window]) restated -> ((x, y), response[, surface])."""
    a = np.ascontiguousarray(src1, dtype=np.float32)
    b = np.ascontiguousarray(src2, dtype=np.float32)
    n = a.shape[0]
    assert a.shape == (n, n) and b.shape == (n, n)
    out = np.zeros(2, np.float64)
    resp = np.zeros(1, np.float64)
    surf = np.zeros((n, n), np.float32) if want_surface else None
    w = None if window is None else np.ascontiguousarray(window, dtype=np.float32)
    rc = lib().mof_oracle_phase_correlate_surface(
        _p(a, C.c_float), n, _p(b, C.c_float), n, n,
        None if w is None else _p(w, C.c_float), 1 if old_mag else 0,
        _p(out, C.c_double), _p(resp, C.c_double),
        None if surf is None else _p(surf, C.c_float))
    if rc:
        raise RuntimeError(f"mof_oracle_phase_correlate failed rc={rc}")
    if want_surface:
        return (float(out[0]), float(out[1])), float(resp[0]), surf
    return (float(out[0]), float(out[1])), float(resp[0])


def hann(n: int) -> np.ndarray:
    w = np.zeros((n, n), np.float32)
    lib().mof_oracle_hann(n, _p(w, C.c_float))
    return w


def fft_pairs(curr: np.ndarray, prev: np.ndarray, N: int, max_px_speed: float = 1e9,
              hann: bool = False, nthreads: int = 1, ocl_compat: bool = False) -> np.ndarray:
    """FftMethod::processImage CPU branch over [n_pairs, F, F] uint8 stacks -> [n_pairs, S, 2].
    ocl_compat: the useOCL = true branch's numerics restated (parity unpinned, see mof_oracle.h)."""
    curr = np.ascontiguousarray(curr, dtype=np.uint8)
    prev = np.ascontiguousarray(prev, dtype=np.uint8)
    assert curr.shape == prev.shape and curr.ndim == 3
    n, H, W = curr.shape
    F = H - (H % 2)
    Neff = N if F % N == 0 else F
    S = (F // Neff) ** 2
    out = np.zeros((n, S, 2), np.float64)
    rc = lib().mof_oracle_fft_pairs_ex(_p(curr, C.c_uint8), _p(prev, C.c_uint8), H * W, W, H, N,
                                       float(max_px_speed), (1 if hann else 0) | (2 if ocl_compat else 0), n,
                                       _p(out, C.c_double), nthreads)
    if rc:
        raise RuntimeError(f"mof_oracle_fft_pairs failed rc={rc}")
    return out


def bm(curr: np.ndarray, prev: np.ndarray, B: int, step: int, R: int, want_sad: bool = False):
    """FastSpacedBM on one frame pair -> (flow_x[gy,gx], flow_y[gy,gx], voted(2,)[, sad])."""
    curr = np.ascontiguousarray(curr, dtype=np.uint8)
    prev = np.ascontiguousarray(prev, dtype=np.uint8)
    H, W = curr.shape
    pitch = B + step
    gx, gy = (W - 2 * R) // pitch, (H - 2 * R) // pitch
    D = 2 * R + 1
    fx = np.zeros((max(gy, 0), max(gx, 0)), np.int8)
    fy = np.zeros_like(fx)
    voted = np.zeros(2, np.float32)
    sad = np.zeros((max(gy, 0), max(gx, 0), D, D), np.int32) if want_sad else None
    rc = lib().mof_oracle_bm(_p(curr, C.c_uint8), _p(prev, C.c_uint8), W, H, W, B, step, R,
                             _p(fx, C.c_int8), _p(fy, C.c_int8), _p(voted, C.c_float),
                             None if sad is None else _p(sad, C.c_int32))
    if rc < 0:
        raise RuntimeError(f"mof_oracle_bm failed rc={rc}")
    return (fx, fy, voted, sad) if want_sad else (fx, fy, voted)


def bm_pairs(curr: np.ndarray, prev: np.ndarray, B: int, step: int, R: int, nthreads: int = 1):
    """-> (flow[n, 2, gy, gx] int8 (x plane then y plane), voted[n, 2] float32)."""
    curr = np.ascontiguousarray(curr, dtype=np.uint8)
    prev = np.ascontiguousarray(prev, dtype=np.uint8)
    n, H, W = curr.shape
    pitch = B + step
    gx, gy = (W - 2 * R) // pitch, (H - 2 * R) // pitch
    flow = np.zeros((n, 2, gy, gx), np.int8)
    voted = np.zeros((n, 2), np.float32)
    rc = lib().mof_oracle_bm_pairs(_p(curr, C.c_uint8), _p(prev, C.c_uint8), H * W, W, H, W, B,
                                   step, R, n, _p(flow, C.c_int8), _p(voted, C.c_float), nthreads)
    if rc < 0:
        raise RuntimeError(f"mof_oracle_bm_pairs failed rc={rc}")
    return flow, voted


def logpolar(src: np.ndarray, M: float, interp: int, dst: np.ndarray | None = None) -> np.ndarray:
    """cv::logPolar(src, dst, (res/2,res/2), M, interp) restated; interp 2=CUBIC, 4=LANCZOS4."""
    src = np.ascontiguousarray(src, dtype=np.uint8)
    res = src.shape[0]
    assert src.shape == (res, res)
    if dst is None:
        dst = np.zeros((res, res), np.uint8)
    rc = lib().mof_oracle_logpolar(_p(src, C.c_uint8), res, res, float(M), interp,
                                   _p(dst, C.c_uint8), res)
    if rc:
        raise RuntimeError(f"mof_oracle_logpolar failed rc={rc}")
    return dst


def scalerot_pairs(curr: np.ndarray, prev: np.ndarray, M: float, nthreads: int = 1) -> np.ndarray:
    curr = np.ascontiguousarray(curr, dtype=np.uint8)
    prev = np.ascontiguousarray(prev, dtype=np.uint8)
    n, res, _ = curr.shape
    out = np.zeros((n, 2), np.float64)
    rc = lib().mof_oracle_scalerot_pairs(_p(curr, C.c_uint8), _p(prev, C.c_uint8), res * res, res,
                                         res, float(M), n, _p(out, C.c_double), nthreads)
    if rc:
        raise RuntimeError(f"mof_oracle_scalerot_pairs failed rc={rc}")
    return out


def interp_table(kind: int) -> np.ndarray:
    ks = 4 if kind == 2 else 8
    ptr = lib().mof_oracle_interp_table(kind)
    return np.ctypeslib.as_array(ptr, shape=(32 * 32, ks, ks)).copy()


def frontend(raw: np.ndarray, F: int, scale_factor: int = 1, cx: int | None = None) -> np.ndarray:
    """src/optic_flow.cpp:1603-1622 over [n, H, W] (MONO8) or [n, H, W, 3] (BGR8) -> [n, F, F]."""
    raw = np.ascontiguousarray(raw, dtype=np.uint8)
    n, H, W = raw.shape[:3]
    ch = 1 if raw.ndim == 3 else raw.shape[3]
    if cx is None:
        cx = (W // scale_factor) // 2
    out = np.empty((n, F, F), np.uint8)
    for k in range(n):
        rc = lib().mof_oracle_frontend(_p(raw[k], C.c_uint8), W, H, ch, W * ch, scale_factor, cx, F,
                                       _p(out[k], C.c_uint8))
        if rc:
            raise RuntimeError(f"mof_oracle_frontend failed rc={rc}")
    return out


def quarter(frames: np.ndarray) -> np.ndarray:
    """cv::resize(im, 1/4) of [n, F, F] gray frames (src/FftMethod.cpp:1931-1932)."""
    frames = np.ascontiguousarray(frames, dtype=np.uint8)
    n, F, _ = frames.shape
    out = np.empty((n, F // 4, F // 4), np.uint8)
    for k in range(n):
        if lib().mof_oracle_quarter(_p(frames[k], C.c_uint8), F, F, _p(out[k], C.c_uint8)):
            raise RuntimeError("mof_oracle_quarter failed")
    return out


def fft_pairs_longrange(curr: np.ndarray, prev: np.ndarray, N: int, max_px_speed: float = 1e9,
                        nthreads: int = 1) -> np.ndarray:
    """FftMethod::processImageLongRange CPU branch (src/FftMethod.cpp:1905-1990) -> [n, S_lr, 2]."""
    F = curr.shape[1] - (curr.shape[1] % 2)
    Neff = N if F % N == 0 else F
    sq_lr = (F // Neff) // 4
    if sq_lr == 0:
        return np.zeros((curr.shape[0], 0, 2))
    c4, p4 = quarter(curr[:, :F, :F]), quarter(prev[:, :F, :F])
    side = sq_lr * Neff
    return fft_pairs(c4[:, :side, :side], p4[:, :side, :side], Neff, max_px_speed, nthreads=nthreads)


def blockmethod(curr: np.ndarray, prev: np.ndarray, B: int, R: int):
    """Integer stage of BlockMethod::processImage (src/BlockMethod.cpp:25-75) on one F x F pair
    -> (flow_x[g,g], flow_y[g,g], voted(2,))."""
    curr = np.ascontiguousarray(curr, dtype=np.uint8)
    prev = np.ascontiguousarray(prev, dtype=np.uint8)
    F = curr.shape[0]
    g = (F - 2 * R) // B
    fx = np.zeros((g, g), np.int8)
    fy = np.zeros_like(fx)
    voted = np.zeros(2, np.float32)
    rc = lib().mof_oracle_blockmethod(_p(curr, C.c_uint8), _p(prev, C.c_uint8), F, F, B, R,
                                      fx.ctypes.data, fy.ctypes.data, voted.ctypes.data)
    if rc < 0:
        raise RuntimeError(f"mof_oracle_blockmethod failed rc={rc}")
    return fx, fy, voted


def up2(img: np.ndarray) -> np.ndarray:
    """cv::resize(img, 2x) for 8-bit gray (INTER_LINEAR), bit-exact restatement."""
    img = np.ascontiguousarray(img, dtype=np.uint8)
    h, w = img.shape
    out = np.zeros((2 * h, 2 * w), np.uint8)
    L = lib()
    L.mof_oracle_up2.restype = None
    L.mof_oracle_up2(_p(img, C.c_uint8), w, h, w, out.ctypes.data_as(C.c_void_p))
    return out


def refine(curr: np.ndarray, fullpix, passes: int = 2) -> np.ndarray:
    """BlockMethod::Refine (src/BlockMethod.cpp:96-147) on [n, h, w] frames with integer flows
    fullpix [n, 2] -> refined [n, 2] (NaN where the reference's cut-out would be empty)."""
    curr = np.ascontiguousarray(curr, dtype=np.uint8)
    if curr.ndim == 2:
        curr = curr[None]
    fullpix = np.asarray(fullpix, dtype=np.int32).reshape(-1, 2)
    n, h, w = curr.shape
    out = np.zeros((n, 2), np.float64)
    L = lib()
    L.mof_oracle_refine.restype = C.c_int
    for f in range(n):
        L.mof_oracle_refine(_p(curr[f], C.c_uint8), w, h, w, int(fullpix[f, 0]), int(fullpix[f, 1]), passes,
                            out[f].ctypes.data_as(C.c_void_p))
    return out


def point_pairs(shifts: np.ndarray, sq: int, N: int):
    """Point-pair assembly of OpticFlow::getRT / get2DT (src/optic_flow.cpp:527-550, :402-416): per frame
    the sections with finite shifts, j outer / i inner, initial = (i*N + N/2, j*N + N/2), shifted =
    initial + shift.  -> (initial [n, S, 2], shifted [n, S, 2] (NaN past the count), counts [n])."""
    shifts = np.asarray(shifts, dtype=np.float64)
    n, S, _ = shifts.shape
    ini = np.full((n, S, 2), np.nan)
    shf = np.full((n, S, 2), np.nan)
    cnt = np.zeros(n, np.int32)
    for f in range(n):
        k = 0
        for j in range(sq):
            for i in range(sq):
                sx, sy = shifts[f, i + sq * j]
                if not (np.isfinite(sx) and np.isfinite(sy)):   # :530 / :407
                    continue
                xi, yi = i * N + N // 2, j * N + N // 2          # :536-537
                ini[f, k] = (xi, yi)
                shf[f, k] = (xi + sx, yi + sy)
                k += 1
        cnt[f] = k
    return ini, shf, cnt
