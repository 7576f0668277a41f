"""The reference's CPU path through the very third-party entry point it calls.  TEST / BASELINE
INFRASTRUCTURE ONLY (same import rule as oracle.py: tests/, smoke() and bench.py's cpu_baseline /
--impl reference legs).

FftMethod::processImage with useOCL=false is, per section,
    shift = -cv::phaseCorrelate(imCurrF(roi), imPrevF(roi))          src/FftMethod.cpp:1836
after two whole-frame convertTo(CV_32FC1) (:1805-1806), followed by the validity rules of
:1840-1856.  The reference cannot be compiled in this image (ROS + OpenCV C++ headers are absent),
but `cv2` 4.13 is the same library one binding away, so this module runs exactly that call sequence:
whole-frame astype(float32), ROI *views* (no copies), cv2.phaseCorrelate, negate, NaN rules.

The reference is single-threaded per frame; a box with C cores can serve C cameras, so the timed
baseline runs one worker process per host core (fork, cv2.setNumThreads(1) in each), every worker
taking a contiguous slice of the frame pairs.
"""
from __future__ import annotations

import multiprocessing as mp
import os

import numpy as np

_G = {}


def available() -> bool:
    try:
        import cv2  # noqa: F401

        return True
    except Exception:
        return False


def version() -> str:
    import cv2

    return cv2.__version__


def fft_method_frame(cv2, curr: np.ndarray, prev: np.ndarray, N: int, max_px_speed: float, out: np.ndarray):
    """One FftMethod::processImage call (CPU branch) on one frame pair -> out[S, 2]."""
    F = curr.shape[0] - (curr.shape[0] % 2)                 # :1706-1708
    if F % N:
        N = F                                               # :1709-1716
    sq = F // N
    cf, pf = curr.astype(np.float32), prev.astype(np.float32)   # convertTo, :1805-1806
    lim = max_px_speed * max_px_speed
    for j in range(sq):
        for i in range(sq):
            (x, y), _ = cv2.phaseCorrelate(cf[j * N:(j + 1) * N, i * N:(i + 1) * N],
                                           pf[j * N:(j + 1) * N, i * N:(i + 1) * N])
            sx, sy = -x, -y                                 # :1836
            ok = not (sx * sx + sy * sy > lim or abs(sx) > N / 2 or abs(sy) > N / 2 or sx != sx or sy != sy)
            out[i + j * sq, 0] = sx if ok else np.nan       # :1840-1856
            out[i + j * sq, 1] = sy if ok else np.nan


def _work(args):
    import cv2

    cv2.setNumThreads(1)
    lo, hi = args
    curr, prev, N, mps = _G["curr"], _G["prev"], _G["N"], _G["mps"]
    F = curr.shape[1] - (curr.shape[1] % 2)
    Ne = N if F % N == 0 else F
    out = np.empty((hi - lo, (F // Ne) ** 2, 2), np.float64)
    for k in range(lo, hi):
        fft_method_frame(cv2, curr[k], prev[k], N, mps, out[k - lo])
    return lo, out


class Pool:
    """Worker processes forked with the inputs in place (no per-step pickling of the frames)."""

    def __init__(self, curr: np.ndarray, prev: np.ndarray, N: int, max_px_speed: float, nproc: int | None = None):
        self.nproc = nproc or os.cpu_count() or 1
        _G.update(curr=curr, prev=prev, N=N, mps=float(max_px_speed))
        self.n = curr.shape[0]
        self.pool = mp.get_context("fork").Pool(self.nproc) if self.nproc > 1 else None
        per = (self.n + self.nproc - 1) // self.nproc
        self.slices = [(lo, min(self.n, lo + per)) for lo in range(0, self.n, per)]

    def run(self) -> np.ndarray:
        parts = self.pool.map(_work, self.slices) if self.pool else [_work(s) for s in self.slices]
        return np.concatenate([p for _, p in sorted(parts, key=lambda t: t[0])], axis=0)

    def close(self):
        if self.pool:
            self.pool.close()
            self.pool.join()
            self.pool = None


def fft_pairs(curr: np.ndarray, prev: np.ndarray, N: int, max_px_speed: float = 1e9, nproc: int = 1) -> np.ndarray:
    """[n, F, F] uint8 stacks -> [n, S, 2], the reference CPU branch through cv2."""
    p = Pool(np.ascontiguousarray(curr), np.ascontiguousarray(prev), N, max_px_speed, nproc)
    try:
        return p.run()
    finally:
        p.close()
