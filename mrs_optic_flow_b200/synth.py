"""Seeded synthetic inputs for the hot path: textured uint8 frames with known
sub-pixel shifts, integer-shift pairs for block matching, rotated/scaled pairs for the
scale/rotation estimator and constant-velocity camera streams.  numpy only.
"""
from __future__ import annotations

import numpy as np


def _blur_fft(noise: np.ndarray, sigma: float) -> np.ndarray:
    """Circular Gaussian blur via the Fourier domain (keeps the texture periodic)."""
    h, w = noise.shape[-2:]
    fy = np.fft.fftfreq(h)[:, None]
    fx = np.fft.rfftfreq(w)[None, :]
    g = np.exp(-2.0 * (np.pi * sigma) ** 2 * (fx * fx + fy * fy))
    return np.fft.irfft2(np.fft.rfft2(noise) * g, s=(h, w))


def texture(rng: np.random.Generator, h: int, w: int, sigma: float = 1.5) -> np.ndarray:
    """White Gaussian noise -> Gaussian blur sigma px (float64, unscaled)."""
    return _blur_fft(rng.standard_normal((h, w)), sigma)


def to_u8(img: np.ndarray, lo: float | None = None, hi: float | None = None) -> np.ndarray:
    lo = img.min() if lo is None else lo
    hi = img.max() if hi is None else hi
    return np.clip(np.rint((img - lo) * (255.0 / (hi - lo))), 0, 255).astype(np.uint8)


def fourier_shift(img: np.ndarray, dx: float, dy: float) -> np.ndarray:
    """Circularly shift `img` so that its content moves by (+dx, +dy) pixels (sub-pixel)."""
    h, w = img.shape
    fy = np.fft.fftfreq(h)[:, None]
    fx = np.fft.rfftfreq(w)[None, :]
    ph = np.exp(-2j * np.pi * (fx * dx + fy * dy))
    return np.fft.irfft2(np.fft.rfft2(img) * ph, s=(h, w))


def shifted_pairs(n_pairs: int, F: int, N: int, seed: int, max_shift: float | None = None,
                  per_section: bool = True):
    """cfg1-style pairs: prev = texture, curr = texture moved by a known sub-pixel shift.

    Each N x N section gets its own texture and shift (per_section=True) so that every section is
    an independent test case.  Returns (curr[n,F,F] u8, prev[n,F,F] u8, true_shift[n,S,2]) with
    section index i + j*sq (i = column tile), matching FftMethod::processImage.
    """
    rng = np.random.default_rng(seed)
    sq = F // N
    S = sq * sq
    ms = N / 4.0 if max_shift is None else max_shift
    curr = np.zeros((n_pairs, F, F), np.uint8)
    prev = np.zeros((n_pairs, F, F), np.uint8)
    true = np.zeros((n_pairs, S, 2), np.float64)
    for p in range(n_pairs):
        if not per_section:
            t = texture(rng, F, F)
            dx, dy = rng.uniform(-ms, ms, 2)
            lo, hi = t.min(), t.max()
            prev[p] = to_u8(t, lo, hi)
            curr[p] = to_u8(fourier_shift(t, dx, dy), lo, hi)
            true[p, :, 0], true[p, :, 1] = dx, dy
            continue
        for j in range(sq):
            for i in range(sq):
                t = texture(rng, N, N)
                dx, dy = rng.uniform(-ms, ms, 2)
                lo, hi = t.min(), t.max()
                prev[p, j * N:(j + 1) * N, i * N:(i + 1) * N] = to_u8(t, lo, hi)
                curr[p, j * N:(j + 1) * N, i * N:(i + 1) * N] = to_u8(
                    fourier_shift(t, dx, dy), lo, hi)
                true[p, i + j * sq] = (dx, dy)
    return curr, prev, true


def bm_pairs(n_pairs: int, F: int, R: int, seed: int, noise_sigma: float = 2.0):
    """cfg2-style pairs: curr = np.roll(prev, integer shift in [-R, R]^2) + noise."""
    rng = np.random.default_rng(seed)
    curr = np.zeros((n_pairs, F, F), np.uint8)
    prev = np.zeros((n_pairs, F, F), np.uint8)
    true = np.zeros((n_pairs, 2), np.int32)
    for p in range(n_pairs):
        t = texture(rng, F, F)
        t = (t - t.min()) * (255.0 / (t.max() - t.min()))
        sx, sy = rng.integers(-R, R + 1, 2)
        pv = t + rng.standard_normal(t.shape) * noise_sigma
        cu = np.roll(t, (-int(sy), -int(sx)), axis=(0, 1)) + rng.standard_normal(t.shape) * noise_sigma
        prev[p] = np.clip(np.rint(pv), 0, 255).astype(np.uint8)
        curr[p] = np.clip(np.rint(cu), 0, 255).astype(np.uint8)
        true[p] = (sx, sy)
    return curr, prev, true


def _bilinear_sample(img: np.ndarray, xs: np.ndarray, ys: np.ndarray) -> np.ndarray:
    h, w = img.shape
    x0 = np.floor(xs).astype(np.int64)
    y0 = np.floor(ys).astype(np.int64)
    fx, fy = xs - x0, ys - y0
    x0c, x1c = np.clip(x0, 0, w - 1), np.clip(x0 + 1, 0, w - 1)
    y0c, y1c = np.clip(y0, 0, h - 1), np.clip(y0 + 1, 0, h - 1)
    return ((1 - fy) * ((1 - fx) * img[y0c, x0c] + fx * img[y0c, x1c])
            + fy * ((1 - fx) * img[y1c, x0c] + fx * img[y1c, x1c]))


def scalerot_pairs(n_pairs: int, res: int, seed: int, max_deg: float = 5.0,
                   max_scale: float = 0.05):
    """cfg3-style pairs: res x res crops of a 2res x 2res smooth texture; curr is the texture
    rotated by U(-max_deg, max_deg) and scaled by U(1-max_scale, 1+max_scale) about the centre."""
    rng = np.random.default_rng(seed)
    curr = np.zeros((n_pairs, res, res), np.uint8)
    prev = np.zeros((n_pairs, res, res), np.uint8)
    true = np.zeros((n_pairs, 2), np.float64)
    big = 2 * res
    yy, xx = np.mgrid[0:res, 0:res].astype(np.float64)
    c = res / 2.0
    for p in range(n_pairs):
        t = texture(rng, big, big, sigma=2.5)
        t = (t - t.min()) * (255.0 / (t.max() - t.min()))
        ang = np.deg2rad(rng.uniform(-max_deg, max_deg))
        sc = rng.uniform(1.0 - max_scale, 1.0 + max_scale)
        off = (big - res) / 2.0
        prev[p] = np.clip(np.rint(_bilinear_sample(t, xx + off, yy + off)), 0, 255).astype(np.uint8)
        ca, sa = np.cos(ang), np.sin(ang)
        dx, dy = (xx - c) / sc, (yy - c) / sc
        xs = ca * dx + sa * dy + c + off
        ys = -sa * dx + ca * dy + c + off
        curr[p] = np.clip(np.rint(_bilinear_sample(t, xs, ys)), 0, 255).astype(np.uint8)
        true[p] = (sc, ang)
    return curr, prev, true


def camera_streams(n_streams: int, n_frames: int, H: int, W: int, seed: int,
                   max_speed: float = 8.0):
    """cfg5-style streams: each stream is a big periodic texture panned at a constant sub-pixel
    velocity; returns frames[n_streams, n_frames, H, W] u8 and velocity[n_streams, 2] (px/frame)."""
    rng = np.random.default_rng(seed)
    frames = np.zeros((n_streams, n_frames, H, W), np.uint8)
    vel = np.zeros((n_streams, 2), np.float64)
    for s in range(n_streams):
        t = texture(rng, H, W)
        lo, hi = t.min(), t.max()
        vx, vy = rng.uniform(-max_speed, max_speed, 2)
        vel[s] = (vx, vy)
        for f in range(n_frames):
            frames[s, f] = to_u8(fourier_shift(t, vx * f, vy * f), lo, hi)
    return frames, vel


def fast_pairs(n_pairs: int, F: int, seed: int, max_shift: int = 12) -> tuple[np.ndarray, np.ndarray]:
    """Cheap large-batch generator for benchmarks: one textured tile bank, integer rolls plus
    per-pair noise (content does not matter for timing; parity is checked on shifted_pairs)."""
    rng = np.random.default_rng(seed)
    bank = [to_u8(texture(rng, F, F)) for _ in range(8)]
    curr = np.empty((n_pairs, F, F), np.uint8)
    prev = np.empty((n_pairs, F, F), np.uint8)
    for p in range(n_pairs):
        t = bank[p % len(bank)]
        sx, sy = rng.integers(-max_shift, max_shift + 1, 2)
        prev[p] = t
        curr[p] = np.roll(t, (int(sy), int(sx)), axis=(0, 1))
        # decorrelate pairs a little so that no two inputs are identical
        curr[p, rng.integers(0, F), :] ^= np.uint8(p & 0xFF)
    return curr, prev
