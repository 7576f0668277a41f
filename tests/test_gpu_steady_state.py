"""GPU parity tests (-m gpu) of the persistent kernels in their STEADY STATE and at BASELINE sizes.

pc_pfa_kernel is a persistent grid-stride kernel: from the second section of a resident slot on it
carries state the first iteration does not have (the aux warp's deferred centroid of the previous
section, taps parked while the main warps already stage the next section's pixels, tensor-memory
slots reused, L2 prefetch of the next image, the lock-step redo of the last section).  Resident
slots on a B200 are 148 (N = 240), 444 (N = 120) and 2220 (N = 60); every test here runs >= 3x
that many sections through ONE launch and compares every section with the CPU oracle at the
path's own tolerance (0.01 px, equal NaN pattern), with black / one-pixel (weak spectrum -> exact
cross-power redo) / over-speed (NaN) frames mixed in so that the slow path and the deferred
centroid interleave with ordinary sections.  BASELINE configs 4 and 5 run at their full size.

Sections whose correlation surface has two top samples within 1e-4 of each other flip their integer
peak on a last-bit difference (in cv2 as well); helpers.assert_fft_parity accepts a miss only after
checking that on the oracle's own surface, and bounds their number.
"""
import os

import numpy as np
import pytest

from helpers import assert_fft_parity, expand_bank
from mrs_optic_flow_b200 import synth

pytestmark = pytest.mark.gpu

NTHR = os.cpu_count() or 1
# (F, N, pairs): sections = pairs * (F/N)^2 >= 3 x resident slots (148 / 444 / 2220 on 148 SMs)
STEADY = {240: (240, 240, 520), 120: (240, 120, 360), 60: (240, 60, 448)}
_cache = {}


@pytest.fixture(scope="module")
def api():
    import torch

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from mrs_optic_flow_b200 import api as _api

    return _api


def mixed_pairs(N):
    """Steady-state batch for section size N with degenerate pairs mixed in at scattered positions."""
    if N in _cache:
        return _cache[N]
    F, _, n = STEADY[N]
    bank_c, bank_p, _ = synth.shifted_pairs(48, F, N, seed=7000 + N)
    both = expand_bank(np.stack([bank_c, bank_p], axis=1), n, N, seed=N)
    curr, prev = np.ascontiguousarray(both[:, 0]), np.ascontiguousarray(both[:, 1])
    rng = np.random.default_rng(50 + N)
    spots = rng.choice(n, size=40, replace=False)
    for t, k in enumerate(spots):
        kind = t % 5
        if kind == 0:                       # both black: every sample ties, first wins
            curr[k] = 0
            prev[k] = 0
        elif kind == 1:                     # black previous frame: zero cross-power
            prev[k] = 0
        elif kind == 2:                     # one bright pixel each: every |P|^2 is tiny -> exact redo
            curr[k] = 0
            prev[k] = 0
            for j in range(F // N):
                for i in range(F // N):
                    y, x = (int(v) for v in rng.integers(N // 4, N // 2, 2))
                    prev[k, j * N + y, i * N + x] = 3 + t % 7
                    curr[k, j * N + y + int(rng.integers(-5, 6)), i * N + x + int(rng.integers(-5, 6))] = 5
        elif kind == 3:                     # weak clutter against a normal frame: fast and exact bins mixed
            curr[k] = 0
            for _ in range(12):
                yy, xx = (int(v) for v in rng.integers(0, F, 2))
                curr[k, yy, xx] = 1 + int(rng.integers(0, 3))
        else:                               # half-section circular shift of every section: over the speed limit
            t_ = prev[k].reshape(F // N, N, F // N, N)
            curr[k] = np.roll(t_, (N // 2 - 3, N // 2 - 2), axis=(1, 3)).reshape(F, F)
    _cache[N] = (curr, prev)
    return curr, prev


def oracle_pairs(oracle_mod, curr, prev, N, mps):
    key = ("exp", N, mps, curr.shape)
    if key not in _cache:
        _cache[key] = oracle_mod.fft_pairs(curr, prev, N, max_px_speed=mps, nthreads=NTHR)
    return _cache[key]


@pytest.mark.parametrize("N,tmem", [(240, True), (240, False), (120, True), (120, False), (60, False)])
def test_pairs_steady_state_vs_oracle(api, oracle_mod, N, tmem, monkeypatch):
    """Pairs mode, one launch, >= 3 sections per resident slot; with the first spectrum in tensor memory
    and (MOF_NO_TMEM=1) through the L2 scratch."""
    import torch

    F, _, n = STEADY[N]
    mps = 40.0   # < N/2 * sqrt(2) for every N: the all-tie (-N/2, -N/2) sections become NaN like the over-speed ones
    curr, prev = mixed_pairs(N)
    exp = oracle_pairs(oracle_mod, curr, prev, N, mps)
    if not tmem:
        monkeypatch.setenv("MOF_NO_TMEM", "1")
    be = api.Backend(F, N, mps, max_batch=n)
    dc, dp = torch.from_numpy(curr).cuda(), torch.from_numpy(prev).cuda()
    got = be.fft_pairs(dc, dp).cpu().numpy()           # device API: all n * S sections in one launch
    worst, wl = assert_fft_parity(oracle_mod, got, exp, curr, prev, N, max_px_speed=mps)
    assert np.isnan(exp).any() and np.isfinite(exp).any()
    got_h = be.fft_pairs(curr, prev)                   # host API, one chunk
    assert np.array_equal(got_h, got, equal_nan=True)
    be.close()
    print(f"N={N} tmem={tmem}: {n * (F // N) ** 2} sections, max err {worst:.2e} px, {wl} near-tie sections")


@pytest.mark.parametrize("N", [240, 120, 60])
def test_pairs_steady_state_unaligned_byte_path(api, oracle_mod, N):
    """The same batches through views whose section origins are not vector aligned (VEC = false)."""
    import torch

    F, _, n = STEADY[N]
    mps = 40.0
    curr, prev = mixed_pairs(N)
    exp = oracle_pairs(oracle_mod, curr, prev, N, mps)
    big_c = torch.zeros((n, F + 3, F + 7), dtype=torch.uint8, device="cuda")
    big_p = torch.zeros_like(big_c)
    big_c[:, 1:F + 1, 3:F + 3] = torch.from_numpy(curr).cuda()
    big_p[:, 1:F + 1, 3:F + 3] = torch.from_numpy(prev).cuda()
    be = api.Backend(F, N, mps)
    got = be.fft_pairs(big_c[:, 1:F + 1, 3:F + 3], big_p[:, 1:F + 1, 3:F + 3]).cpu().numpy()
    assert_fft_parity(oracle_mod, got, exp, curr, prev, N, max_px_speed=mps)
    be.close()


@pytest.mark.parametrize("N", [240, 120, 60])
def test_stream_mode_steady_state_vs_oracle(api, oracle_mod, N):
    """Stateful mode: >= 3 x resident slots of sections per push, 4 pushes (first call zeros, then frame
    t against frame t-1 frame by frame, src/FftMethod.cpp:1791-1793, :1872), with a reset and a
    set_prev in the middle and a black frame on some streams."""
    F, _, n = STEADY[N]
    T = 4
    bank, _ = synth.camera_streams(24, T, F, F, seed=900 + N, max_speed=6.0)
    frames = expand_bank(bank, n, N, seed=2 * N + 1)          # [n, T, F, F]
    frames[5::37, 2] = 0                                      # a black frame in the middle of some streams
    be = api.Backend(F, N, 80.0, n_streams=n, max_batch=n)
    ids = np.random.default_rng(N).permutation(n).astype(np.int32)
    out0 = be.fft_stream_push(ids, frames[ids, 0])
    assert np.all(out0 == 0.0) or np.allclose(out0, 0.0, atol=1e-5)
    for t in range(1, T):
        got = be.fft_stream_push(ids, frames[ids, t])
        exp = oracle_mod.fft_pairs(frames[ids, t], frames[ids, t - 1], N, max_px_speed=80.0, nthreads=NTHR)
        assert_fft_parity(oracle_mod, got, exp, frames[ids, t], frames[ids, t - 1], N, max_px_speed=80.0)
        if t == 1:   # stream ids[0] starts over; stream ids[1] gets an explicit previous frame
            be.fft_stream_reset(int(ids[0]))
            be.fft_stream_set_prev(int(ids[1]), frames[ids[1], 0])
            sub = ids[:2].copy()
            g2 = be.fft_stream_push(sub, frames[sub, 3])
            assert np.allclose(g2[0], 0.0, atol=1e-5)
            e2 = oracle_mod.fft_pairs(frames[sub[1:], 3], frames[sub[1:], 0], N, max_px_speed=80.0)
            assert_fft_parity(oracle_mod, g2[1:], e2, frames[sub[1:], 3], frames[sub[1:], 0], N, max_px_speed=80.0)
            # put both streams back on the common timeline
            for s_ in sub:
                be.fft_stream_set_prev(int(s_), frames[s_, 1])
    be.close()


def test_cfg4_full_size_vs_oracle(api, oracle_mod):
    """BASELINE config 4 in full: 4096 independent pairs of 480 x 480 with 8 x 8 sections of 60 px --
    262 144 sections in one launch (118 per resident slot), every one compared with the oracle."""
    import torch

    n, F, N = 4096, 480, 60
    bank_c, bank_p, _ = synth.shifted_pairs(24, F, N, seed=4)
    both = expand_bank(np.stack([bank_c, bank_p], axis=1), n, N, seed=44)
    curr, prev = np.ascontiguousarray(both[:, 0]), np.ascontiguousarray(both[:, 1])
    del both
    be = api.Backend(F, N, 80.0)
    got = be.fft_pairs(torch.from_numpy(curr).cuda(), torch.from_numpy(prev).cuda()).cpu().numpy()
    be.close()
    exp = oracle_mod.fft_pairs(curr, prev, N, max_px_speed=80.0, nthreads=NTHR)
    worst, wl = assert_fft_parity(oracle_mod, got, exp, curr, prev, N, max_px_speed=80.0, max_outlier_frac=0.002)
    print(f"cfg4: {n * 64} sections, max err {worst:.2e} px, {wl} near-tie sections")


def test_cfg5_full_size_raw_streams_vs_oracle(api, oracle_mod):
    """BASELINE config 5 in full on one GPU: 64 camera streams x 64 frames of 752 x 480 BGR8 through
    mof_fft_stream_push_raw (front end: crop 480^2 at cx = 376, RGB2GRAY; then the stateful push,
    N = 120) against the oracle's front end + FftMethod loop, frame by frame."""
    n_streams, T, W, H, F, N = 64, 64, 752, 480, 480, 120
    rng = np.random.default_rng(5)
    base = np.stack([synth.to_u8(synth.texture(rng, H, W)) for _ in range(8)])
    vel = rng.integers(-8, 9, (n_streams, 2))
    gains = np.array([0.9, 1.0, 0.8])
    be = api.Backend(F, N, 80.0, n_streams=n_streams, max_batch=n_streams)
    ids = np.arange(n_streams, dtype=np.int32)
    prev_gray = None
    worst = 0.0
    for t in range(T):
        raw = np.empty((n_streams, H, W, 3), np.uint8)
        for s_ in range(n_streams):
            g = np.roll(base[s_ % 8], (int(vel[s_, 1]) * t, int(vel[s_, 0]) * t), axis=(0, 1))
            g = g ^ np.uint8((s_ * 5 + t) & 0x0F)            # per-stream, per-frame low-bit noise
            raw[s_] = np.clip(g[..., None] * gains, 0, 255).astype(np.uint8)
        got = be.fft_stream_push_raw(ids, raw, 1, 376)
        gray = oracle_mod.frontend(raw, F, 1, 376)
        if t == 0:
            assert np.allclose(got, 0.0, atol=1e-5)
        else:
            exp = oracle_mod.fft_pairs(gray, prev_gray, N, max_px_speed=80.0, nthreads=NTHR)
            w, _ = assert_fft_parity(oracle_mod, got, exp, gray, prev_gray, N, max_px_speed=80.0)
            worst = max(worst, w)
        prev_gray = gray
    be.close()
    print(f"cfg5: {n_streams} streams x {T} frames, max err {worst:.2e} px")


def test_longrange_steady_state_vs_oracle(api, oracle_mod):
    """Long-range mode (src/FftMethod.cpp:1905-1990) with 1800 pairs of 480 x 480: 7200 sections of
    60 px after the 1/4 down-scale, > 3 per resident slot."""
    import torch

    n, F, N = 1800, 480, 60
    bank_c, bank_p, _ = synth.shifted_pairs(12, F, 240, seed=61, max_shift=20.0)
    both = expand_bank(np.stack([bank_c, bank_p], axis=1), n, 240, seed=62)
    curr, prev = np.ascontiguousarray(both[:, 0]), np.ascontiguousarray(both[:, 1])
    be = api.Backend(F, N, 80.0)
    got = be.fft_pairs_longrange(torch.from_numpy(curr).cuda(), torch.from_numpy(prev).cuda()).cpu().numpy()
    be.close()
    exp = oracle_mod.fft_pairs_longrange(curr, prev, N, max_px_speed=80.0, nthreads=NTHR)
    c4, p4 = oracle_mod.quarter(curr), oracle_mod.quarter(prev)
    assert_fft_parity(oracle_mod, got, exp, c4, p4, N, max_px_speed=80.0)


def test_scalerot_steady_state_vs_oracle(api, oracle_mod):
    """scaleRotationEstimator pairs at 456 pairs: lp_kernel's 148 persistent CTAs and the N = 240
    correlation behind it both iterate >= 3 times.  Tolerances 1e-3 (scale) / 1e-3 rad."""
    import torch

    n, res = 456, 240
    bc, bp, _ = synth.scalerot_pairs(24, res, seed=33)
    curr, prev = np.resize(bc, (n, res, res)).copy(), np.resize(bp, (n, res, res)).copy()
    for k in range(n):   # make every pair distinct
        curr[k, (k * 7) % res, :] ^= np.uint8((k * 13 + 1) & 0xFF)
    be = api.Backend(res, res, 1e9, sr_resolution=res, sr_m=49.9, max_batch=n)
    got = be.scalerot_pairs(torch.from_numpy(curr).cuda(), torch.from_numpy(prev).cuda()).cpu().numpy()
    be.close()
    exp = oracle_mod.scalerot_pairs(curr, prev, 49.9, nthreads=NTHR)
    err = np.abs(got - exp)
    off = (err[:, 0] >= 1e-3) | (err[:, 1] >= 1e-3)
    assert off.sum() <= 2, (off.sum(), err.max(axis=0))   # near-tie surfaces only (log-polar images are bit-exact)
