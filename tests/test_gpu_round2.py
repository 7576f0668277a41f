"""GPU tests (-m gpu) of the round-2 additions: T frames per stream in one launch
(mof_fft_stream_push_many*), the OCL-compatible numerics (mof_config.ocl_compat), and several devices
behind one handle (mof_multi_*)."""
import os

import numpy as np
import pytest

from helpers import assert_fft_parity, expand_bank
from mrs_optic_flow_b200 import synth

pytestmark = pytest.mark.gpu
NTHR = os.cpu_count() or 1


@pytest.fixture(scope="module")
def api():
    import torch

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from mrs_optic_flow_b200 import api as _api

    return _api


def _streams(n, T, F, N, seed):
    bank, _ = synth.camera_streams(8, T, F, F, seed=seed, max_speed=6.0)
    return expand_bank(bank, n, N, seed=seed + 1)          # [n, T, F, F]


@pytest.mark.parametrize("F,N,n,T", [(240, 120, 40, 9), (240, 240, 20, 7), (240, 60, 12, 6), (480, 120, 64, 16),
                                     (240, 240, 200, 12),   # 400 work items on 148 slots: chains of 6 frames, 3 per slot
                                     # remainder pieces bundled behind the whole segments (stream_push_many_impl):
                                     (240, 240, 51, 14),    # 2 segments of 6 + pieces of 2, two to a bundle, the last bundle half full
                                     (240, 240, 77, 11),    # one segment of 8 + pieces of 3, two to a bundle
                                     (480, 120, 8, 64)])    # cfg5's share of one GPU out of eight: 3 x 20 + pieces of 4, four to a bundle
def test_push_many_equals_frame_by_frame_and_oracle(api, oracle_mod, F, N, n, T):
    """One push_many call == T successive pushes (same kernels, same data: bitwise), and every frame
    matches the oracle's FftMethod loop against the previous frame (first frame: zeros).  The (stream,
    section) sequences are cut into time segments inside the launch: a segment that starts in the middle
    transforms the frame before it again, which must not change any number."""
    allf = _streams(n, T + 3, F, N, seed=300 + N + T)
    frames, more = allf[:, :T], allf[:, T:]
    ids = np.random.default_rng(T).permutation(n).astype(np.int32)
    a = api.Backend(F, N, 80.0, n_streams=n, max_batch=n * T)
    got = a.fft_stream_push_many(ids, frames[ids])
    assert got.shape == (n, T, (F // N) ** 2, 2)
    b = api.Backend(F, N, 80.0, n_streams=n, max_batch=n)
    step = np.stack([b.fft_stream_push(ids, frames[ids, t]) for t in range(T)], axis=1)
    assert np.array_equal(got, step, equal_nan=True)
    assert np.allclose(got[:, 0], 0.0, atol=1e-5)
    for t in range(1, T):
        exp = oracle_mod.fft_pairs(frames[ids, t], frames[ids, t - 1], N, max_px_speed=80.0, nthreads=NTHR)
        assert_fft_parity(oracle_mod, got[:, t], exp, frames[ids, t], frames[ids, t - 1], N, max_px_speed=80.0)
    # the state carries over: a second call continues from frame T-1, like the next single push would
    g2 = a.fft_stream_push_many(ids, more[ids])
    s2 = np.stack([b.fft_stream_push(ids, more[ids, t]) for t in range(3)], axis=1)
    assert np.array_equal(g2, s2, equal_nan=True)
    exp = oracle_mod.fft_pairs(more[ids, 0], frames[ids, T - 1], N, max_px_speed=80.0, nthreads=NTHR)
    assert_fft_parity(oracle_mod, g2[:, 0], exp, more[ids, 0], frames[ids, T - 1], N, max_px_speed=80.0)
    # chunked host path (max_batch < n*T) and the device-pointer path give the same numbers
    import torch

    c = api.Backend(F, N, 80.0, n_streams=n, max_batch=n * 2)
    assert np.array_equal(c.fft_stream_push_many(ids, frames[ids]), got, equal_nan=True)
    d = api.Backend(F, N, 80.0, n_streams=n, max_batch=1)
    dev = d.fft_stream_push_many(ids, torch.from_numpy(frames[ids]).cuda()).cpu().numpy()
    assert np.array_equal(dev, got, equal_nan=True)
    for be in (a, b, c, d):
        be.close()


def test_push_many_raw_equals_frame_by_frame(api, oracle_mod):
    """cfg5 geometry, raw BGR8 frames: push_many_raw == successive push_raw calls, bitwise."""
    n, T, W, H, F, N = 6, 5, 752, 480, 480, 120
    rng = np.random.default_rng(21)
    base = synth.to_u8(synth.texture(rng, H, W))
    raw = np.empty((n, T, H, W, 3), np.uint8)
    for k in range(n):
        for t in range(T):
            g = np.roll(base, (2 * t * (k % 3 - 1), 3 * t), axis=(0, 1)) ^ np.uint8((k + 3 * t) & 7)
            raw[k, t] = np.stack([g, g // 2 + 40, 255 - g], axis=-1)
    ids = np.arange(n, dtype=np.int32)
    a = api.Backend(F, N, 80.0, n_streams=n, max_batch=n * T)
    got = a.fft_stream_push_many_raw(ids, raw, 1, 376)
    b = api.Backend(F, N, 80.0, n_streams=n, max_batch=n)
    step = np.stack([b.fft_stream_push_raw(ids, raw[:, t], 1, 376) for t in range(T)], axis=1)
    assert np.array_equal(got, step, equal_nan=True)
    gray = np.stack([oracle_mod.frontend(raw[:, t], F, 1, 376) for t in range(T)], axis=1)
    for t in range(1, T):
        exp = oracle_mod.fft_pairs(gray[:, t], gray[:, t - 1], N, max_px_speed=80.0, nthreads=NTHR)
        assert_fft_parity(oracle_mod, got[:, t], exp, gray[:, t], gray[:, t - 1], N, max_px_speed=80.0)
    a.close()
    b.close()


@pytest.mark.parametrize("N", [60, 120, 240])
def test_push_many_weak_spectra_take_the_exact_cross_power(api, oracle_mod, N):
    """Sequences of sparse frames (single pixels, weak clutter, black): bins below the fast-normalisation
    threshold are redone with the exact magSpectrums / divSpectrums form.  At N = 240 the previous spectrum of
    a redone butterfly is the copy the thread holds in registers (tensor memory already has the current one)."""
    rng = np.random.default_rng(200 + N)
    n, T, F = 5, 6, N
    frames = np.zeros((n, T, F, F), np.uint8)
    for k in range(n):
        y, x = (int(v) for v in rng.integers(N // 3, N // 2, 2))
        for t in range(T):
            if k == 4 and t in (2, 3):
                continue                                   # black frames inside a sequence
            frames[k, t, y + t * (k - 2), x + 2 * t] = 3 + k + t
            if k >= 2:
                for _ in range(4):
                    yy, xx = (int(v) for v in rng.integers(0, N, 2))
                    if frames[k, t, yy, xx] == 0:
                        frames[k, t, yy, xx] = 1
    ids = np.arange(n, dtype=np.int32)
    a = api.Backend(F, N, 80.0, n_streams=n, max_batch=n * T)
    got = a.fft_stream_push_many(ids, frames)
    a.close()
    b = api.Backend(F, N, 80.0, n_streams=n, max_batch=n)
    step = np.stack([b.fft_stream_push(ids, frames[:, t]) for t in range(T)], axis=1)
    b.close()
    assert np.array_equal(got, step, equal_nan=True)
    for t in range(1, T):
        exp = oracle_mod.fft_pairs(frames[:, t], frames[:, t - 1], N, max_px_speed=80.0)
        assert np.array_equal(np.isnan(got[:, t]), np.isnan(exp)), (t, got[:, t], exp)
        assert np.nanmax(np.abs(got[:, t] - exp), initial=0.0) < 0.01, (t, got[:, t], exp)
    # the clean streams are pure translations of one pixel: (dx, dy) = (2, k - 2) per frame
    for k in (0, 1):
        assert np.allclose(got[k, 1:, 0], np.array([2.0, k - 2.0]), atol=0.02), got[k]


@pytest.mark.parametrize("F,N,n,T,mb", [(240, 240, 4, 20, 8), (240, 120, 5, 33, 40), (240, 60, 3, 16, 3)])
def test_push_many_host_pipeline_equals_single_chunk(api, F, N, n, T, mb):
    """mof_fft_stream_push_many with host frames cuts the sequence into chunks and copies chunk i + 1 while
    chunk i runs; the result must not depend on the chunking."""
    frames = _streams(n, T, F, N, seed=400 + N)
    ids = np.arange(n, dtype=np.int32)
    a = api.Backend(F, N, 80.0, n_streams=n, max_batch=mb)
    got = a.fft_stream_push_many(ids, frames)
    got2 = a.fft_stream_push_many(ids, frames[:, :3])
    a.close()
    b = api.Backend(F, N, 80.0, n_streams=n, max_batch=n * T)
    import torch
    exp = b.fft_stream_push_many(ids, torch.from_numpy(frames).cuda()).cpu().numpy()
    exp2 = b.fft_stream_push_many(ids, torch.from_numpy(frames[:, :3].copy()).cuda()).cpu().numpy()
    b.close()
    assert np.array_equal(got, exp, equal_nan=True)
    assert np.array_equal(got2, exp2, equal_nan=True)


@pytest.mark.gpu
@pytest.mark.parametrize("W,H,F,N,k,cx,T,full", [(752, 480, 480, 120, 1, 376, 20, False),   # 10 chunks, all rows in the window
                                                 (752, 480, 480, 120, 1, 376, 20, True),    # whole frames staged
                                                 (640, 512, 240, 120, 2, 150, 17, False),   # scale 2, rows and columns cropped
                                                 (400, 300, 240, 60, 1, 137, 3, False)])    # odd byte offset, one chunk
def test_push_many_raw_host_pipeline_equals_device_path(api, monkeypatch, W, H, F, N, k, cx, T, full):
    """mof_fft_stream_push_many_raw (host frames): only the crop window crosses PCIe, chunk i + 1 is copied
    while chunk i runs.  Must equal the device-resident call on the same frames bit for bit."""
    import torch

    if full:
        monkeypatch.setenv("MOF_FE_FULL_H2D", "1")
    n = 3
    rng = np.random.default_rng(77 + W)
    base = synth.to_u8(synth.texture(rng, H, W))
    raw = np.empty((n, T, H, W, 3), np.uint8)
    for s_ in range(n):
        for t in range(T):
            g = np.roll(base, (t * (s_ - 1), 2 * t), axis=(0, 1)) ^ np.uint8((s_ + 5 * t) & 7)
            raw[s_, t] = np.stack([g, g // 3 + 17, 255 - g], axis=-1)
    ids = np.arange(n, dtype=np.int32)
    host = api.Backend(F, N, 80.0, n_streams=n, max_batch=2 * n)
    got = host.fft_stream_push_many_raw(ids, raw, k, cx)
    got2 = host.fft_stream_push_many_raw(ids, raw[:, :2], k, cx)          # state chains across calls
    host.close()
    devb = api.Backend(F, N, 80.0, n_streams=n, max_batch=n * T)
    exp = devb.fft_stream_push_many_raw(ids, torch.from_numpy(raw).cuda(), k, cx).cpu().numpy()
    exp2 = devb.fft_stream_push_many_raw(ids, torch.from_numpy(raw[:, :2].copy()).cuda(), k, cx).cpu().numpy()
    devb.close()
    assert not np.isnan(got[:, 1:]).all()
    assert np.array_equal(got, exp, equal_nan=True)
    assert np.array_equal(got2, exp2, equal_nan=True)


@pytest.mark.parametrize("F,N,n", [(240, 60, 24), (240, 120, 500), (240, 240, 460), (480, 120, 8)])
def test_ocl_compat_vs_restated_opencl_numerics(api, oracle_mod, F, N, n):
    """mof_config.ocl_compat against the oracle's restatement of cl/FftMethod.cl (PARITY UNPINNED: no
    OpenCL device can run the reference kernel here).  Shifts inside the search radius follow the 7x7
    positive-only FP32 centroid; the batch sizes put several sections on every resident slot."""
    bank_c, bank_p, _ = synth.shifted_pairs(min(n, 32), F, N, seed=1300 + N, max_shift=min(N / 4.0, 40.0))
    both = expand_bank(np.stack([bank_c, bank_p], axis=1), n, N, seed=5)
    curr, prev = np.ascontiguousarray(both[:, 0]), np.ascontiguousarray(both[:, 1])
    be = api.Backend(F, N, 80.0, max_batch=64, ocl_compat=True)
    got = be.fft_pairs(curr, prev)
    be.close()
    exp = oracle_mod.fft_pairs(curr, prev, N, max_px_speed=80.0, nthreads=NTHR, ocl_compat=True)
    assert np.array_equal(np.isnan(got), np.isnan(exp))
    err = np.abs(got - exp).max(axis=2)
    off = np.nan_to_num(err, nan=0.0) >= 0.01
    assert off.mean() <= 0.01, (off.sum(), np.nanmax(err))      # argmax flips on near-ties only
    cpu = oracle_mod.fft_pairs(curr, prev, N, max_px_speed=80.0, nthreads=NTHR)
    assert np.nanmedian(np.abs(exp - cpu)) > 1e-3                # it really is a different estimator


def test_ocl_compat_known_answers(api, oracle_mod):
    """Identical frames -> 0; an integer circular shift inside the radius -> that shift; a shift beyond
    SEARCH_RADIUS = 55 px cannot be seen (the peak is masked away); black frames -> no positive sample."""
    N = 240
    rng = np.random.default_rng(3)
    tex = synth.to_u8(synth.texture(rng, N, N))
    curr = np.stack([tex, np.roll(tex, (-7, 12), axis=(0, 1)), np.roll(tex, (0, 80), axis=(0, 1)), np.zeros_like(tex)])
    prev = np.stack([tex, tex, tex, np.zeros_like(tex)])
    be = api.Backend(N, N, 1000.0, ocl_compat=True)
    got = be.fft_pairs(curr, prev)[:, 0]
    be.close()
    exp = oracle_mod.fft_pairs(curr, prev, N, max_px_speed=1000.0, ocl_compat=True)[:, 0]
    assert np.allclose(got[0], 0.0, atol=1e-3) and np.allclose(got[1], [12.0, -7.0], atol=1e-3)
    assert abs(got[2, 0] - 80.0) > 20.0                          # 80 px > radius 55: not found
    keep = [0, 1, 3]   # (the masked-away peak leaves an argmax among noise samples: not comparable)
    assert np.array_equal(np.isnan(got[keep]), np.isnan(exp[keep])) and np.nanmax(np.abs(got[keep] - exp[keep])) < 0.01
    cpu = api.Backend(N, N, 1000.0)
    assert abs(cpu.fft_pairs(curr[2:3], prev[2:3])[0, 0, 0] - 80.0) < 0.5   # the CPU branch does find it
    cpu.close()


@pytest.mark.parametrize("F,N", [(240, 120), (240, 240), (160, 80)])   # (160, 80): the generic-size kernel behind mof_multi
def test_multi_context_matches_single_context_bitwise(api, F, N):
    """mof_multi_* with the device list (0, 0, 0) -- three contexts with their own worker threads on one
    GPU exercise the sharding exactly as three GPUs would: contiguous pair slices, stream s on context
    s % 3 -- against one plain context, bitwise."""
    import torch

    ndev = torch.cuda.device_count()
    devices = [g % ndev for g in range(3)]
    S = (F // N) ** 2
    curr, prev, _ = synth.shifted_pairs(50, F, N, seed=9)
    one = api.Backend(F, N, 80.0, max_batch=16, n_streams=10)
    multi = api.MultiBackend(devices, F, N, 80.0, max_batch=16, n_streams=10)
    assert np.array_equal(multi.fft_pairs(curr, prev), one.fft_pairs(curr, prev), equal_nan=True)
    assert multi.fft_pairs(curr[:0], prev[:0]).shape == (0, S, 2)
    assert np.array_equal(multi.fft_pairs(curr[:2], prev[:2]), one.fft_pairs(curr[:2], prev[:2]), equal_nan=True)
    frames = _streams(10, 6, F, N, seed=90)
    ids = np.arange(10, dtype=np.int32)                          # interleaved over the contexts
    for t in range(3):
        assert np.array_equal(multi.fft_stream_push(ids, frames[:, t]), one.fft_stream_push(ids, frames[:, t]),
                              equal_nan=True)
    assert np.array_equal(multi.fft_stream_push_many(ids, frames[:, 3:]), one.fft_stream_push_many(ids, frames[:, 3:]),
                          equal_nan=True)
    grouped = np.array([0, 3, 6, 9, 1, 4, 7, 2, 5, 8], np.int32)  # ordered by owning context: zero-copy slices
    more = _streams(10, 4, F, N, seed=91)
    assert np.array_equal(multi.fft_stream_push_many(grouped, more[grouped]), one.fft_stream_push_many(grouped, more[grouped]),
                          equal_nan=True)
    odd = np.array([7, 2, 8], np.int32)                           # irregular subset
    assert np.array_equal(multi.fft_stream_push(odd, frames[odd, 0]), one.fft_stream_push(odd, frames[odd, 0]), equal_nan=True)
    with pytest.raises(api.MofError):
        multi.fft_stream_push(np.array([10], np.int32), frames[:1, 0])
    multi.close()
    one.close()


@pytest.mark.skipif(not os.environ.get("MOF_TEST_MULTI_GPU"), reason="needs >= 2 GPUs (gpurun --gpus 2); set MOF_TEST_MULTI_GPU=1")
def test_multi_gpu_matches_one_gpu_bitwise(api):
    import torch

    G = torch.cuda.device_count()
    assert G >= 2
    F, N = 240, 240
    curr, prev = synth.fast_pairs(301, F, seed=4)
    one = api.Backend(F, N, 80.0, max_batch=64)
    multi = api.MultiBackend(list(range(G)), F, N, 80.0, max_batch=64)
    assert np.array_equal(multi.fft_pairs(curr, prev), one.fft_pairs(curr, prev), equal_nan=True)
    multi.close()
    one.close()


@pytest.mark.parametrize("F,B,step,R", [(240, 120, 24, 21), (480, 120, 24, 21), (128, 16, 0, 24), (256, 30, 2, 9)])
def test_bm_tma_staging_equals_byte_staging_and_oracle(api, oracle_mod, monkeypatch, F, B, step, R):
    """FastSpacedBM with the block / search window staged by the TMA unit (cp.async.bulk.tensor, the default
    when base and strides are 16-byte aligned) and by the byte loop (MOF_BM_NO_TMA=1): bit-exact with each
    other and with the oracle, including blocks whose window box hangs over the image edge (zero fill)."""
    curr, prev, _ = synth.bm_pairs(7, F, R, seed=F + B + R)
    be = api.Backend(F, F, 1e9, bm_block=B, bm_step=step, bm_radius=R, max_batch=8)
    flow_t, voted_t = be.bm_pairs(curr, prev)
    monkeypatch.setenv("MOF_BM_NO_TMA", "1")
    flow_b, voted_b = be.bm_pairs(curr, prev)
    be.close()
    oflow, ovoted = oracle_mod.bm_pairs(curr, prev, B, step, R)
    assert np.array_equal(flow_t, flow_b) and np.array_equal(voted_t, voted_b)
    assert np.array_equal(flow_t, oflow) and np.array_equal(voted_t, ovoted)


@pytest.mark.parametrize("F,R,gate", [(240, 21, True), (480, 21, True), (240, 20, True), (384, 21, False)])
def test_bm_row_wide_kernel_equals_group_kernel_and_oracle(api, oracle_mod, monkeypatch, F, R, gate):
    """Default FastSpacedBM geometry (120 px blocks, 41 / 43 search positions per axis): the row-wide kernel
    (a thread owns one search row with all 44 candidates over 15 block rows; partial sums meet through
    shared-memory atomics) against the previous kernel (MOF_BM_NO_WIDE=1) and the oracle, bit-exact; with
    TMA and byte staging, one and nine blocks per frame, ties (flat frames) and the noise gate on and off."""
    curr, prev, _ = synth.bm_pairs(9, F, R, seed=F + R)
    curr[7] = 77                      # flat frames: every candidate ties -> first minimum (sy outer, sx inner)
    prev[7] = 77
    prev[8] = curr[8]                 # identical frames: SAD(R, R) = 0
    be = api.Backend(F, F, 1e9, bm_block=120, bm_step=24, bm_radius=R, max_batch=16, bm_no_gate=not gate)
    flow_w, voted_w = be.bm_pairs(curr, prev)
    monkeypatch.setenv("MOF_BM_NO_TMA", "1")
    flow_wb, voted_wb = be.bm_pairs(curr, prev)
    monkeypatch.delenv("MOF_BM_NO_TMA")
    monkeypatch.setenv("MOF_BM_NO_WIDE", "1")
    flow_g, voted_g = be.bm_pairs(curr, prev)
    be.close()
    assert np.array_equal(flow_w, flow_g) and np.array_equal(voted_w, voted_g)
    assert np.array_equal(flow_w, flow_wb) and np.array_equal(voted_w, voted_wb)
    if gate:
        oflow, ovoted = oracle_mod.bm_pairs(curr, prev, 120, 24, R)
        assert np.array_equal(flow_w, oflow) and np.array_equal(voted_w, ovoted)


@pytest.mark.parametrize("kind", ["fft60", "fft120", "fft240"])
def test_full_size_cv2_golden_fft(api, oracle_mod, golden_dir, kind):
    """BASELINE config 1 at SURVEY 8d's size: 256 pairs per section size against the committed outputs of
    the reference's own cv2.phaseCorrelate call sequence (tests/golden/full_cv2.npz)."""
    from helpers import full_golden

    N = int(kind[3:])
    curr, prev, exp = full_golden(golden_dir, kind)
    be = api.Backend(240, N, 80.0, max_batch=64)
    got = be.fft_pairs(curr, prev)
    be.close()
    assert_fft_parity(oracle_mod, got, exp, curr, prev, N, max_px_speed=80.0)


def test_full_size_goldens_bm_and_scalerot(api, golden_dir):
    """BASELINE configs 2 and 3 at SURVEY 8d's sizes: 256 FastSpacedBM pairs bit-exact against brute-force
    numpy, 128 scale / rotation pairs within 1e-3 / 1e-3 rad of cv2.logPolar + cv2.phaseCorrelate."""
    from helpers import full_golden

    curr, prev, exp = full_golden(golden_dir, "bm")
    be = api.Backend(240, 240, 1e9, bm_block=120, bm_step=24, bm_radius=21, max_batch=64)
    flow, voted = be.bm_pairs(curr, prev)
    be.close()
    assert np.array_equal(flow[:, :, 0, 0], exp) and np.array_equal(voted, exp.astype(np.float32))
    curr, prev, exp = full_golden(golden_dir, "scalerot")
    be = api.Backend(240, 240, 1e9, sr_resolution=240, sr_m=49.9, max_batch=32)
    got = be.scalerot_pairs(curr, prev)
    be.close()
    off = (np.abs(got[:, 0] - exp[:, 0]) >= 1e-3) | (np.abs(got[:, 1] - exp[:, 1]) >= 1e-3)
    assert off.sum() <= 1, (off.sum(), np.abs(got - exp).max(axis=0))
