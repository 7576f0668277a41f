"""GPU parity tests (-m gpu): the CUDA backend, called through the C ABI (ctypes -> libmof_b200.so),
against the CPU oracle on the same seeded inputs, against the committed golden vectors (outputs of
cv2.phaseCorrelate / cv2.logPolar, brute-force numpy for block matching), and -- at full batch
sizes -- through size-independent properties.

Tolerances (BASELINE.json north_star): FFT shifts 0.01 px per section; scale 1e-3; rotation
1e-3 rad; block matching bit-exact.  Well-conditioned sections are additionally held to 2e-3 px.
"""
import os

import numpy as np
import pytest

from mrs_optic_flow_b200 import synth

pytestmark = pytest.mark.gpu

FFT_TOL = 0.01
FFT_TIGHT = 2e-3


def _ill(curr, prev, N):
    from helpers import ill_conditioned_sections

    return ill_conditioned_sections(curr, prev, N)


@pytest.fixture(scope="module")
def api():
    import torch

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from mrs_optic_flow_b200 import api as _api

    return _api


@pytest.mark.parametrize("F,N,n", [(240, 60, 6), (240, 120, 8), (240, 240, 12), (480, 120, 3),
                                   (480, 60, 2), (120, 60, 4), (60, 60, 8)])
def test_fft_pairs_vs_oracle(api, oracle_mod, F, N, n):
    curr, prev, true = synth.shifted_pairs(n, F, N, seed=1000 + F + N)
    be = api.Backend(F, N, 80.0, max_batch=4)  # max_batch < n exercises the chunked double buffering
    got = be.fft_pairs(curr, prev)
    exp = oracle_mod.fft_pairs(curr, prev, N, max_px_speed=80.0)
    assert got.shape == exp.shape == (n, (F // N) ** 2, 2)
    assert np.array_equal(np.isnan(got), np.isnan(exp))
    err = np.abs(got - exp).max(axis=2)
    bad = _ill(curr, prev, N)
    assert np.nanmax(err) < FFT_TOL, np.nanmax(err)
    assert np.nanmax(err[~bad]) < FFT_TIGHT, np.nanmax(err[~bad])
    be.close()


@pytest.mark.parametrize("key", ["F240_N60", "F240_N120", "F240_N240", "F60_N60", "F120_N120"])
def test_fft_pairs_vs_cv2_golden(api, golden_dir, key):
    g = np.load(os.path.join(golden_dir, "fft_cv2.npz"))
    F, N = int(key.split("_")[0][1:]), int(key.split("_N")[1])
    curr, prev, exp = g[key + "_curr"], g[key + "_prev"], g[key + "_cv2"]
    be = api.Backend(F, N, 1e9, max_batch=16)
    got = be.fft_pairs(curr, prev)
    err = np.abs(got - exp).max(axis=2)
    assert np.nanmax(err) < FFT_TOL, np.nanmax(err)
    assert np.nanmax(err[~_ill(curr, prev, N)]) < FFT_TIGHT
    be.close()


def test_fft_known_answers_and_nan(api, golden_dir):
    g = np.load(os.path.join(golden_dir, "fft_cv2.npz"))
    curr, prev = g["ka_curr"], g["ka_prev"]
    be = api.Backend(120, 120, 80.0)
    got = be.fft_pairs(curr, prev).reshape(3, 2)
    assert np.allclose(got[0], 0.0, atol=1e-5)             # identical frames
    assert np.allclose(got[1], [3.0, -2.0], atol=1e-4)      # integer circular shift
    assert np.allclose(got, g["ka_cv2"].reshape(3, 2), atol=1e-4)
    be.close()
    be = api.Backend(120, 120, 20.0)                        # 50 px > max_px_speed -> NaN
    got = be.fft_pairs(curr, prev).reshape(3, 2)
    assert np.isnan(got[2]).all() and not np.isnan(got[:2]).any()
    be.close()


def test_fft_geometry_fixups_and_strides(api, oracle_mod):
    # odd frame size -> F-1; non-divisible N -> one window (src/FftMethod.cpp:1706-1716); padded rows
    curr, prev, _ = synth.shifted_pairs(3, 120, 120, seed=5)
    big_c = np.zeros((3, 130, 136), np.uint8)
    big_p = np.zeros((3, 130, 136), np.uint8)
    big_c[:, :120, :120], big_p[:, :120, :120] = curr, prev
    ref = oracle_mod.fft_pairs(curr, prev, 120, max_px_speed=80.0)
    be = api.Backend(121, 50, 80.0)
    assert (be.F, be.N, be.S) == (120, 120, 1)
    got = be.fft_pairs(big_c, big_p)  # row stride 136, frame stride 130*136
    assert np.nanmax(np.abs(got - ref)) < FFT_TIGHT
    be.close()
    # 100 = 2^2 5^2 is outside the prime-factor plan: the generic-size kernel takes it (tests/test_gpu_generic_sizes.py);
    # 98 = 2 * 7^2 is a size cv2 would zero-pad: refused loudly
    c100, p100 = np.ascontiguousarray(curr[:, :100, :100]), np.ascontiguousarray(prev[:, :100, :100])
    g = api.Backend(100, 100, 80.0)
    assert np.nanmax(np.abs(g.fft_pairs(c100, p100) - oracle_mod.fft_pairs(c100, p100, 100, max_px_speed=80.0))) < 0.01
    g.close()
    with pytest.raises(api.MofError):
        api.Backend(98, 98, 80.0).fft_pairs(curr[:, :98, :98], prev[:, :98, :98])


def test_fft_device_pointer_api_matches_host_api(api):
    import torch

    curr, prev, _ = synth.shifted_pairs(5, 240, 120, seed=77)
    be = api.Backend(240, 120, 80.0)
    host = be.fft_pairs(curr, prev)
    dc, dp = torch.from_numpy(curr).cuda(), torch.from_numpy(prev).cuda()
    dev = be.fft_pairs(dc, dp)
    torch.cuda.synchronize()
    assert np.array_equal(host, dev.cpu().numpy())  # same kernel, same data -> bitwise
    be.close()


def test_fft_hann_option(api, oracle_mod):
    curr, prev, _ = synth.shifted_pairs(4, 120, 120, seed=31)
    be = api.Backend(120, 120, 80.0, hann=True)
    got = be.fft_pairs(curr, prev)
    exp = oracle_mod.fft_pairs(curr, prev, 120, max_px_speed=80.0, hann=True)
    assert np.nanmax(np.abs(got - exp)) < FFT_TIGHT
    be.close()


def test_fft_stream_semantics(api, oracle_mod):
    """processImage statefulness: first call -> prev := curr (zero shift), then frame t against
    frame t-1 (src/FftMethod.cpp:1791-1793, :1872); setImPrev seeds the previous frame."""
    frames, vel = synth.camera_streams(3, 5, 240, 240, seed=8, max_speed=6.0)
    be = api.Backend(240, 120, 80.0, n_streams=3, max_batch=3)
    ids = np.array([2, 0, 1], np.int32)  # deliberately permuted
    outs = [be.fft_stream_push(ids, frames[ids, t]) for t in range(5)]
    assert np.allclose(outs[0], 0.0, atol=1e-5)
    for t in range(1, 5):
        exp = oracle_mod.fft_pairs(frames[ids, t], frames[ids, t - 1], 120, max_px_speed=80.0)
        assert np.nanmax(np.abs(outs[t] - exp)) < FFT_TIGHT
    # same numbers as the stateless pair API
    pair = be.fft_pairs(frames[ids, 4], frames[ids, 3])
    assert np.nanmax(np.abs(outs[4] - pair)) < 1e-6
    # reset -> first-call behaviour again; set_prev -> correlate against the seeded frame
    be.fft_stream_reset(0)
    z = be.fft_stream_push(np.array([0], np.int32), frames[0:1, 2])
    assert np.allclose(z, 0.0, atol=1e-5)
    be.fft_stream_set_prev(1, frames[1, 0])
    s = be.fft_stream_push(np.array([1], np.int32), frames[1:2, 1])
    exp = oracle_mod.fft_pairs(frames[1:2, 1], frames[1:2, 0], 120, max_px_speed=80.0)
    assert np.nanmax(np.abs(s - exp)) < FFT_TIGHT
    be.close()


def test_fft_reference_class_interface(api, oracle_mod):
    frames, _ = synth.camera_streams(1, 3, 240, 240, seed=12, max_speed=5.0)
    m = api.FftMethod(240, 120, 80.0, False, False, False, False, None, 30, "", False)
    first = m.processImage(frames[0, 0], False, False, (120, 120), 0.0, (0.0, 0.0), [])
    assert len(first) == 4 and np.allclose(first, 0.0, atol=1e-5)
    second = m.processImage(frames[0, 1])
    exp = oracle_mod.fft_pairs(frames[0, 1:2], frames[0, 0:1], 120, max_px_speed=80.0)[0]
    assert np.nanmax(np.abs(np.array(second) - exp)) < FFT_TIGHT


def test_fft_full_size_properties(api):
    """Batch-size-independent checks at a large batch: swapping the roles of the two frames negates
    the shift, permuting the batch permutes the result bitwise, integer circular shifts are
    recovered exactly."""
    n, F, N = 1024, 240, 240
    curr, prev = synth.fast_pairs(n, F, seed=3)
    be = api.Backend(F, N, 1e9, max_batch=256)
    a = be.fft_pairs(curr, prev)
    b = be.fft_pairs(prev, curr)
    assert np.nanmax(np.abs(a + b)) < 1e-3
    perm = np.random.default_rng(0).permutation(n)
    c = be.fft_pairs(curr[perm], prev[perm])
    assert np.array_equal(c, a[perm])
    # integer rolls of the tile bank (fast_pairs): shift is an integer up to noise from one row
    assert np.nanmax(np.abs(a - np.rint(a))) < 0.1
    be.close()


# ---- block matching -------------------------------------------------------------------------------
@pytest.mark.parametrize("name", ["small", "default", "grid480"])
def test_bm_bit_exact_vs_golden_and_oracle(api, oracle_mod, golden_dir, name):
    g = np.load(os.path.join(golden_dir, "bm_bruteforce.npz"))
    F, B, step, R = [int(v) for v in g[name + "_geom"]]
    curr, prev = g[name + "_curr"], g[name + "_prev"]
    be = api.Backend(F, F, 1e9, bm_block=B, bm_step=step, bm_radius=R, max_batch=2)
    flow, voted = be.bm_pairs(curr, prev)
    assert np.array_equal(flow[:, 0], g[name + "_fx"]) and np.array_equal(flow[:, 1], g[name + "_fy"])
    assert np.array_equal(voted, g[name + "_voted"])
    oflow, ovoted = oracle_mod.bm_pairs(curr, prev, B, step, R)
    assert np.array_equal(flow, oflow) and np.array_equal(voted, ovoted)
    be.close()


@pytest.mark.parametrize("F,B,step,R", [(240, 120, 24, 21), (200, 30, 10, 12), (150, 23, 5, 7),
                                        (128, 16, 0, 24)])
def test_bm_bit_exact_random_geometries(api, oracle_mod, F, B, step, R):
    curr, prev, _ = synth.bm_pairs(5, F, R, seed=F + B)
    be = api.Backend(F, F, 1e9, bm_block=B, bm_step=step, bm_radius=R, max_batch=8)
    flow, voted = be.bm_pairs(curr, prev)
    oflow, ovoted = oracle_mod.bm_pairs(curr, prev, B, step, R)
    assert np.array_equal(flow, oflow) and np.array_equal(voted, ovoted)
    be.close()


def test_bm_reference_class_interface(api, golden_dir):
    g = np.load(os.path.join(golden_dir, "bm_bruteforce.npz"))
    m = api.FastSpacedBMMethod(120, 21, 24, 0, 0, 0, 0, 0, 0, 0, 0, 0, False, None, frame_size=240)
    m.setImPrev(g["default_prev"][0])
    out = m.processImage(g["default_curr"][0], False, False, (120, 120), 0.0, (0.0, 0.0))
    assert out == [tuple(float(v) for v in g["default_voted"][0])]


# ---- scale / rotation -----------------------------------------------------------------------------
def test_logpolar_bit_exact_and_scalerot_vs_golden(api, oracle_mod, golden_dir):
    g = np.load(os.path.join(golden_dir, "scalerot_cv2.npz"))
    res, M = int(g["res"]), float(g["M"])
    be = api.Backend(res, res, 1e9, sr_resolution=res, sr_m=M, max_batch=4)
    lp = be.logpolar(g["prev"], 2)
    assert np.array_equal(lp, g["lp_cubic"])          # uint8, bit-exact vs cv2.logPolar
    lp2 = be.logpolar(g["curr"], 4, dst=lp)
    assert np.array_equal(lp2, g["lp_lanczos"])
    got = be.scalerot_pairs(g["curr"], g["prev"])
    assert np.abs(got[:, 0] - g["cv2"][:, 0]).max() < 1e-3
    assert np.abs(got[:, 1] - g["cv2"][:, 1]).max() < 1e-3
    exp = oracle_mod.scalerot_pairs(g["curr"], g["prev"], M)
    assert np.abs(got - exp).max() < 1e-4
    be.close()


def test_scalerot_stateful_class(api, oracle_mod):
    curr, prev, _ = synth.scalerot_pairs(2, 240, seed=301)
    est = api.scaleRotationEstimator(240, 49.9, False, None, 30)
    assert est.processImage(prev[0], False, False) == (1.0, 0.0)       # first call (:48, :73)
    s, r = est.processImage(curr[0])
    exp = oracle_mod.scalerot_pairs(curr[:1], prev[:1], 49.9)[0]
    assert abs(s - exp[0]) < 1e-4 and abs(r - exp[1]) < 1e-4


@pytest.mark.parametrize("N", [60, 120, 240])
def test_fft_unaligned_device_views_take_the_byte_path(api, oracle_mod, N):
    """Section origins that are not vector aligned (odd base offset, odd row stride) go through the
    byte-wise staging path of the kernel (VEC = false) and must give the same answers."""
    import torch

    F = 240
    curr, prev, _ = synth.shifted_pairs(3, F, N, seed=600 + N)
    big_c = torch.zeros((3, F + 3, F + 7), dtype=torch.uint8, device="cuda")
    big_p = torch.zeros_like(big_c)
    big_c[:, 1:F + 1, 3:F + 3] = torch.from_numpy(curr).cuda()
    big_p[:, 1:F + 1, 3:F + 3] = torch.from_numpy(prev).cuda()
    be = api.Backend(F, N, 80.0)
    got = be.fft_pairs(big_c[:, 1:F + 1, 3:F + 3], big_p[:, 1:F + 1, 3:F + 3]).cpu().numpy()
    aligned = be.fft_pairs(curr, prev)
    exp = oracle_mod.fft_pairs(curr, prev, N, max_px_speed=80.0)
    assert np.array_equal(np.isnan(got), np.isnan(exp))
    assert np.nanmax(np.abs(got - exp)) < FFT_TOL
    # both staging paths feed the same transforms; the aligned pairs path keeps the first spectrum in
    # tensor memory and is a different instantiation, so FMA contraction may differ in the last bits
    assert np.allclose(got, aligned, rtol=0, atol=1e-5, equal_nan=True), np.nanmax(np.abs(got - aligned))
    be.close()


def test_fft_empty_and_degenerate_inputs(api, oracle_mod):
    """n_pairs = 0 is a no-op.  A black frame has an exactly zero spectrum, so the correlation
    surface is exactly zero, every sample ties and cv::minMaxLoc keeps the first: shift (-N/2, -N/2)
    in cv2, in the oracle and here.  Constant non-zero frames correlate pure rounding noise (cv2 and
    the oracle already disagree with each other there): only required not to fault."""
    be = api.Backend(120, 60, 80.0, max_batch=4)
    z = np.zeros((0, 120, 120), np.uint8)
    assert be.fft_pairs(z, z).shape == (0, 4, 2)
    rng = np.random.default_rng(9)
    tex = synth.to_u8(synth.texture(rng, 120, 120))
    black = np.zeros((120, 120), np.uint8)
    curr = np.stack([black, tex, black, tex])
    prev = np.stack([black, black, tex, tex])
    got = be.fft_pairs(curr, prev)
    exp = oracle_mod.fft_pairs(curr, prev, 60, max_px_speed=80.0)
    assert np.array_equal(np.isnan(got), np.isnan(exp)), (got, exp)
    assert np.nanmax(np.abs(got - exp), initial=0.0) < FFT_TOL, (got, exp)
    assert np.all(got[:3] == -30.0)
    flat = np.stack([np.full((120, 120), 255, np.uint8), np.full((120, 120), 17, np.uint8)])
    flat2 = np.stack([np.full((120, 120), 255, np.uint8), np.full((120, 120), 200, np.uint8)])
    r = be.fft_pairs(flat, flat2)
    assert r.shape == (2, 4, 2) and np.all(np.isnan(r) | (np.abs(r) <= 30.0))
    be.close()


@pytest.mark.gpu
@pytest.mark.parametrize("N", [60, 120, 240])
def test_fft_weak_spectra_take_the_exact_cross_power(api, oracle_mod, N):
    """Bins with |F1 * conj(F2)|^2 below 1024 leave the fast normalisation and are redone with the
    exact magSpectrums/divSpectrums form (src/FftMethod.cpp:113-114).  Sparse frames make such bins:
    all of them (two single pixels), a mix (a pixel against a pixel plus weak clutter), or every bin
    exactly zero (black).  At N = 240 the previous spectrum of these redone butterflies comes back
    from tensor memory."""
    rng = np.random.default_rng(100 + N)
    F = N
    n = 6
    curr = np.zeros((n, F, F), np.uint8)
    prev = np.zeros((n, F, F), np.uint8)
    want = []
    for k in range(n):
        y, x = (int(v) for v in rng.integers(N // 4, N // 2, 2))
        dy, dx = (int(v) for v in rng.integers(-7, 8, 2))
        prev[k, y, x] = 3 + k
        curr[k, y + dy, x + dx] = 5
        if k >= 2:   # weak clutter: |F| now varies from bin to bin around the threshold
            for _ in range(5):
                yy, xx = (int(v) for v in rng.integers(0, N, 2))
                if curr[k, yy, xx] == 0:
                    curr[k, yy, xx] = 1
        want.append((dx, dy))
    be = api.Backend(F, N, 80.0, max_batch=n)
    got = be.fft_pairs(curr, prev)
    exp = oracle_mod.fft_pairs(curr, prev, N, max_px_speed=80.0)
    assert np.array_equal(np.isnan(got), np.isnan(exp)), (got, exp)
    assert np.nanmax(np.abs(got - exp), initial=0.0) < FFT_TOL, (got, exp)
    # the two clean pairs are pure translations of one pixel
    assert np.allclose(got[:2, 0], np.array(want[:2], float), atol=0.02), (got[:2], want[:2])
    be.close()
    # every bin exactly zero: all samples tie and the first wins, shift (-N/2, -N/2); the speed limit
    # (src/FftMethod.cpp:1849-1853) is lifted so that it is not turned into NaN at N >= 120
    be = api.Backend(F, N, 1000.0, max_batch=2)
    black = np.zeros((2, F, F), np.uint8)
    assert np.all(be.fft_pairs(black, black) == -N / 2)
    assert np.all(oracle_mod.fft_pairs(black, black, N, max_px_speed=1000.0) == -N / 2)
    be.close()
