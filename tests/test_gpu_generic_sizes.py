"""GPU tests (-m gpu) of the section sizes outside the prime-factor plan (pg_kernels.cu): the reference's
own fallback N := frameSize (src/FftMethod.cpp:1709-1716, one 480 px section at default.yaml's frame size)
and user sizes 30 / 80 / 96 / 160.  Same bar as the fast path: 0.01 px against the oracle, cv2 as arbiter."""
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


# n chosen so that every resident CTA (148 .. 592) walks several sections
@pytest.mark.parametrize("F,N,n", [(480, 480, 460), (480, 160, 200), (480, 96, 60), (240, 80, 400),
                                   (240, 30, 40), (300, 100, 300), (288, 144, 300), (512, 256, 300),
                                   (768, 768, 150)])
def test_generic_pairs_vs_oracle(api, oracle_mod, F, N, n):
    bank_c, bank_p, _ = synth.shifted_pairs(min(n, 24), F, N, seed=4000 + N, max_shift=min(N / 4.0, 40.0))
    both = expand_bank(np.stack([bank_c, bank_p], axis=1), n, N, seed=9)
    curr, prev = np.ascontiguousarray(both[:, 0]), np.ascontiguousarray(both[:, 1])
    be = api.Backend(F, N, 80.0, max_batch=64)
    assert (be.F, be.N) == (F, N)
    got = be.fft_pairs(curr, prev)
    be.close()
    exp = oracle_mod.fft_pairs(curr, prev, N, max_px_speed=80.0, nthreads=NTHR)
    # 30 px sections: 1 % of the 5x5 centroids are ill-conditioned (every one is re-checked against cv2 itself)
    err, _ = assert_fft_parity(oracle_mod, got, exp, curr, prev, N, max_px_speed=80.0,
                               max_outlier_frac=0.03 if N < 60 else 0.005)
    assert err < 0.01


def test_generic_device_api_and_degenerate_frames(api, oracle_mod):
    import torch

    F, N, n = 480, 480, 20
    curr, prev, _ = synth.shifted_pairs(n, F, N, seed=77, max_shift=30.0)
    curr[3] = 0                      # black current frame
    prev[5] = prev[5, 0, 0]          # constant previous frame
    curr[7] = prev[7]                # identical frames: zero shift
    be = api.Backend(F, N, 80.0, max_batch=8)
    host = be.fft_pairs(curr, prev)
    dev = be.fft_pairs(torch.from_numpy(curr).cuda(), torch.from_numpy(prev).cuda()).cpu().numpy()
    be.close()
    assert np.array_equal(host, dev, equal_nan=True)
    assert np.allclose(host[7], 0.0, atol=1e-4)
    exp = oracle_mod.fft_pairs(curr, prev, N, max_px_speed=80.0, nthreads=NTHR)
    assert_fft_parity(oracle_mod, host, exp, curr, prev, N, max_px_speed=80.0, max_outlier_frac=0.2)


@pytest.mark.parametrize("F,N", [(480, 480), (240, 80)])
def test_generic_stream_semantics(api, oracle_mod, F, N):
    """first call -> zero shift, then frame t against frame t - 1; set_prev / reset; push_many == pushes."""
    frames, _ = synth.camera_streams(3, 6, F, F, seed=8 + N, max_speed=6.0)
    be = api.Backend(F, N, 80.0, n_streams=3, max_batch=3)
    ids = np.array([2, 0, 1], np.int32)
    outs = [be.fft_stream_push(ids, frames[ids, t]) for t in range(4)]
    assert np.allclose(outs[0], 0.0, atol=1e-4)
    for t in range(1, 4):
        exp = oracle_mod.fft_pairs(frames[ids, t], frames[ids, t - 1], N, max_px_speed=80.0)
        assert_fft_parity(oracle_mod, outs[t], exp, frames[ids, t], frames[ids, t - 1], N, max_px_speed=80.0)
    pair = be.fft_pairs(frames[ids, 3], frames[ids, 2])
    assert np.array_equal(outs[3], pair, equal_nan=True)
    many = be.fft_stream_push_many(ids, np.ascontiguousarray(frames[ids][:, 4:6]))     # frames 4, 5 after 3
    for k, t in enumerate((4, 5)):
        exp = oracle_mod.fft_pairs(frames[ids, t], frames[ids, t - 1], N, max_px_speed=80.0)
        assert_fft_parity(oracle_mod, many[:, k], exp, frames[ids, t], frames[ids, t - 1], N, max_px_speed=80.0)
    be.fft_stream_reset(0)
    z = be.fft_stream_push(np.array([0], np.int32), frames[0:1, 2])
    assert np.allclose(z, 0.0, atol=1e-4)
    be.fft_stream_set_prev(1, frames[1, 0])
    s = be.fft_stream_push(np.array([1], np.int32), frames[1:2, 1])
    exp = oracle_mod.fft_pairs(frames[1:2, 1], frames[1:2, 0], N, max_px_speed=80.0)
    assert_fft_parity(oracle_mod, s, exp, frames[1:2, 1], frames[1:2, 0], N, max_px_speed=80.0)
    be.close()


def test_generic_many_streams_in_one_push(api, oracle_mod):
    """more streams than ids fit in the kernel parameters (batches of 64)."""
    F, N, n = 160, 80, 150
    bank, _ = synth.camera_streams(6, 3, F, F, seed=31, max_speed=5.0)
    frames = expand_bank(bank, n, N, seed=3)
    be = api.Backend(F, N, 80.0, n_streams=n, max_batch=n)
    ids = np.random.default_rng(1).permutation(n).astype(np.int32)
    be.fft_stream_push(ids, frames[:, 0])
    got = be.fft_stream_push(ids, frames[:, 1])
    got2 = be.fft_stream_push(ids[::-1].copy(), frames[::-1, 2].copy())[::-1]
    be.close()
    exp = oracle_mod.fft_pairs(frames[:, 1], frames[:, 0], N, max_px_speed=80.0, nthreads=NTHR)
    assert_fft_parity(oracle_mod, got, exp, frames[:, 1], frames[:, 0], N, max_px_speed=80.0)
    exp2 = oracle_mod.fft_pairs(frames[:, 2], frames[:, 1], N, max_px_speed=80.0, nthreads=NTHR)
    assert_fft_parity(oracle_mod, got2, exp2, frames[:, 2], frames[:, 1], N, max_px_speed=80.0)


def test_reference_constructor_fallback_gives_one_480_section(api, oracle_mod):
    """samplePointSize = 100 does not divide frameSize = 480: the constructor sets N := 480
    (src/FftMethod.cpp:1709-1716) and processImage returns ONE shift per frame."""
    frames, _ = synth.camera_streams(1, 3, 480, 480, seed=12, max_speed=5.0)
    m = api.FftMethod(480, 100, 80.0, False, False, False, False, None, 30, "", False)
    assert (m.frameSize, m.samplePointSize, m.sqNum) == (480, 480, 1)
    first = m.processImage(frames[0, 0], False, False, (240, 240), 0.0, (0.0, 0.0), [])
    assert len(first) == 1 and np.allclose(first, 0.0, atol=1e-4)
    second = m.processImage(frames[0, 1])
    exp = oracle_mod.fft_pairs(frames[0, 1:2], frames[0, 0:1], 100, max_px_speed=80.0)[0]
    assert np.nanmax(np.abs(np.array(second) - exp)) < 0.01


def test_generic_hann_and_longrange(api, oracle_mod):
    curr, prev, _ = synth.shifted_pairs(6, 160, 80, seed=5, max_shift=12.0)
    be = api.Backend(160, 80, 80.0, hann=True)
    got = be.fft_pairs(curr, prev)
    be.close()
    exp = oracle_mod.fft_pairs(curr, prev, 80, max_px_speed=80.0, hann=True)
    assert np.nanmax(np.abs(got - exp)) < 0.01
    # long range: quarter frame, same section size (src/FftMethod.cpp:1905-1990)
    c2, p2, _ = synth.shifted_pairs(4, 640, 80, seed=6, max_shift=30.0)
    be = api.Backend(640, 80, 80.0)
    lr = be.fft_pairs_longrange(c2, p2)
    be.close()
    exp = oracle_mod.fft_pairs_longrange(c2, p2, 80, max_px_speed=80.0)
    assert lr.shape == exp.shape == (4, 4, 2)
    assert np.nanmax(np.abs(lr - exp)) < 0.01


def test_sizes_the_generic_path_does_not_cover_fail_loudly(api):
    for F, N in ((490, 70), (450, 90), (1000, 1000)):   # 7 | 70: cv2 zero-pads; 90 = 2 * 45 is fine ...
        if N == 90:
            be = api.Backend(F, N, 80.0)                # ... so it must work
            be.close()
            continue
        be = api.Backend(F, N, 80.0)
        z = np.zeros((1, F, F), np.uint8)
        with pytest.raises(Exception):
            be.fft_pairs(z, z)
        be.close()
    with pytest.raises(Exception):
        api.Backend(480, 480, 80.0, ocl_compat=True)


@pytest.mark.parametrize("res", [160, 96, 200])
def test_scalerot_at_resolutions_outside_the_plan(api, oracle_mod, res):
    """scaleRotationEstimator at a resolution the prime-factor plan does not cover: log-polar kernel as always,
    the correlation of the two log-polar images on the generic-size kernel with the estimator's tail
    (src/scaleRotationEstimator.cpp:117-126).  1e-3 scale / 1e-3 rad against the oracle."""
    curr, prev, _ = synth.scalerot_pairs(6, res, seed=60 + res)
    be = api.Backend(res, res, 1e9, sr_resolution=res, sr_m=49.9, max_batch=8)
    got = be.scalerot_pairs(curr, prev)
    be.close()
    exp = oracle_mod.scalerot_pairs(curr, prev, 49.9)
    assert np.array_equal(np.isnan(got), np.isnan(exp))
    assert float(np.nanmax(np.abs(got - exp))) < 1e-3, (got, exp)
