"""Camera front end (src/optic_flow.cpp:1603-1622) and long-range mode
(FftMethod::processImageLongRange, src/FftMethod.cpp:1905-1990).

CPU part: the oracle restatement against tests/golden/frontend_cv2.npz (outputs of cv2.resize /
cv2.cvtColor / cv2.phaseCorrelate, minted by oracle/gen_golden.py frontend).  GPU part (-m gpu):
the CUDA kernels through the C ABI against the goldens and the oracle.  Byte outputs are bit-exact;
long-range shifts within 0.01 px (BASELINE.json north_star).
"""
import os

import numpy as np
import pytest

from mrs_optic_flow_b200 import synth

FE_CASES = [(1, 101), (2, 47), (4, 25)]


@pytest.fixture(scope="module")
def fe_golden(golden_dir):
    return np.load(os.path.join(golden_dir, "frontend_cv2.npz"))


# ---- oracle vs cv2 goldens (CPU) -------------------------------------------------------------------
@pytest.mark.parametrize("k,cx", FE_CASES)
def test_oracle_frontend_vs_cv2_golden(oracle_mod, fe_golden, k, cx):
    got = oracle_mod.frontend(fe_golden["raw"], 24, k, cx)
    assert np.array_equal(got, fe_golden[f"gray_k{k}_cx{cx}"])


def test_oracle_frontend_mono(oracle_mod, fe_golden):
    mono = np.ascontiguousarray(fe_golden["raw"][..., 1])
    assert np.array_equal(oracle_mod.frontend(mono, 24, 2, 50), fe_golden["mono_k2_cx50"])


def test_oracle_longrange_vs_cv2_golden(oracle_mod, fe_golden):
    got = oracle_mod.fft_pairs_longrange(fe_golden["lr_curr"], fe_golden["lr_prev"], 60, 80.0)
    assert got.shape == fe_golden["lr_cv2"].shape == (3, 1, 2)
    assert np.abs(got - fe_golden["lr_cv2"]).max() < 2e-3
    # the down-scaled shift is a quarter of the true one (coarse check of the generator itself)
    assert np.abs(got - fe_golden["lr_true"] / 4).max() < 0.25


def test_oracle_frontend_rejects_bad_geometry(oracle_mod):
    raw = np.zeros((1, 64, 90, 3), np.uint8)
    with pytest.raises(RuntimeError):
        oracle_mod.frontend(raw, 24, 3, 15)       # ratio 3 is not supported
    with pytest.raises(RuntimeError):
        oracle_mod.frontend(raw, 24, 1, 5)        # crop leaves the frame


# ---- CUDA vs goldens / oracle (GPU) ----------------------------------------------------------------
@pytest.fixture(scope="module")
def api():
    import torch

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from mrs_optic_flow_b200 import api as _api

    return _api


@pytest.mark.gpu
@pytest.mark.parametrize("k,cx", FE_CASES)
def test_frontend_vs_cv2_golden(api, fe_golden, k, cx):
    be = api.Backend(24, 24, 80.0)
    got = be.frontend(fe_golden["raw"], k, cx)
    assert np.array_equal(got, fe_golden[f"gray_k{k}_cx{cx}"])
    mono = np.ascontiguousarray(fe_golden["raw"][..., 1])
    assert np.array_equal(be.frontend(mono, 2, 50), fe_golden["mono_k2_cx50"])
    be.close()


@pytest.mark.gpu
@pytest.mark.parametrize("k", [1, 2, 4])
def test_frontend_camera_frame_vs_oracle(api, oracle_mod, k):
    """752 x 480 BGR8 camera frames (cfg5), F = 480 / k, device-pointer entry point, odd cx."""
    import torch

    rng = np.random.default_rng(50 + k)
    raw = rng.integers(0, 256, (3, 480, 752, 3), dtype=np.uint8)
    F = 480 // k
    cx = 376 // k + 1
    be = api.Backend(F, F if F in (60, 120, 240) else 120, 80.0)
    exp = oracle_mod.frontend(raw, F, k, cx)
    got = be.frontend(torch.from_numpy(raw).cuda(), k, cx).cpu().numpy()
    assert np.array_equal(got, exp)
    assert np.array_equal(be.frontend(raw, k, cx), exp)   # host-pointer entry point
    be.close()


@pytest.mark.gpu
def test_frontend_errors(api):
    be = api.Backend(24, 24, 80.0)
    raw = np.zeros((1, 64, 90, 3), np.uint8)
    with pytest.raises(api.MofError):
        be.frontend(raw, 3, 15)
    with pytest.raises(api.MofError):
        be.frontend(raw, 1, 5)
    be.close()


@pytest.mark.gpu
def test_stream_push_raw_matches_gray_push(api, oracle_mod):
    """raw BGR frames -> front end + stateful push == oracle front end -> oracle FftMethod."""
    n_streams, F, N = 3, 240, 120
    rng = np.random.default_rng(77)
    c, p, _ = synth.shifted_pairs(n_streams, F, N, seed=78)

    def embed(g):   # gray frame -> 376 x 300 BGR frame whose crop at cx converts back to (nearly) g
        raw = rng.integers(0, 256, (g.shape[0], 300, 376, 3), dtype=np.uint8)
        raw[:, 30:30 + F, 60:60 + F, :] = g[..., None]
        return raw

    raw0, raw1 = embed(p), embed(c)
    be = api.Backend(F, N, 80.0, n_streams=n_streams, max_batch=n_streams)
    ids = np.arange(n_streams, dtype=np.int32)
    first = be.fft_stream_push_raw(ids, raw0, 1, 60 + F // 2)
    assert np.allclose(first, 0.0, atol=1e-5)   # first call: prev := curr (src/FftMethod.cpp:1791-1793)
    got = be.fft_stream_push_raw(ids, raw1, 1, 60 + F // 2)
    g0 = oracle_mod.frontend(raw0, F, 1, 60 + F // 2)
    g1 = oracle_mod.frontend(raw1, F, 1, 60 + F // 2)
    assert np.array_equal(g0, p) and np.array_equal(g1, c)   # gray of replicated channels is the identity
    exp = oracle_mod.fft_pairs(g1, g0, N, max_px_speed=80.0)
    assert np.array_equal(np.isnan(got), np.isnan(exp))
    assert np.nanmax(np.abs(got - exp)) < 0.01
    be.close()


@pytest.mark.gpu
def test_longrange_vs_cv2_golden(api, fe_golden):
    be = api.Backend(240, 60, 80.0, max_batch=2)
    assert be.longrange_geometry() == (60, 1)
    got = be.fft_pairs_longrange(fe_golden["lr_curr"], fe_golden["lr_prev"])
    assert got.shape == (3, 1, 2)
    assert np.abs(got - fe_golden["lr_cv2"]).max() < 0.01
    be.close()


@pytest.mark.gpu
def test_longrange_grid_vs_oracle(api, oracle_mod):
    """F = 480, N = 60: sqNum 8 -> sqNum_lr 2, four 60 px sections of the 120 px down-scaled frame."""
    import torch

    curr, prev, _ = synth.shifted_pairs(3, 480, 240, seed=91, max_shift=40.0)
    be = api.Backend(480, 60, 80.0, max_batch=2)
    assert be.longrange_geometry() == (120, 4)
    exp = oracle_mod.fft_pairs_longrange(curr, prev, 60, 80.0)
    got = be.fft_pairs_longrange(curr, prev)
    assert got.shape == exp.shape == (3, 4, 2)
    assert np.array_equal(np.isnan(got), np.isnan(exp))
    assert np.nanmax(np.abs(got - exp)) < 0.01
    dgot = be.fft_pairs_longrange(torch.from_numpy(curr).cuda(), torch.from_numpy(prev).cuda()).cpu().numpy()
    assert np.array_equal(np.isnan(dgot), np.isnan(exp)) and np.nanmax(np.abs(dgot - exp)) < 0.01
    be.close()


@pytest.mark.gpu
def test_longrange_small_grid_is_empty(api):
    be = api.Backend(240, 120, 80.0)   # sqNum 2 -> sqNum_lr 0: the reference returns an empty vector
    assert be.longrange_geometry() == (60, 0)
    z = np.zeros((1, 240, 240), np.uint8)
    assert be.fft_pairs_longrange(z, z).shape == (1, 0, 2)
    m = api.FftMethod(240, 120, 80.0)
    assert m.processImageLongRange(z[0]) == []
    be.close()


@pytest.mark.gpu
def test_frontend_bgr_word_path_all_alignments(api, oracle_mod):
    """Scale 1, BGR8, word-aligned rows: the kernel reads aligned words and funnel-shifts by the byte
    offset of the crop origin.  Every offset (cx mod 4), crops touching the left and the right image
    edge (the last group of the last row must not read past the buffer), a frame height equal to the
    crop (y0 = 0), and a row stride that is not a multiple of 4 (byte path) must all be bit-exact."""
    import torch

    rng = np.random.default_rng(4242)
    F = 60
    be = api.Backend(F, F, 80.0)
    for W, H in ((100, 60), (100, 75), (61, 64), (60, 60), (63, 60)):
        raw = rng.integers(0, 256, (3, H, W, 3), dtype=np.uint8)
        d_raw = torch.from_numpy(raw).cuda()
        for cx in range(F // 2, W - F // 2 + 1):
            exp = oracle_mod.frontend(raw, F, 1, cx)
            got = be.frontend(d_raw, 1, cx).cpu().numpy()
            assert np.array_equal(got, exp), (W, H, cx)
    be.close()
