"""BlockMethod::Refine (src/BlockMethod.cpp:96-147): 2x cv::resize + 3x3 SAD search per pass on the
up-scaled CURRENT frame.  Goldens (tests/golden/refine_cv2.npz) were made by oracle/gen_golden.py
with the cv2 calls the reference makes, statement by statement.  Bit-exact everywhere."""
import os

import numpy as np
import pytest


@pytest.fixture(scope="module")
def rf_golden(golden_dir):
    return np.load(os.path.join(golden_dir, "refine_cv2.npz"))


def test_up2_oracle_matches_cv2_resize(rf_golden, oracle_mod):
    for k in range(3):
        assert np.array_equal(oracle_mod.up2(rf_golden[f"up_in{k}"]), rf_golden[f"up_out{k}"])


def test_refine_oracle_matches_cv2_golden(rf_golden, oracle_mod):
    got = oracle_mod.refine(rf_golden["frames"], rf_golden["full"])
    assert np.array_equal(got, rf_golden["refined"])
    # non-negative offsets compare the image with itself at zero displacement: unchanged
    assert np.array_equal(got[1], [3.0, 2.0]) and np.array_equal(got[6], [7.0, 0.0])
    # an offset that leaves no cut-out: the reference's cv::Rect would throw -> NaN
    assert np.isnan(oracle_mod.refine(rf_golden["frames"][:1], [[60, 0]])).all()


@pytest.mark.gpu
def test_refine_cuda_bit_exact(rf_golden, oracle_mod):
    import torch

    assert torch.cuda.is_available()
    from mrs_optic_flow_b200 import api

    frames, full = rf_golden["frames"], rf_golden["full"]
    be = api.Backend(96, 96, 1e9, bm_block=16, bm_step=0, bm_radius=8, bm_no_gate=True)
    got = be.bm_refine(frames, full)
    assert np.array_equal(got, rf_golden["refined"])
    # device pointers, padded rows, odd width (byte-store path of the up-scale), three passes
    rng = np.random.default_rng(77)
    odd = rng.integers(0, 256, (5, 61, 75), dtype=np.uint8)
    fo = rng.integers(-6, 7, (5, 2)).astype(np.int32)
    big = torch.zeros((5, 64, 80), dtype=torch.uint8, device="cuda")
    big[:, :61, :75] = torch.from_numpy(odd).cuda()
    for passes in (1, 2, 3):
        dev = be.bm_refine(big[:, :61, :75], torch.from_numpy(fo), passes)
        assert np.array_equal(dev.cpu().numpy(), oracle_mod.refine(odd, fo, passes))
    # empty batch, and an offset without a cut-out
    assert be.bm_refine(frames[:0], full[:0]).shape == (0, 2)
    assert np.isnan(be.bm_refine(frames[:1], np.array([[60, 0]], np.int32))).all()
    # larger frames (the default 240 px frame with |flow| up to the scan radius)
    big_f = rng.integers(0, 256, (6, 240, 240), dtype=np.uint8)
    bf = rng.integers(-21, 22, (6, 2)).astype(np.int32)
    assert np.array_equal(be.bm_refine(big_f, bf), oracle_mod.refine(big_f, bf))
    be.close()
