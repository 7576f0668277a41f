"""BlockMethod::processImage, integer stage (src/BlockMethod.cpp:25-80): SAD over B x B blocks at
n*B + R, cv::minMaxLoc first minimum, per-axis histogram argmax.  Golden = the very cv2 calls the
reference makes (oracle/gen_golden.py blockmethod).  Bit-exact."""
import os

import numpy as np
import pytest


@pytest.fixture(scope="module")
def bm_golden(golden_dir):
    return np.load(os.path.join(golden_dir, "blockmethod_cv2.npz"))


def test_oracle_blockmethod_vs_cv2_golden(oracle_mod, bm_golden):
    F, B, R = (int(v) for v in bm_golden["geom"])
    for p in range(bm_golden["curr"].shape[0]):
        fx, fy, voted = oracle_mod.blockmethod(bm_golden["curr"][p], bm_golden["prev"][p], B, R)
        assert np.array_equal(fx, bm_golden["fx"][p]) and np.array_equal(fy, bm_golden["fy"][p])
        assert np.array_equal(voted, bm_golden["voted"][p])
    # flat pair: all candidates tie, the first one (-R, -R) wins -- no noise gate in BlockMethod
    assert np.all(bm_golden["fx"][2] == -R) and np.all(bm_golden["fy"][2] == -R)


@pytest.mark.gpu
def test_blockmethod_cuda_vs_cv2_golden(bm_golden, oracle_mod):
    import torch

    assert torch.cuda.is_available()
    from mrs_optic_flow_b200 import api

    F, B, R = (int(v) for v in bm_golden["geom"])
    be = api.Backend(F, F, 1e9, bm_block=B, bm_step=0, bm_radius=R, bm_no_gate=True)
    flow, voted = be.bm_pairs(bm_golden["curr"], bm_golden["prev"])
    assert np.array_equal(flow[:, 0], bm_golden["fx"]) and np.array_equal(flow[:, 1], bm_golden["fy"])
    assert np.array_equal(voted, bm_golden["voted"])
    be.close()
    # the reference class interface: imPrev starts black (src/BlockMethod.cpp:19-20)
    m = api.BlockMethod(F, B, R, 2 * R + 1, 0, 0)
    first = m.processImage(bm_golden["prev"][0])
    exp = oracle_mod.blockmethod(bm_golden["prev"][0], np.zeros((F, F), np.uint8), B, R)[2]
    assert m.voted == (int(exp[0]), int(exp[1]))
    assert first == [tuple(oracle_mod.refine(bm_golden["prev"][0], [m.voted])[0])]   # Refine, :78
    second = m.processImage(bm_golden["curr"][0])
    assert m.voted == tuple(int(v) for v in bm_golden["voted"][0])
    assert second == [tuple(oracle_mod.refine(bm_golden["curr"][0], [m.voted])[0])]
