"""Point pairs for OpticFlow::getRT / get2DT (src/optic_flow.cpp:527-550, :402-416): NaN filter and
section-centre geometry.  Oracle = a direct restatement of the loop; the CUDA kernel must match it
exactly (the only arithmetic is one double addition per coordinate)."""
import numpy as np
import pytest


def _case(n, sq, seed):
    rng = np.random.default_rng(seed)
    sh = rng.uniform(-20, 20, (n, sq * sq, 2))
    bad = rng.random((n, sq * sq)) < 0.3
    sh[bad] = np.nan
    sh[0, :, :] = np.nan            # a frame with no valid section
    if n > 1:
        sh[1] = rng.uniform(-5, 5, (sq * sq, 2))   # a frame with all sections valid
    if n > 2 and sq > 1:
        sh[2, 0, 0] = np.inf        # infinite is rejected like NaN (std::isfinite)
        sh[2, 1, 1] = np.nan        # one bad component rejects the section
    return sh


def test_oracle_point_pairs_known_answer(oracle_mod):
    sh = np.array([[[1.5, -2.0], [np.nan, np.nan], [0.25, 0.5], [3.0, np.nan]]])
    ini, shf, cnt = oracle_mod.point_pairs(sh, 2, 120)
    assert cnt.tolist() == [2]
    assert ini[0, :2].tolist() == [[60.0, 60.0], [60.0, 180.0]]       # sections (i=0,j=0) and (i=0,j=1)
    assert shf[0, :2].tolist() == [[61.5, 58.0], [60.25, 180.5]]
    assert np.isnan(ini[0, 2:]).all() and np.isnan(shf[0, 2:]).all()


@pytest.mark.gpu
@pytest.mark.parametrize("n,sq,N", [(5, 4, 120), (3, 8, 60), (7, 1, 240), (4, 2, 60)])
def test_point_pairs_cuda_matches_oracle(oracle_mod, n, sq, N):
    import torch

    assert torch.cuda.is_available()
    from mrs_optic_flow_b200 import api

    sh = _case(n, sq, 10 * n + sq)
    be = api.Backend(240, 120, 80.0)
    ini, shf, cnt = oracle_mod.point_pairs(sh, sq, N)
    for got in (be.point_pairs(sh, sq, N),
                tuple(t.cpu().numpy() for t in be.point_pairs(torch.from_numpy(sh).cuda(), sq, N))):
        assert np.array_equal(got[2], cnt)
        assert np.array_equal(got[0], ini, equal_nan=True) and np.array_equal(got[1], shf, equal_nan=True)
    be.close()
