"""GPU test (-m gpu) of mof_fft_pairs_noalloc_device: the pairs launch for callers that cannot allocate (its
scratch is a static __device__ array inside the library; no context).  Called through the C ABI and compared
with the existing entry point (mof_fft_pairs_device on a context) on two calls in a row.  In a file of its own
that pytest collects last."""
import ctypes as C

import numpy as np
import pytest

from mrs_optic_flow_b200 import synth

pytestmark = pytest.mark.gpu
TOL_PX = 0.01   # the tolerance of the FFT parity tests (north_star); the kernels are the same, so 0 is expected


def _noalloc(L, torch, curr, prev, F, N, max_px_speed, out):
    err = C.create_string_buffer(512)
    st = torch.cuda.current_stream(curr.device).cuda_stream
    rc = L.mof_fft_pairs_noalloc_device(curr.device.index or 0, curr.data_ptr(), prev.data_ptr(), curr.stride(0),
                                        curr.stride(1), F, N, max_px_speed, curr.shape[0], out.data_ptr(), st,
                                        err, len(err))
    return rc, err.value.decode()


@pytest.mark.parametrize("F,N,n", [(240, 240, 300), (480, 120, 70), (240, 60, 200)])
def test_noalloc_pairs_equal_context_pairs_on_two_calls_in_a_row(F, N, n):
    import torch

    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    from mrs_optic_flow_b200 import _lib, api

    L = _lib.load()
    dev = torch.device("cuda", 0)
    S = (F // N) ** 2
    be = api.Backend(F, N, 80.0, max_batch=8)
    # every buffer exists before the calls under test: they allocate nothing themselves
    outs = [torch.full((n, S, 2), -7.0, dtype=torch.float64, device=dev) for _ in range(2)]
    pairs = []
    for k in range(2):   # second call: other frames, so stale scratch contents would show
        c, p, _ = synth.shifted_pairs(n, F, N, seed=900 + 10 * k + N, max_shift=min(N / 4.0, 30.0))
        if k == 1:
            c[3] = 0          # a black frame: the exact cross-power path and NaN results
        pairs.append((torch.from_numpy(c).to(dev), torch.from_numpy(p).to(dev)))
    torch.cuda.synchronize()
    for k in range(2):
        rc, msg = _noalloc(L, torch, pairs[k][0], pairs[k][1], F, N, 80.0, outs[k])
        assert rc == 0, msg
    torch.cuda.synchronize()
    for k in range(2):
        exp = be.fft_pairs(pairs[k][0], pairs[k][1]).cpu().numpy()
        got = outs[k].cpu().numpy()
        assert np.array_equal(np.isnan(got), np.isnan(exp))
        assert np.nanmax(np.abs(got - exp), initial=0.0) < TOL_PX
        assert np.isfinite(got).mean() > 0.9
    be.close()


def test_noalloc_rejects_what_the_static_scratch_is_not_sized_for():
    import torch

    from mrs_optic_flow_b200 import _lib

    L = _lib.load()
    dev = torch.device("cuda", 0)
    z = torch.zeros((1, 480, 480), dtype=torch.uint8, device=dev)
    out = torch.zeros((1, 1, 2), dtype=torch.float64, device=dev)
    rc, msg = _noalloc(L, torch, z, z, 480, 480, 80.0, out)        # generic-size kernel: N x N complex scratch per CTA
    assert rc == 1 and "480" in msg and str(36 * 1024 * 1024) in msg
    rc, msg = _noalloc(L, torch, z, z, 480, 100, 80.0, out)        # fix-up N := F = 480: same
    assert rc == 1 and "480" in msg
    err = C.create_string_buffer(256)
    rc = L.mof_fft_pairs_noalloc_device(0, None, None, 0, 240, 240, 240, 80.0, 1, None, None, err, len(err))
    assert rc == 1 and b"bad argument" in err.value
    rc = L.mof_fft_pairs_noalloc_device(0, z.data_ptr(), z.data_ptr(), 480 * 480, 480, 240, 240, 80.0, 0, out.data_ptr(),
                                        None, None, 0)
    assert rc == 0                                                  # empty batch, no error buffer: fine
