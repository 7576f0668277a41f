#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_round2.py tests/test_gpu_parity.py tests/test_frontend.py tests/test_gpu_steady_state.py -m gpu -q -k "push or stream or raw or cfg5 or prev" > gpurun_out/r2y_pytest.log 2>&1
tail -n 3 gpurun_out/r2y_pytest.log | cut -c1-300
for e in "" "MOF_NO_INLINE_IDS=1"; do
env $e python - <<'PY'
import os, sys, time; sys.path.insert(0, '.')
import numpy as np, torch
from mrs_optic_flow_b200 import api
dev = torch.device("cuda", 0)
for n in (1, 8, 64):
    raws = torch.randint(0, 256, (n, 4, 480, 752, 3), dtype=torch.uint8, device=dev)
    be = api.Backend(480, 120, 80.0, n_streams=n)
    ids = np.arange(n, dtype=np.int32)
    out = torch.empty((n, be.S, 2), dtype=torch.float64, device=dev)
    for k in range(10):
        be.fft_stream_push_raw(ids, raws[:, k % 4], 1, 376, out=out)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for k in range(200):
        be.fft_stream_push_raw(ids, raws[:, k % 4], 1, 376, out=out)
        torch.cuda.synchronize()
    lat = (time.perf_counter() - t0) / 200 * 1e6
    t0 = time.perf_counter()
    for k in range(200):
        be.fft_stream_push_raw(ids, raws[:, k % 4], 1, 376, out=out)
    torch.cuda.synchronize()
    thr = (time.perf_counter() - t0) / 200 * 1e6
    print(f"inline_ids={'off' if os.environ.get('MOF_NO_INLINE_IDS') else 'on'} streams={n}: push_raw latency {lat:.1f} us (call + sync), back-to-back {thr:.1f} us per call")
    be.close()
PY
done > gpurun_out/r2y_latency.txt 2>&1
cat gpurun_out/r2y_latency.txt | cut -c1-200
