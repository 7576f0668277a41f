#!/bin/bash
mkdir -p gpurun_out
{
for lib in "" "$PWD/ab_variant.so"; do
MOF_B200_LIB=$lib python - <<'PY'
import os, sys; sys.path.insert(0, '.')
import torch, bench
from mrs_optic_flow_b200 import api
dev = torch.device("cuda", 0)
tag = "FAKE no spectrum traffic" if os.environ.get("MOF_B200_LIB") else "real"
import numpy as np
# cfg5 correlation part only: 64 streams x 64 gray frames 480^2, N=120, push_many
n, T, F, N = 64, 64, 480, 120
hc, hp = bench.make_inputs_host(64, F, seed=9)
da, _ = bench.device_batch(torch, hc, hp, n * T, dev)
frames = da.view(n, T, F, F)
be = api.Backend(F, N, 80.0, n_streams=n)
ids = np.arange(n, dtype=np.int32)
out = torch.empty((n, T, be.S, 2), dtype=torch.float64, device=dev)
ms = bench.timed_device_loop(torch, lambda: be.fft_stream_push_many(ids, frames, out=out), 6, 3, lambda: None)
print(tag, "cfg5 gray push_many N=120:", n * T * 6 / (ms * 1e-3), "frames/s")
be.close()
print(tag, bench.stream240_many_workload(torch, api, 0, dev))
print(tag, bench.stream240_workload(torch, api, 0, dev))
PY
done
} > gpurun_out/r3f_fake_nospec.txt 2>&1
cut -c1-260 gpurun_out/r3f_fake_nospec.txt
