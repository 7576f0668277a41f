#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_round2.py tests/test_gpu_steady_state.py -m gpu -q -k "push_many" > gpurun_out/r3g_pytest.log 2>&1
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/r3g_pytest.log | cut -c1-200
{
for e in "" "MOF_NO_TMEM=1"; do
env $e python - <<'PY'
import os, sys; sys.path.insert(0, '.')
import torch, bench
from mrs_optic_flow_b200 import api
dev = torch.device("cuda", 0)
tag = "scratch chain (MOF_NO_TMEM=1)" if os.environ.get("MOF_NO_TMEM") else "spectrum chain in tensor memory"
print(tag, bench.stream240_many_workload(torch, api, 0, dev))
PY
done
python tools/prof_fft.py 240 240 8192 10 2>&1 | tail -n 1
} > gpurun_out/r3g_seqtm.txt 2>&1
cut -c1-260 gpurun_out/r3g_seqtm.txt
