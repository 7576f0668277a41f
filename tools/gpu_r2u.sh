#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_frontend.py tests/test_gpu_round2.py tests/test_gpu_steady_state.py -m gpu -q -k "frontend or raw or cfg5" > gpurun_out/r2u_pytest.log 2>&1
{
python - <<'PY'
import sys; sys.path.insert(0, '.')
import torch, bench
from mrs_optic_flow_b200 import api
print("A 16 px/thread:", bench.raw_stream_workload(torch, api, 0, torch.device("cuda", 0)))
PY
MOF_FE_NO_BGR16=1 python - <<'PY'
import sys; sys.path.insert(0, '.')
import torch, bench
from mrs_optic_flow_b200 import api
print("B 4 px/thread: ", bench.raw_stream_workload(torch, api, 0, torch.device("cuda", 0)))
PY
} > gpurun_out/r2u_ab.txt 2>&1
timeout 300 python bench.py --workload streams --steps 5 --warmup 3 > gpurun_out/r2u_bench_streams_1gpu.json 2>/dev/null
MOF_FE_NO_BGR16=1 timeout 300 python bench.py --workload streams --steps 5 --warmup 3 > gpurun_out/r2u_bench_streams_1gpu_old.json 2>/dev/null
tail -n 1 gpurun_out/r2u_pytest.log; cut -c1-330 gpurun_out/r2u_ab.txt; python -c "
import json
for f in ('r2u_bench_streams_1gpu','r2u_bench_streams_1gpu_old'):
    d=json.load(open('gpurun_out/'+f+'.json')); print(f, d['value'], d['ms_per_step'])"
