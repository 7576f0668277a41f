#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu --timeout 900 > gpurun_out/r3r_pytest.log 2>&1
echo "default build: $(grep -E 'passed|failed' gpurun_out/r3r_pytest.log | tail -n 1)"; grep -E "^(FAILED|ERROR)" gpurun_out/r3r_pytest.log | head -5 | cut -c1-200
MOF_B200_LIB=$PWD/ab_variant.so timeout 900 python -m pytest tests/test_gpu_round2.py tests/test_gpu_steady_state.py tests/test_gpu_parity.py tests/test_frontend.py -x -q -m gpu -k "stream or push or raw or cfg5" > gpurun_out/r3r_pytest_var.log 2>&1
echo "skew in every mode: $(grep -E 'passed|failed' gpurun_out/r3r_pytest_var.log | tail -n 1)"; grep -E "^(FAILED|ERROR)" gpurun_out/r3r_pytest_var.log | head -5 | cut -c1-200
{
for lib in "" "$PWD/ab_variant.so"; do
MOF_B200_LIB=$lib python - <<'PY'
import os, sys; sys.path.insert(0, '.')
import torch, bench
from mrs_optic_flow_b200 import api
dev = torch.device("cuda", 0)
tag = "B (team skew in every mode)" if os.environ.get("MOF_B200_LIB") else "A (team skew in pairs mode only)"
for f in (bench.stream_workload, bench.raw_stream_workload):
    r = f(torch, api, 0, dev)
    print(tag, r["workload"][:60], round(r["frames_per_s"]))
PY
MOF_B200_LIB=$lib timeout 300 python bench.py --workload streams --steps 5 --warmup 3 --no-e2e 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('   bench streams', d['value'], d['ms_per_step'])"
done
} > gpurun_out/r3r_teamskew_streams_ab.txt 2>&1
cat gpurun_out/r3r_teamskew_streams_ab.txt | cut -c1-200
