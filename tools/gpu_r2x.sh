#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_round2.py -m gpu -q -k "push or stream" > gpurun_out/r2x_pytest.log 2>&1
tail -n 3 gpurun_out/r2x_pytest.log | cut -c1-300
python - > gpurun_out/r2x_e2e_stream240.txt 2>&1 <<'PY'
import sys; sys.path.insert(0, '.')
import torch, bench
from mrs_optic_flow_b200 import api
print(bench.stream240_e2e_workload(torch, api, 0, torch.device("cuda", 0)))
PY
cut -c1-600 gpurun_out/r2x_e2e_stream240.txt
