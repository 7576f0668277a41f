#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_round2.py -m gpu -q --timeout 300 -k "bm_tma" > gpurun_out/r2f_bm.log 2>&1
echo "bm rc $?" >> gpurun_out/r2f_bm.log
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/r2f_pytest.log 2>&1
echo "pytest rc $?" >> gpurun_out/r2f_pytest.log
{
MOF_BM_NO_TMA=1 python - <<'PY'
import sys; sys.path.insert(0, '.')
import torch, bench
from mrs_optic_flow_b200 import api
print("BM byte staging:", bench.bm_workload(torch, api, 0, torch.device("cuda", 0)))
PY
python - <<'PY'
import sys; sys.path.insert(0, '.')
import torch, bench
from mrs_optic_flow_b200 import api
print("BM TMA staging: ", bench.bm_workload(torch, api, 0, torch.device("cuda", 0)))
PY
timeout 120 python tools/prof_fft.py 240 240 8192 8 2>&1 | tail -1
} > gpurun_out/r2f_ab.txt 2>&1
tail -3 gpurun_out/r2f_bm.log; grep -E "passed|failed" gpurun_out/r2f_pytest.log | tail -2; grep -c "^FAILED" gpurun_out/r2f_pytest.log; cat gpurun_out/r2f_ab.txt | cut -c1-200
