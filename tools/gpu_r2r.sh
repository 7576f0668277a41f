#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/r2r_pytest.log 2>&1
{
for cfg in "120 480 2368" "120 240 8192"; do
  echo "A (team staging buffer, aux warp fetches one image ahead): $(timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
  echo "B (staging inside the section array, fetch at the row stage): $(MOF_B200_LIB=$PWD/ab_noxstage.so timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
done
python - <<'PY'
import sys; sys.path.insert(0, '.')
import torch, bench
from mrs_optic_flow_b200 import api
print("A streams:", bench.stream_workload(torch, api, 0, torch.device("cuda", 0)))
PY
MOF_B200_LIB=$PWD/ab_noxstage.so python - <<'PY'
import sys; sys.path.insert(0, '.')
import torch, bench
from mrs_optic_flow_b200 import api
print("B streams:", bench.stream_workload(torch, api, 0, torch.device("cuda", 0)))
PY
} > gpurun_out/r2r_ab.txt 2>&1
timeout 300 python bench.py --workload streams --steps 5 --warmup 3 > gpurun_out/r2r_bench_streams_1gpu.json 2>/dev/null
grep -E "passed|failed" gpurun_out/r2r_pytest.log | tail -n 1; grep "^FAILED" gpurun_out/r2r_pytest.log | head -5; cut -c1-200 gpurun_out/r2r_ab.txt; python -c "
import json; d=json.load(open('gpurun_out/r2r_bench_streams_1gpu.json')); print(d['value'], d['ms_per_step'])"
