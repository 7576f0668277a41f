#!/bin/bash
# fused Refine after the parallel final reduction: tests, A/B, then one ncu capture with source counters
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu --timeout 600 -k "refine or Refine or blockmethod or BlockMethod" 2>&1 | grep -E "passed|failed" | tail -n 1
MOF_RF_NO_FUSED=1 timeout 600 python -m pytest tests -x -q -m gpu --timeout 600 -k "refine or Refine" 2>&1 | grep -E "passed|failed" | tail -n 1
cat > /tmp/prof_rf.py <<'PY'
import os, sys
sys.path.insert(0, os.getcwd())
import torch, bench
from mrs_optic_flow_b200 import api
print(bench.refine_workload(torch, api, 0, torch.device("cuda", 0)))
PY
for i in 1 2; do
echo "fused:    $(timeout 200 python /tmp/prof_rf.py 2>&1 | tail -1)"
echo "unfused:  $(MOF_RF_NO_FUSED=1 timeout 200 python /tmp/prof_rf.py 2>&1 | tail -1)"; echo "no strips: $(MOF_B200_LIB=$PWD/ab_variant.so timeout 200 python /tmp/prof_rf.py 2>&1 | tail -1)"
done | tee gpurun_out/r4v_ab_refine.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:rf_sad -s 2 -c 1 -o gpurun_out/r4v_rf python /tmp/prof_rf.py > gpurun_out/r4v_ncu.log 2>&1
ls -la gpurun_out/r4v_rf.ncu-rep
