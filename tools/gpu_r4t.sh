#!/bin/bash
# BlockMethod::Refine: fused kernel (2x image in shared memory) vs MOF_RF_NO_FUSED=1 (up-scale kernel + SAD kernel through HBM) vs the first version (ab_variant.so)
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu --timeout 600 -k "refine or Refine or blockmethod or BlockMethod" > gpurun_out/r4t_pytest.log 2>&1
grep -E "passed|failed" gpurun_out/r4t_pytest.log | tail -n 1; grep "^FAILED\|Error\|assert" gpurun_out/r4t_pytest.log | head -8 | cut -c1-200
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
echo "unfused:  $(MOF_RF_NO_FUSED=1 timeout 200 python /tmp/prof_rf.py 2>&1 | tail -1)"
echo "first:    $(MOF_B200_LIB=$PWD/ab_variant.so timeout 200 python /tmp/prof_rf.py 2>&1 | tail -1)"
done | tee gpurun_out/r4t_ab_refine.txt
