#!/bin/bash
# row-wide FastSpacedBM kernel: bit-exactness tests, then A/B against the previous kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu --timeout 600 -k "bm or blockmethod or BlockMethod" > gpurun_out/r4c_pytest.log 2>&1
grep -E "passed|failed" gpurun_out/r4c_pytest.log | tail -n 1; grep "^FAILED\|Error" gpurun_out/r4c_pytest.log | head -8 | cut -c1-200
{ echo "wide:"; timeout 200 python tools/prof_bm.py 4096 2>&1 | tail -2; echo "previous kernel (MOF_BM_NO_WIDE=1):"; MOF_BM_NO_WIDE=1 timeout 200 python tools/prof_bm.py 4096 2>&1 | tail -2; } | tee gpurun_out/r4c_ab_bm_wide.txt
