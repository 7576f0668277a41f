#!/bin/bash
# row-wide FastSpacedBM: window views on the FMA pipe (IMAD.HI) vs PRMT; bit-exactness tests first
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu --timeout 600 -k "bm or blockmethod or BlockMethod" > gpurun_out/r4e_pytest.log 2>&1
grep -E "passed|failed" gpurun_out/r4e_pytest.log | tail -n 1; grep "^FAILED\|Error" gpurun_out/r4e_pytest.log | head -8 | cut -c1-200
{ echo "views with IMAD.HI / IMAD (FMA pipe):"; timeout 200 python tools/prof_bm.py 4096 2>&1 | tail -2; echo "views with PRMT (integer pipe), -DMOF_BMW_VIEW_IMAD=0:"; MOF_B200_LIB=$PWD/ab_variant.so timeout 200 python tools/prof_bm.py 4096 2>&1 | tail -2; } | tee gpurun_out/r4e_ab_bm_views.txt
