#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu --timeout 600 -k "refine or Refine or blockmethod or BlockMethod or bm" 2>&1 | grep -E "passed|failed" | tail -n 1
MOF_RF_NO_FUSED=1 timeout 600 python -m pytest tests -x -q -m gpu --timeout 600 -k "refine or Refine" 2>&1 | grep -E "passed|failed" | tail -n 1
timeout 300 python __graft_entry__.py smoke 2>&1 | tail -n 1 | cut -c1-200
