#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/r3n_alone.log 2>&1
echo "alone: $(grep -E 'passed|failed' gpurun_out/r3n_alone.log | tail -n 1)"; grep -E "^(FAILED|ERROR)|^E " gpurun_out/r3n_alone.log | head -10 | cut -c1-250
timeout 1800 python -m pytest tests -x -q -m gpu --timeout 900 > gpurun_out/r3n_full.log 2>&1
echo "full: $(grep -E 'passed|failed' gpurun_out/r3n_full.log | tail -n 1)"; grep -E "^(FAILED|ERROR)" gpurun_out/r3n_full.log | head -10 | cut -c1-250
tail -n 4 gpurun_out/r3n_full.log | cut -c1-200
