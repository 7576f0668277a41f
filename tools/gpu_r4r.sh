#!/bin/bash
# N = 240 scratch variants without the OCL / window code (EXTRAS = false): tests that touch them, then the stream workloads
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu --timeout 600 -k "stream or push or ocl or hann or unaligned or NO_TMEM or tmem or steady or seq or raw or cfg5 or adapter or class" > gpurun_out/r4r_pytest.log 2>&1
grep -E "passed|failed" gpurun_out/r4r_pytest.log | tail -n 1; grep "^FAILED\|Error" gpurun_out/r4r_pytest.log | head -8 | cut -c1-200
echo "in-tree:"; timeout 300 python tools/prof_streams.py 2>&1 | tail -4
echo "MOF_EXTRAS_ALWAYS=1 (previous variants):"; MOF_EXTRAS_ALWAYS=1 timeout 300 python tools/prof_streams.py 2>&1 | tail -4
