#!/bin/bash
# N = 60: step descriptors filled once per step (solo warps); N=60 tests, then A/B
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu --timeout 600 -k "60 or cfg4 or steady or stream or longrange or fft_pairs" > gpurun_out/r4k_pytest.log 2>&1
grep -E "passed|failed" gpurun_out/r4k_pytest.log | tail -n 1; grep "^FAILED\|Error" gpurun_out/r4k_pytest.log | head -8 | cut -c1-200
for i in 1 2; do
for cfg in "60 480 4096" "60 240 8192"; do
  echo "A (once): $(timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
  echo "B (twice): $(MOF_B200_LIB=$PWD/ab_variant.so timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
done
done 2>&1 | tee gpurun_out/r4k_ab_solo_step_once.txt
