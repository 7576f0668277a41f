#!/bin/bash
mkdir -p gpurun_out
{
for cfg in "60 480 4096" "60 240 8192"; do
  echo "A (one lock-step group of 15 warps): $(timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -n 1)"
  echo "B (two groups, 8 + 7 warps, a row stage apart): $(MOF_B200_LIB=$PWD/ab_variant.so timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -n 1)"
done
} > gpurun_out/r3o_solo_groups_ab.txt 2>&1
cat gpurun_out/r3o_solo_groups_ab.txt | cut -c1-250
MOF_B200_LIB=$PWD/ab_variant.so timeout 900 python -m pytest tests/test_gpu_steady_state.py tests/test_gpu_parity.py -m gpu -q -k "60" > gpurun_out/r3o_pytest.log 2>&1
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/r3o_pytest.log | cut -c1-200
