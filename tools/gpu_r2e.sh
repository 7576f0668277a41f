#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -s > gpurun_out/r2e_pytest.log 2>&1
echo "pytest rc $?" >> gpurun_out/r2e_pytest.log
{
for cfg in "240 240 4736" "240 240 8192"; do
  echo "A (in-tree): $(timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
  echo "D (cp.async by main warps): $(MOF_B200_LIB=$PWD/ab_cpasync.so timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
done
} > gpurun_out/r2e_ab.txt 2>&1
timeout 300 python tools/stage_clocks.py 240 240 > gpurun_out/r2e_stage_clocks_pc240.txt 2>&1
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/r2e_pytest.log | cut -c1-150; cat gpurun_out/r2e_ab.txt; head -5 gpurun_out/r2e_stage_clocks_pc240.txt | cut -c1-250
