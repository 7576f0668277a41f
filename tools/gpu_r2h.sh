#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_steady_state.py -m gpu -q --timeout 900 -k "240 or pairs" > gpurun_out/r2h_pytest.log 2>&1
echo "pytest rc $?" >> gpurun_out/r2h_pytest.log
{
for rep in 1 2; do
for cfg in "240 240 8192"; do
  echo "A (in-tree: X16 loads one round ahead): $(timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
  echo "B (+ IY16 units one ahead):             $(MOF_B200_LIB=$PWD/ab_y16.so timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
  echo "C (no pipelining):                      $(MOF_B200_LIB=$PWD/ab_nopipe.so timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
done; done
} > gpurun_out/r2h_ab.txt 2>&1
grep -E "passed|failed" gpurun_out/r2h_pytest.log | tail -1; cat gpurun_out/r2h_ab.txt
