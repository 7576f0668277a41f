#!/bin/bash
mkdir -p gpurun_out
{
for cfg in "60 480 2368" "60 240 8192"; do
  echo "A (per-section realign):  $(timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
  echo "B (lock-step every stage group): $(MOF_B200_LIB=$PWD/ab_solo1.so timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
done
} > gpurun_out/r2k_ab.txt 2>&1
MOF_B200_LIB=$PWD/ab_solo1.so timeout 600 python -m pytest tests/test_gpu_steady_state.py tests/test_gpu_parity.py -m gpu -q -k "60" > gpurun_out/r2k_pytest.log 2>&1
tail -n 1 gpurun_out/r2k_pytest.log; cat gpurun_out/r2k_ab.txt
