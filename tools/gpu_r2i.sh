#!/bin/bash
mkdir -p gpurun_out
{
for cfg in "120 480 2368" "120 240 8192"; do
  echo "A (lock-step teams):        $(timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
  echo "B (team barriers, realign): $(MOF_B200_LIB=$PWD/ab_ls0.so timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
  echo "C (free-running teams):     $(MOF_B200_LIB=$PWD/ab_ls2.so timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
done
echo "A N=240: $(timeout 120 python tools/prof_fft.py 240 240 8192 10 2>&1 | tail -1)"
} > gpurun_out/r2i_ab.txt 2>&1
MOF_B200_LIB=$PWD/ab_ls2.so timeout 600 python -m pytest tests/test_gpu_steady_state.py -m gpu -q -k "120" > gpurun_out/r2i_pytest_ls2.log 2>&1
tail -1 gpurun_out/r2i_pytest_ls2.log; cat gpurun_out/r2i_ab.txt
