#!/bin/bash
mkdir -p gpurun_out
{
for cfg in "60 480 2368" "60 240 8192"; do
  echo "A  (fetch at the row stage):          $(timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
  for v in 1 2 3; do echo "PV$v (early requests, bits = $v):       $(MOF_B200_LIB=$PWD/ab_pv$v.so timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"; done
done
} > gpurun_out/r2t_ab.txt 2>&1
MOF_B200_LIB=$PWD/ab_pv3.so timeout 600 python -m pytest tests/test_gpu_steady_state.py tests/test_gpu_parity.py -m gpu -q -k "60" > gpurun_out/r2t_pytest.log 2>&1
tail -n 1 gpurun_out/r2t_pytest.log; cat gpurun_out/r2t_ab.txt
