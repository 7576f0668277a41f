#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/r2m_pytest.log 2>&1
echo "pytest rc $?" >> gpurun_out/r2m_pytest.log
{
for cfg in "60 480 2368" "60 240 8192"; do
  echo "A (1 x 15, L2 scratch):  $(timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
  echo "T (1 x 15, spectrum in TMEM): $(MOF_B200_LIB=$PWD/ab_solotm.so timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
done
} > gpurun_out/r2m_ab.txt 2>&1
MOF_B200_LIB=$PWD/ab_solotm.so timeout 600 python -m pytest tests/test_gpu_steady_state.py tests/test_gpu_parity.py -m gpu -q -k "60" > gpurun_out/r2m_pytest_tm60.log 2>&1
grep -E "passed|failed" gpurun_out/r2m_pytest.log | tail -1; grep -c "^FAILED" gpurun_out/r2m_pytest.log; tail -n 1 gpurun_out/r2m_pytest_tm60.log; cat gpurun_out/r2m_ab.txt
