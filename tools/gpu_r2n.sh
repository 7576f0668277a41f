#!/bin/bash
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/r2n_pytest.log 2>&1
echo "pytest rc $?" >> gpurun_out/r2n_pytest.log
{
for cfg in "240 240 8192" "120 480 2368" "60 480 2368" "60 240 8192"; do
  echo "$(timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
done
} > gpurun_out/r2n_perf.txt 2>&1
timeout 200 python tools/prof_fft.py 60 480 592 3 > gpurun_out/r2n_plain60.log 2>&1 &&
timeout 900 ncu --set full --clock-control none -k regex:pc_pfa -s 1 -c 1 -o gpurun_out/r2n_pc60 python tools/prof_fft.py 60 480 592 3 > gpurun_out/r2n_ncu60.log 2>&1
grep -E "passed|failed" gpurun_out/r2n_pytest.log | tail -n 1; grep -c "^FAILED" gpurun_out/r2n_pytest.log; cat gpurun_out/r2n_perf.txt
