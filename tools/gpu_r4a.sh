#!/bin/bash
# PARK2 (second image parked in dead columns, N = 240 pairs): parity subset, then A/B against the previous kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_steady_state.py -x -q -m gpu --timeout 600 -k "fft_pairs or known or geometry or device_pointer or weak or full_size_properties or pairs_steady or empty" > gpurun_out/r4a_pytest.log 2>&1
grep -E "passed|failed" gpurun_out/r4a_pytest.log | tail -n 1; grep "^FAILED\|Error" gpurun_out/r4a_pytest.log | head -8 | cut -c1-200
for i in 1 2; do
for cfg in "240 240 8288"; do
  echo "A(new): $(timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
  echo "B(old): $(MOF_B200_LIB=$PWD/ab_variant.so timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
done
done 2>&1 | tee gpurun_out/r4a_ab.txt
