#!/bin/bash
# PARK2 with the lane-derived invariants recomputed (no spills): parity subset, A/B/C, stage clocks of the new kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_steady_state.py -x -q -m gpu --timeout 600 -k "fft_pairs or known or geometry or device_pointer or weak or full_size_properties or pairs_steady or empty" > gpurun_out/r4b_pytest.log 2>&1
grep -E "passed|failed" gpurun_out/r4b_pytest.log | tail -n 1; grep "^FAILED\|Error" gpurun_out/r4b_pytest.log | head -8 | cut -c1-200
for i in 1 2; do
for cfg in "240 240 8288"; do
  echo "A(park2+opaque): $(timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
  echo "B(old): $(MOF_B200_LIB=$PWD/ab_variant.so timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
  echo "C(old+opaque): $(MOF_B200_LIB=$PWD/ab_variant2.so timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
done
done 2>&1 | tee gpurun_out/r4b_ab.txt
timeout 300 python tools/stage_clocks.py 240 240 > gpurun_out/r4b_stage_clocks_pc240.txt 2>&1
head -4 gpurun_out/r4b_stage_clocks_pc240.txt | cut -c1-260
