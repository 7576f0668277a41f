#!/bin/bash
# opaque lane index at every stage call (nothing lane-derived hoisted across stages): A = in-tree, B = N=60/120, C = all sizes
mkdir -p gpurun_out
for i in 1 2; do
for cfg in "120 480 2368" "120 240 4736" "60 480 4096" "60 240 8192" "240 240 8288"; do
  echo "cfg $cfg"
  echo "  A (in-tree):       $(timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
  echo "  B (stages 60/120): $(MOF_B200_LIB=$PWD/ab_B.so timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
  echo "  C (stages all):    $(MOF_B200_LIB=$PWD/ab_C.so timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
done
done 2>&1 | tee gpurun_out/r4m_ab_opaque_stages.txt
