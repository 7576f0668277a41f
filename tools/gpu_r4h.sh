#!/bin/bash
# next image hinted into L1 instead of L2 (prefetch.global.L1): A = in-tree (L2 hints), B/C/D = variants
mkdir -p gpurun_out
for cfg in "120 480 2368" "120 240 4736" "60 480 2368" "60 240 4736"; do
  echo "cfg $cfg"
  echo "  A (L2 hints):      $(timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
  echo "  B (L1 loads N=120): $(MOF_B200_LIB=$PWD/ab_l1_8.so timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
  echo "  C (L1 loads 60/120):$(MOF_B200_LIB=$PWD/ab_l1_12.so timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
done 2>&1 | tee gpurun_out/r4i_ab_l1_loads.txt
