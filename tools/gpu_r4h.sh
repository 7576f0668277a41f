#!/bin/bash
# next image hinted into L1 instead of L2 (prefetch.global.L1): A = in-tree (L2 hints), B/C/D = variants
mkdir -p gpurun_out
for cfg in "120 480 2368" "120 240 4736" "60 480 2368" "60 240 4736" "240 240 8288"; do
  echo "cfg $cfg"
  echo "  A (L2 hints):      $(timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
  echo "  B (L1 at N=120):   $(MOF_B200_LIB=$PWD/ab_l1_8.so timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
  echo "  C (L1 at N=60/120):$(MOF_B200_LIB=$PWD/ab_l1_12.so timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
  echo "  D (L1 everywhere): $(MOF_B200_LIB=$PWD/ab_l1_28.so timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
done 2>&1 | tee gpurun_out/r4h_ab_prefetch_l1.txt
