#!/bin/bash
# lane-derived invariants recomputed at N = 60 / 120 (no spills), with and without the dedicated staging buffer at N = 120
mkdir -p gpurun_out
for i in 1 2; do
for cfg in "120 480 2368" "120 240 4736" "60 480 4096" "60 240 8192"; do
  echo "cfg $cfg"
  echo "  A (in-tree):            $(timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
  echo "  B (recompute):          $(MOF_B200_LIB=$PWD/ab_B.so timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
  echo "  C (recompute + XSTAGE): $(MOF_B200_LIB=$PWD/ab_C.so timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
done
done 2>&1 | tee gpurun_out/r4l_ab_opaque.txt
