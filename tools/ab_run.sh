#!/bin/bash
# on the GPU box: in-tree library (A) vs ab_variant.so (B), kernel-only throughput, same inputs
for cfg in "240 240 4736" "120 480 2368" "60 480 2368" "60 240 4736"; do
  echo "A: $(timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
  echo "B: $(MOF_B200_LIB=$PWD/ab_variant.so timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
done
