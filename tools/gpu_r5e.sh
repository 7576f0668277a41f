#!/bin/bash
# packed (v, -v) constants from constant memory (MOF_CONST_BANK=1, ab_c.so) against the tree (UMOV pairs)
mkdir -p gpurun_out
{
for rep in 1 2; do
for cfg in "240 240 8288" "120 480 2368" "60 480 2368" "120 240 8192"; do
  echo "A (tree) $cfg: $(timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
  echo "B (MOF_CONST_BANK=1) $cfg: $(MOF_B200_LIB=$PWD/ab_c.so timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
done
done
} > gpurun_out/r5e_ab_const_bank.txt 2>&1
cat gpurun_out/r5e_ab_const_bank.txt | cut -c1-200
