#!/bin/bash
# N = 120: the team staging buffer (MOF_XSTAGE) measured again on the final kernel (no spills in the pairs variant, free-running teams)
mkdir -p gpurun_out
{
for cfg in "120 480 2368" "120 240 8192"; do
  echo "A (tree) $cfg: $(timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
  echo "B (MOF_XSTAGE=1) $cfg: $(MOF_B200_LIB=$PWD/ab_c.so timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
  echo "C (MOF_XSTAGE=1, MOF_PREFETCH_L1=0) $cfg: $(MOF_B200_LIB=$PWD/ab_d.so timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
done
} > gpurun_out/r5d_ab_xstage_again.txt 2>&1
cat gpurun_out/r5d_ab_xstage_again.txt | cut -c1-200
