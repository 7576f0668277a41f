#!/bin/bash
mkdir -p gpurun_out
{
for cfg in "60 480 4096" "60 240 8192"; do
  echo "G=1 (one lock-step group of 15 warps): $(timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -n 1)"
  for g in 2 3 5; do
    echo "G=$g: $(MOF_B200_LIB=$PWD/ab_g$g.so timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -n 1)"
  done
done
for cfg in "120 480 2368" "120 240 8192"; do
  echo "A (teams re-aligned once per section): $(timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -n 1)"
  echo "B (free-running teams, MOF_LOCKSTEP=2): $(MOF_B200_LIB=$PWD/ab_l2.so timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -n 1)"
  echo "C (free-running teams, one row stage apart at the start): $(MOF_B200_LIB=$PWD/ab_ts.so timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -n 1)"
done
} > gpurun_out/r3p_groups_ab.txt 2>&1
cat gpurun_out/r3p_groups_ab.txt | cut -c1-220
