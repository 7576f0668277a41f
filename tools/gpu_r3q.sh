#!/bin/bash
mkdir -p gpurun_out
{
for cfg in "60 480 4096" "60 240 8192" "120 480 2368" "120 240 8192"; do
  echo "base: $(timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -n 1)"
  echo "3 groups / teams skewed, released after the first row stage: $(MOF_B200_LIB=$PWD/ab_s1.so timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -n 1)"
  echo "3 groups / teams skewed, released after the first pass: $(MOF_B200_LIB=$PWD/ab_s2.so timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -n 1)"
  echo "4 groups (N=60) / 3 teams, released after the first pass: $(MOF_B200_LIB=$PWD/ab_s3.so timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -n 1)"
done
} > gpurun_out/r3q_skew_ab.txt 2>&1
cat gpurun_out/r3q_skew_ab.txt | cut -c1-220
