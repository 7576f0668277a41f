#!/bin/bash
mkdir -p gpurun_out
{
for cfg in "60 480 2368" "60 240 8192"; do
  echo "A (3 CTAs x 5 warps, lock-step at team barriers): $(timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
  echo "S1 (A + block barrier between all stages):        $(MOF_B200_LIB=$PWD/ab_s1.so timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
  echo "S2 (1 CTA x 15 warps):                            $(MOF_B200_LIB=$PWD/ab_s2.so timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
  echo "S3 (S2 + block barrier between all stages):       $(MOF_B200_LIB=$PWD/ab_s3.so timeout 120 python tools/prof_fft.py $cfg 10 2>&1 | tail -1)"
done
} > gpurun_out/r2l_ab.txt 2>&1
cat gpurun_out/r2l_ab.txt
