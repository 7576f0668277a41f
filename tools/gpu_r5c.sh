#!/bin/bash
# generic-size kernel experiments (not adopted, profiles/r5c_*): expects ab_old.so (tree), ab_c/d/e.so (variants) built with tools/ab.sh
# everywhere; D = fused, register-DFT order everywhere; E = C without the fusion (bitwise reference for C)
mkdir -p gpurun_out
{
for cfg in "480 480 888" "160 480 444" "96 480 512" "80 240 2048" "30 240 1024" "768 768 444" "256 512 512" "100 300 1024" "600 600 444" "720 720 444" "450 450 444"; do
  echo "A $cfg: $(MOF_B200_LIB=$PWD/ab_old.so timeout 120 python tools/prof_fft.py $cfg 4 2>&1 | tail -1)"
  echo "B $cfg: $(timeout 120 python tools/prof_fft.py $cfg 4 2>&1 | tail -1)"
  echo "C $cfg: $(MOF_B200_LIB=$PWD/ab_c.so timeout 120 python tools/prof_fft.py $cfg 4 2>&1 | tail -1)"
  echo "D $cfg: $(MOF_B200_LIB=$PWD/ab_d.so timeout 120 python tools/prof_fft.py $cfg 4 2>&1 | tail -1)"
  echo "E $cfg: $(MOF_B200_LIB=$PWD/ab_e.so timeout 120 python tools/prof_fft.py $cfg 4 2>&1 | tail -1)"
done
for cfg in "480 480 600" "96 480 300" "30 240 300" "256 512 400"; do
  set -- $cfg
  MOF_B200_LIB=$PWD/ab_c.so timeout 120 python tools/pg_dump.py $cfg /tmp/c.npy > /dev/null 2>&1
  MOF_B200_LIB=$PWD/ab_e.so timeout 120 python tools/pg_dump.py $cfg /tmp/e.npy > /dev/null 2>&1
  python -c "import numpy as np; c=np.load('/tmp/c.npy'); e=np.load('/tmp/e.npy'); print('fused vs three-pass, N=$1 F=$2 pairs=$3: bitwise equal =', bool((c.view(np.uint64)==e.view(np.uint64)).all()), 'finite', float(np.isfinite(c).mean()))"
done
} > gpurun_out/r5c_ab_pg_fused_columns.txt 2>&1
cat gpurun_out/r5c_ab_pg_fused_columns.txt | cut -c1-170
timeout 900 python -m pytest tests/test_gpu_generic_sizes.py -m gpu -q 2>&1 | tail -3
