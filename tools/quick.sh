#!/bin/bash
# quick GPU check: parity tests + kernel-only throughput at the three section sizes
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for cfg in "240 240 4736" "120 480 1184" "60 480 1184" "60 240 4736"; do
  timeout 120 python tools/prof_fft.py $cfg 5 2>&1 | tail -1
done
