#!/bin/bash
# ncu captures (each after its plain run exited 0): N=240 pairs kernel with source counters, row-wide FastSpacedBM kernel
mkdir -p gpurun_out
timeout 200 python tools/prof_fft.py 240 240 2368 3 > gpurun_out/r4d_plain240.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:pc_pfa -s 1 -c 1 -o gpurun_out/r4d_pc240 python tools/prof_fft.py 240 240 2368 3 > gpurun_out/r4d_ncu240.log 2>&1
tail -n 1 gpurun_out/r4d_plain240.log
timeout 200 python tools/prof_bm.py 2368 > gpurun_out/r4d_plainbm.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:bm_sad_wide -s 1 -c 1 -o gpurun_out/r4d_bmwide python tools/prof_bm.py 2368 > gpurun_out/r4d_ncubm.log 2>&1
tail -n 2 gpurun_out/r4d_plainbm.log
ls -la gpurun_out/*.ncu-rep
