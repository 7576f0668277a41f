#!/bin/bash
# ncu source-level captures of the log-polar kernels (after the plain run exited 0)
mkdir -p gpurun_out
timeout 200 python tools/prof_side.py 592 > gpurun_out/r4g_plain_side.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:lp_kernel -s 2 -c 2 -o gpurun_out/r4g_lp python tools/prof_side.py 592 > gpurun_out/r4g_ncu_lp.log 2>&1
head -3 gpurun_out/r4g_plain_side.log
ls -la gpurun_out/r4g_lp.ncu-rep
