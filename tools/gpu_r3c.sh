#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/prof_new.py > gpurun_out/r3c_plain.log 2>&1 &&
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"fe_bgr16|pg_kernel" -c 2 -o gpurun_out/r3c_new python tools/prof_new.py once > gpurun_out/r3c_ncu.log 2>&1
cat gpurun_out/r3c_plain.log | cut -c1-200; tail -n 3 gpurun_out/r3c_ncu.log | cut -c1-200
