#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/stage_clocks.py 120 480 > gpurun_out/r4x_stage_clocks_pc120.txt 2>&1
head -20 gpurun_out/r4x_stage_clocks_pc120.txt | cut -c1-230
