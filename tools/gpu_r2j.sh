#!/bin/bash
# ncu captures of the current kernels (one launch each): N=240 pairs (TM), N=120 pairs (TM), N=60 pairs
mkdir -p gpurun_out
M=smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_ffma_pred_on.sum,smsp__sass_thread_inst_executed_op_fp32_pred_on.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_lsu.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active
timeout 200 python tools/prof_fft.py 240 240 2368 3 > gpurun_out/r2j_plain240.log 2>&1 &&
timeout 900 ncu --set full --metrics $M --clock-control none --import-source on -k regex:pc_pfa -s 1 -c 1 -o gpurun_out/r2j_pc240 python tools/prof_fft.py 240 240 2368 3 > gpurun_out/r2j_ncu240.log 2>&1
timeout 200 python tools/prof_fft.py 120 480 592 3 > gpurun_out/r2j_plain120.log 2>&1 &&
timeout 900 ncu --set full --metrics $M --clock-control none --import-source on -k regex:pc_pfa -s 1 -c 1 -o gpurun_out/r2j_pc120 python tools/prof_fft.py 120 480 592 3 > gpurun_out/r2j_ncu120.log 2>&1
timeout 200 python tools/prof_fft.py 60 480 592 3 > gpurun_out/r2j_plain60.log 2>&1 &&
timeout 900 ncu --set full --metrics $M --clock-control none --import-source on -k regex:pc_pfa -s 1 -c 1 -o gpurun_out/r2j_pc60 python tools/prof_fft.py 60 480 592 3 > gpurun_out/r2j_ncu60.log 2>&1
tail -1 gpurun_out/r2j_plain240.log gpurun_out/r2j_plain120.log gpurun_out/r2j_plain60.log
