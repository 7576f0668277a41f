#!/bin/bash
# ncu captures of the final N = 120 and N = 60 pairs kernels (one launch each)
mkdir -p gpurun_out
M=smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_ffma_pred_on.sum,smsp__sass_thread_inst_executed_op_fp32_pred_on.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_lsu.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active
timeout 200 python tools/prof_fft.py 120 480 592 3 > gpurun_out/r5j_plain120.log 2>&1 &&
timeout 600 ncu --set full --metrics $M --clock-control none --import-source on -k regex:pc_pfa -s 1 -c 1 -o gpurun_out/r5j_pc120 python tools/prof_fft.py 120 480 592 3 > gpurun_out/r5j_ncu120.log 2>&1
timeout 200 python tools/prof_fft.py 60 480 592 3 > gpurun_out/r5j_plain60.log 2>&1 &&
timeout 600 ncu --set full --metrics $M --clock-control none --import-source on -k regex:pc_pfa -s 1 -c 1 -o gpurun_out/r5j_pc60 python tools/prof_fft.py 60 480 592 3 > gpurun_out/r5j_ncu60.log 2>&1
tail -q -n 1 gpurun_out/r5j_plain120.log gpurun_out/r5j_plain60.log
