#!/bin/bash
# round-2 first GPU pass: full parity suite, both bench arms, per-stage clocks, one ncu capture with FP32 op counters
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -s > gpurun_out/r2a_pytest.log 2>&1
echo "pytest rc $?" >> gpurun_out/r2a_pytest.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2a_bench_1gpu.json 2> gpurun_out/r2a_bench.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2a_bench_reference_1gpu.json 2>> gpurun_out/r2a_bench.err
timeout 300 python tools/stage_clocks.py 240 240 > gpurun_out/r2a_stage_clocks_pc240.txt 2>&1
M=smsp__sass_thread_inst_executed_op_fadd_pred_on.sum,smsp__sass_thread_inst_executed_op_fmul_pred_on.sum,smsp__sass_thread_inst_executed_op_ffma_pred_on.sum,smsp__sass_thread_inst_executed_op_fp32_pred_on.sum,sm__inst_executed_pipe_fma.sum,sm__inst_executed_pipe_fmaheavy.sum,sm__inst_executed_pipe_fmalite.sum,sm__inst_executed_pipe_alu.sum,sm__inst_executed_pipe_lsu.sum,sm__inst_executed_pipe_xu.sum,sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active,sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active,smsp__inst_executed_pipe_fma.sum,sm__sass_thread_inst_executed_op_fp32_pred_on.sum
timeout 200 python tools/prof_fft.py 240 240 2368 3 > gpurun_out/r2a_plain.log 2>&1 &&
timeout 900 ncu --set full --metrics $M --clock-control none --import-source on -k regex:pc_pfa -s 1 -c 1 -o gpurun_out/r2a_pc240 python tools/prof_fft.py 240 240 2368 3 > gpurun_out/r2a_ncu.log 2>&1
tail -3 gpurun_out/r2a_pytest.log; cat gpurun_out/r2a_plain.log | tail -1
