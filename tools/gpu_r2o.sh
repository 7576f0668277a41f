#!/bin/bash
# round-2 final 1-GPU numbers: smoke, both bench arms, launch list of the bench command, FP32 microbenchmark
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2o_smoke.log 2>&1; echo "smoke rc $?" >> gpurun_out/r2o_smoke.log
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/r2o_bench_reference_1gpu.json 2> gpurun_out/r2o_bench.err
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/r2o_bench_1gpu.json 2>> gpurun_out/r2o_bench.err
timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu --no-extras --no-parity > gpurun_out/r2o_plain.log 2>&1 &&
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r2o_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-extras --no-parity > gpurun_out/r2o_ncu_launches.log 2>&1
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/f32x2 tools/ubench/f32x2.cu 2>/dev/null && /tmp/f32x2 > gpurun_out/r2o_ubench_f32x2.txt 2>&1
tail -2 gpurun_out/r2o_smoke.log; tail -3 gpurun_out/r2o_bench.err; cat gpurun_out/r2o_ubench_f32x2.txt | tail -6
