#!/bin/bash
mkdir -p gpurun_out
true
true
timeout 400 python bench.py --workload streams --steps 5 --warmup 3 > gpurun_out/r2w_bench_streams_1gpu.json 2>gpurun_out/r2w_err.txt
MOF_FE_FULL_H2D=1 timeout 400 python bench.py --workload streams --steps 5 --warmup 3 > gpurun_out/r2w_bench_streams_1gpu_fullh2d.json 2>>gpurun_out/r2w_err.txt
tail -n 5 gpurun_out/r2w_err.txt | cut -c1-300
python -c "
import json
for f in ('r2w_bench_streams_1gpu','r2w_bench_streams_1gpu_fullh2d'):
    d=json.load(open('gpurun_out/'+f+'.json')); print(f, d['value'], d['ms_per_step'], d.get('e2e'))"
