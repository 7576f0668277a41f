#!/bin/bash
# cfg5 streams on one 8-GPU box after the bundled cut: 1 GPU and 8 GPUs (device-resident leg only), uniform cut beside it
mkdir -p gpurun_out
L8="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29508"
timeout 200 python bench.py --gpus 1 --workload streams --steps 5 --warmup 3 --no-e2e > gpurun_out/r5g_bench_streams_1gpu.json 2>> gpurun_out/r5g_err.log
timeout 200 $L8 bench.py --gpus 8 --workload streams --steps 5 --warmup 3 --no-e2e > gpurun_out/r5g_bench_streams_8gpu.json 2>> gpurun_out/r5g_err.log
MOF_SEQ_UNIFORM=1 timeout 200 $L8 bench.py --gpus 8 --workload streams --steps 5 --warmup 3 --no-e2e > gpurun_out/r5g_bench_streams_8gpu_uniform.json 2>> gpurun_out/r5g_err.log
for f in gpurun_out/r5g_bench_*; do echo $f; python -c "
import json,sys
d=json.load(open('$f')); print({k:d.get(k) for k in ('value','n_gpus','ms_per_step','single_frame_push_latency_us')})"; done; tail -n 5 gpurun_out/r5g_err.log | cut -c1-200
