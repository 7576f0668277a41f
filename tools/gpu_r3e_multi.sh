#!/bin/bash
# multi-GPU pass (one 8-GPU box): bitwise multi-device test, cfg5 streams at 2/4/8, headline pairs at 8,
# single-process multi-context e2e at 8
mkdir -p gpurun_out
MOF_TEST_MULTI_GPU=1 timeout 300 python -m pytest tests/test_gpu_round2.py -m gpu -q -k "multi" > gpurun_out/r3e_pytest_multi.log 2>&1
echo "rc $?" >> gpurun_out/r3e_pytest_multi.log
for n in 1 2 4 8; do
  if [ $n = 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n))"; fi
  timeout 300 $L bench.py --gpus $n --workload streams --steps 5 --warmup 3 > gpurun_out/r3e_bench_streams_${n}gpu.json 2>> gpurun_out/r3e_err.log
done
timeout 300 python bench.py --gpus 8 --workload multi --steps 5 --warmup 3 > gpurun_out/r3e_bench_multictx_8gpu.json 2>> gpurun_out/r3e_err.log
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29600 bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu --no-extras > gpurun_out/r3e_bench_8gpu.json 2>> gpurun_out/r3e_err.log
tail -n 2 gpurun_out/r3e_pytest_multi.log | cut -c1-200; for f in gpurun_out/r3e_bench_*; do echo $f; python -c "
import json,sys
d=json.load(open('$f')); print({k:d.get(k) for k in ('value','n_gpus','ms_per_step','single_frame_push_latency_us')}, (d.get('e2e') or {}).get('value'))"; done; tail -n 5 gpurun_out/r3e_err.log | cut -c1-200
