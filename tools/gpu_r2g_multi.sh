#!/bin/bash
# multi-GPU pass (one 8-GPU box): bitwise multi-device test, cfg5 streams at 1/2/4/8, headline pairs at 8,
# single-process multi-context e2e at 1/2/4/8 and the box's concurrent H2D ceiling
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2g_topo.txt 2>&1
MOF_TEST_MULTI_GPU=1 timeout 300 python -m pytest tests/test_gpu_round2.py -m gpu -q -k "multi" > gpurun_out/r2g_pytest_multi.log 2>&1
echo "rc $?" >> gpurun_out/r2g_pytest_multi.log
timeout 120 tools/ubench/h2d_sweep 256 > gpurun_out/r2g_h2d_sweep.json 2>&1
for n in 1 2 4 8; do
  if [ $n = 1 ]; then L="python"; else L="python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500+n))"; fi
  timeout 300 $L bench.py --gpus $n --workload streams --steps 5 --warmup 3 > gpurun_out/r2g_bench_streams_${n}gpu.json 2>> gpurun_out/r2g_err.log
  timeout 300 python bench.py --gpus $n --workload multi --steps 5 --warmup 3 > gpurun_out/r2g_bench_multictx_${n}gpu.json 2>> gpurun_out/r2g_err.log
done
timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29600 bench.py --gpus 8 --steps 10 --warmup 3 > gpurun_out/r2g_bench_8gpu.json 2>> gpurun_out/r2g_err.log
tail -2 gpurun_out/r2g_pytest_multi.log; cat gpurun_out/r2g_h2d_sweep.json; for f in gpurun_out/r2g_bench_*; do echo $f; python -c "
import json,sys
d=json.load(open('$f')); print({k:d.get(k) for k in ('value','n_gpus','ms_per_step','single_frame_push_latency_us')}, (d.get('e2e') or {}).get('value'))"; done; tail -5 gpurun_out/r2g_err.log
