#!/bin/bash
# 2-GPU sanity pass: the multi-GPU tests, then bench.py under torchrun (pairs and streams workloads)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu --timeout 600 -k "multi" > gpurun_out/r5k_pytest_multi.log 2>&1
grep -E "passed|failed" gpurun_out/r5k_pytest_multi.log | tail -n 1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu --no-extras > gpurun_out/r5k_bench_2gpu.json 2> gpurun_out/r5k_err.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --workload streams --steps 5 --warmup 3 > gpurun_out/r5k_bench_streams_2gpu.json 2>> gpurun_out/r5k_err.txt
tail -n 2 gpurun_out/r5k_err.txt | cut -c1-200
python - <<'PY'
import json
for f in ('r5k_bench_2gpu.json', 'r5k_bench_streams_2gpu.json'):
    try:
        d = json.loads(open('gpurun_out/' + f).read().strip().splitlines()[-1])
        print(f, d['n_gpus'], d['value'], d.get('e2e', {}).get('value'), d.get('scaling'))
    except Exception as e:
        print(f, 'unreadable', e)
PY
