#!/bin/bash
# round-end pass after the push_many tape cut: full GPU suite, smoke, headline bench, reference arm, streams bench, launch list
mkdir -p gpurun_out
R=r5i
timeout 1500 python -m pytest tests -x -q -m gpu --timeout 900 > gpurun_out/${R}_pytest.log 2>&1
grep -E "passed|failed" gpurun_out/${R}_pytest.log | tail -n 1; grep "^FAILED" gpurun_out/${R}_pytest.log | head -8 | cut -c1-200
timeout 300 python __graft_entry__.py smoke > gpurun_out/${R}_smoke.log 2>&1; tail -n 1 gpurun_out/${R}_smoke.log | cut -c1-300
timeout 600 python bench.py > gpurun_out/${R}_bench_1gpu.json 2> gpurun_out/${R}_bench_err.txt
timeout 400 python bench.py --impl reference > gpurun_out/${R}_bench_reference_1gpu.json 2>> gpurun_out/${R}_bench_err.txt
timeout 400 python bench.py --workload streams --steps 5 --warmup 3 > gpurun_out/${R}_bench_streams_1gpu.json 2>> gpurun_out/${R}_bench_err.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${R}_launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu --no-extras --no-parity > gpurun_out/${R}_ncu_bench.log 2>&1
tail -n 3 gpurun_out/${R}_bench_err.txt | cut -c1-200
python - <<'PY'
import json
R = "r5i"
d = json.load(open(f'gpurun_out/{R}_bench_1gpu.json'))
print('value', d['value'], 'fp32', d['roofline_fp32']['frac'], 'e2e', d['e2e']['value'], 'parity', d.get('parity_max_err_px'), 'cpu', d['cpu_baseline']['value'], d['cpu_baseline']['kind'])
for w in d.get('other_workloads', []):
    print({k: (round(v, 3) if isinstance(v, float) else v) for k, v in w.items() if k in ('workload', 'pairs_per_s', 'frames_per_s', 'blocks_per_s', 'error', 'fp32_frac_nominal', 'h2d_gbs')})
r = json.load(open(f'gpurun_out/{R}_bench_reference_1gpu.json')); print('reference', r.get('value'), r.get('cpu_baseline', {}).get('cores'))
s = json.load(open(f'gpurun_out/{R}_bench_streams_1gpu.json')); print('streams', s['value'], s.get('e2e', {}).get('value'), s.get('single_frame_push_latency_us'))
PY
