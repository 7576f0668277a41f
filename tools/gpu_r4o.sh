#!/bin/bash
# cfg5 streams bench and stream-mode kernels: in-tree (recompute at N=60/120) vs -DMOF_OPAQUE_WORDC=0
mkdir -p gpurun_out
for i in 1 2 3; do
echo "new: $(timeout 300 python bench.py --workload streams --steps 5 --warmup 3 2>/dev/null | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["value"], d.get("e2e",{}).get("value"), d.get("single_frame_push_latency_us"))')"
echo "old: $(MOF_B200_LIB=$PWD/ab_variant.so timeout 300 python bench.py --workload streams --steps 5 --warmup 3 2>/dev/null | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["value"], d.get("e2e",{}).get("value"), d.get("single_frame_push_latency_us"))')"
done | tee gpurun_out/r4o_ab_streams_wordc.txt
