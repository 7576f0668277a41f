#!/bin/bash
# cfg5 with pinned HOST frames: one 3-D copy per chunk (tree) against one 2-D copy per stream and chunk (MOF_FE_H2D_PER_STREAM=1)
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu -k "raw" 2>&1 | tail -2
{
for rep in 1 2; do
  echo "3-D copy per chunk: $(timeout 300 python bench.py --workload streams --steps 5 --warmup 3 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e'])")"
  echo "2-D copy per stream and chunk: $(MOF_FE_H2D_PER_STREAM=1 timeout 300 python bench.py --workload streams --steps 5 --warmup 3 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e'])")"
done
} > gpurun_out/r5l_ab_h2d_3d.txt 2>&1
cut -c1-400 gpurun_out/r5l_ab_h2d_3d.txt
