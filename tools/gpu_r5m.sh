#!/bin/bash
# host-frame push_many: number of chunks per call (MOF_PIPE_CHUNKS), gray 240^2 frames and cfg5 raw frames
mkdir -p gpurun_out
{
for k in 8 4 3 2; do
  echo "gray F=240 N=240, 296 streams x 16 frames, $k chunks: $(MOF_PIPE_CHUNKS=$k timeout 120 python tools/prof_e2e_streams.py 2>&1 | tail -n 1)"
  echo "cfg5 raw, 64 streams x 16 frames, $k chunks: $(MOF_PIPE_CHUNKS=$k timeout 200 python bench.py --workload streams --steps 3 --warmup 3 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['e2e']['value'])")"
done
} > gpurun_out/r5m_pipe_chunks.txt 2>&1
cat gpurun_out/r5m_pipe_chunks.txt
