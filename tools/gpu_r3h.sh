#!/bin/bash
mkdir -p gpurun_out
{
for lib in "" "$PWD/ab_variant.so"; do
MOF_B200_LIB=$lib python - <<'PY'
import os, sys; sys.path.insert(0, '.')
import torch, bench
from mrs_optic_flow_b200 import api
dev = torch.device("cuda", 0)
tag = "B (no spectrum hints)" if os.environ.get("MOF_B200_LIB") else "A (L2 hints for the next step's cached spectrum)"
print(tag, bench.stream240_workload(torch, api, 0, dev))
print(tag, bench.stream_workload(torch, api, 0, dev))
print(tag, bench.raw_stream_workload(torch, api, 0, dev))
PY
done
} > gpurun_out/r3h_spec_prefetch_ab.txt 2>&1
cut -c1-300 gpurun_out/r3h_spec_prefetch_ab.txt
