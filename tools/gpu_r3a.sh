#!/bin/bash
mkdir -p gpurun_out
python - > gpurun_out/r3a_generic_sizes.txt 2>&1 <<'PY'
import sys; sys.path.insert(0, '.')
import torch, bench
from mrs_optic_flow_b200 import api
dev = torch.device("cuda", 0)
for F, N, n in ((480, 480, 1024), (480, 160, 1024), (480, 96, 512), (240, 80, 2048), (240, 30, 1024), (768, 768, 512)):
    r = bench.side_workload(torch, api, 0, dev, f"generic F={F} N={N}", F, N, n)
    print({k: (round(v, 4) if isinstance(v, float) else v) for k, v in r.items()})
PY
cut -c1-400 gpurun_out/r3a_generic_sizes.txt
