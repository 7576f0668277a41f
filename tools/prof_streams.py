"""Stream-mode side workloads of bench.py on their own (A/B of kernel variants through MOF_B200_LIB):
N=240 stream mode (4096 streams, one frame per step), push_many at N=240, 64-stream push at N=120, raw cfg5 push."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
from mrs_optic_flow_b200 import api

dev = torch.device("cuda", 0)
for fn in (bench.stream240_workload, bench.stream240_many_workload, bench.stream_workload, bench.raw_stream_workload):
    w = fn(torch, api, 0, dev)
    print({k: (round(v, 1) if isinstance(v, float) else v) for k, v in w.items() if k in ("workload", "frames_per_s", "ms_per_step")})
