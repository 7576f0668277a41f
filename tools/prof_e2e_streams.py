"""The two host-frame push_many lines of bench.py on their own (sweeps of host-pipeline switches):
e2e stream mode F=240 N=240 (gray frames) and, with 'raw', cfg5's raw 752x480 BGR8 frames."""
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

import bench
from mrs_optic_flow_b200 import api

dev = torch.device("cuda", 0)
w = bench.stream240_e2e_workload(torch, api, 0, dev)
print({k: (round(v, 1) if isinstance(v, float) else v) for k, v in w.items() if k in ("frames_per_s", "h2d_gbs")})
