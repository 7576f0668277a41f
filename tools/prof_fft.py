"""Small driver for ncu: runs the batched FftMethod kernel on device-resident synthetic pairs.
usage: python tools/prof_fft.py [N=240] [F=240] [pairs=592] [launches=3]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from bench import device_batch, make_inputs_host
from mrs_optic_flow_b200 import api

N = int(sys.argv[1]) if len(sys.argv) > 1 else 240
F = int(sys.argv[2]) if len(sys.argv) > 2 else 240
n = int(sys.argv[3]) if len(sys.argv) > 3 else 592
launches = int(sys.argv[4]) if len(sys.argv) > 4 else 3
dev = torch.device("cuda", 0)
hc, hp = make_inputs_host(32, F, seed=1)
dc, dp = device_batch(torch, hc, hp, n, dev)
be = api.Backend(F, N, 80.0)
out = torch.empty((n, be.S, 2), dtype=torch.float64, device=dev)
be.enable_kernel_timing(True)
for _ in range(launches):
    be.fft_pairs(dc, dp, out=out)
torch.cuda.synchronize()
ms, k = be.fft_kernel_time()
print(f"N={N} F={F} pairs={n}: {ms:.3f} ms/launch over {k} launches -> {n / ms * 1e3:.0f} pairs/s; "
      f"finite {float(torch.isfinite(out).all(dim=(1, 2)).float().mean()):.3f}")
be.close()
