"""Driver for ncu: the round-2 kernels that have no profile yet -- fe_bgr16_kernel (1024 camera frames) and
pg_kernel (generic section size, F = N = 480).  usage: python tools/prof_new.py [once]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from bench import device_batch, make_inputs_host
from mrs_optic_flow_b200 import api

once = len(sys.argv) > 1 and sys.argv[1] == "once"
dev = torch.device("cuda", 0)


def timed(fn, reps=3):
    if once:
        fn()
        torch.cuda.synchronize()
        return float("nan")
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


nf = 1024
raw = torch.randint(0, 256, (nf, 480, 752, 3), dtype=torch.uint8, device=dev)
be = api.Backend(480, 120, 80.0)
gray = torch.empty((nf, 480, 480), dtype=torch.uint8, device=dev)
ms = timed(lambda: be.frontend(raw, 1, 376, out=gray))
print(f"front end: {nf} frames 752x480x3 -> 480^2 gray, {ms * 1e3:.1f} us -> {nf * 480 * 480 * 4 / ms / 1e6:.0f} GB/s algorithmic")
be.close()
del raw, gray

n = 444
hc, hp = make_inputs_host(16, 480, seed=3)
dc, dp = device_batch(torch, hc, hp, n, dev)
be = api.Backend(480, 480, 80.0)
out = torch.empty((n, 1, 2), dtype=torch.float64, device=dev)
ms = timed(lambda: be.fft_pairs(dc, dp, out=out))
print(f"generic N=480: {n} pairs, {ms:.3f} ms -> {n / ms * 1e3:.0f} pairs/s")
be.close()
