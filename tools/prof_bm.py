"""FastSpacedBM kernel-only throughput on device-resident synthetic pairs (default geometry F=240 B=120 step=24 R=21,
and F=480: 3 x 3 blocks).  usage: python tools/prof_bm.py [pairs=4096]   (MOF_BM_NO_WIDE=1: previous kernel)"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from bench import device_batch
from mrs_optic_flow_b200 import api, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
dev = torch.device("cuda", 0)


def timed(fn, reps=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


for F, npairs in ((240, n), (480, n // 8)):
    c, p, _ = synth.bm_pairs(16, F, 21, seed=2)
    dc, dp = device_batch(torch, c, p, npairs, dev)
    be = api.Backend(F, F, 1e9, bm_block=120, bm_step=24, bm_radius=21, max_batch=npairs)
    nb = ((F - 42) // 144) ** 2
    ms = timed(lambda: be.bm_pairs(dc, dp))
    print(f"bm F={F}: {npairs} pairs x {nb} blocks, {ms:.3f} ms -> {npairs * nb / ms * 1e3:.0f} blocks/s "
          f"= {npairs * nb * 26.6256e6 / ms * 1e3 / 1e12:.1f} T pixel-ops/s")
    be.close()
