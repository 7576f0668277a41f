"""Small driver for ncu: the side kernels (log-polar, FastSpacedBM, front end) on device-resident
synthetic inputs.  usage: python tools/prof_side.py [pairs=296]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from bench import device_batch, make_inputs_host
from mrs_optic_flow_b200 import api, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 296
dev = torch.device("cuda", 0)


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


hc, hp = make_inputs_host(32, 240, seed=3)
dc, dp = device_batch(torch, hc, hp, n, dev)
be = api.Backend(240, 240, 1e9, sr_resolution=240, sr_m=49.9, max_batch=n)
ms = timed(lambda: be.scalerot_pairs(dc, dp))
print(f"scalerot: {n} pairs, {ms:.3f} ms -> {n / ms * 1e3:.0f} pairs/s")
be.close()

c, p, _ = synth.bm_pairs(16, 240, 21, seed=2)
dc, dp = device_batch(torch, c, p, n, dev)
be = api.Backend(240, 240, 1e9, bm_block=120, bm_step=24, bm_radius=21)
ms = timed(lambda: be.bm_pairs(dc, dp))
print(f"bm: {n} pairs, {ms:.3f} ms -> {n / ms * 1e3:.0f} blocks/s")
be.close()

raw = torch.randint(0, 256, (64, 480, 752, 3), dtype=torch.uint8, device=dev)
be = api.Backend(480, 120, 80.0, n_streams=64)
gray = torch.empty((64, 480, 480), dtype=torch.uint8, device=dev)
ms = timed(lambda: be.frontend(raw, 1, 376, out=gray))
print(f"front end: 64 frames 752x480x3 -> 480^2 gray, {ms * 1e3:.1f} us -> {64 * 480 * 480 * 4 / ms / 1e6:.0f} GB/s")
ids = np.arange(64, dtype=np.int32)
ms = timed(lambda: be.fft_stream_push_raw(ids, raw, 1, 376))
print(f"raw push: 64 frames, {ms * 1e3:.1f} us")
be.close()
