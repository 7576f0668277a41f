"""Small run of the round-2 additions for compute-sanitizer memcheck: 16-pixel front end (tight crop staging and
device frames), generic-size kernel (pairs, stream push, cache update), push_many with the spectrum chain in
tensor memory, crop-window host pipeline.  usage: python tools/san_new.py"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from mrs_optic_flow_b200 import api, synth

rng = np.random.default_rng(0)
# front end: device frames (fast path needs slack inside the frame) and host frames (tight crop + FE_SLACK)
raw = rng.integers(0, 256, (3, 4, 480, 752, 3), dtype=np.uint8)
be = api.Backend(480, 120, 80.0, n_streams=3, max_batch=6)
ids = np.arange(3, dtype=np.int32)
a = be.fft_stream_push_many_raw(ids, raw, 1, 376)
b = be.fft_stream_push_many_raw(ids, torch.from_numpy(raw).cuda(), 1, 376).cpu().numpy()
g = be.frontend(raw[:, 0], 1, 376)
print("front end / raw pipeline", a.shape, b.shape, g.shape, np.isfinite(a[:, 1:]).mean())
be.close()
# last frame of an allocation: crop ends in the frame's last row
last = torch.from_numpy(raw[0, 0]).cuda()[None].contiguous()
be = api.Backend(480, 120, 80.0)
print("front end, exact-size device allocation", be.frontend(last, 1, 376).shape)
be.close()
# generic sizes
for F, N in ((160, 80), (480, 480), (90, 30)):
    c, p, _ = synth.shifted_pairs(5, F, N, seed=N, max_shift=5.0)
    be = api.Backend(F, N, 80.0, n_streams=5, max_batch=5)
    r = be.fft_pairs(c, p)
    ids5 = np.arange(5, dtype=np.int32)
    be.fft_stream_push(ids5, c)
    s = be.fft_stream_push(ids5, p)
    m = be.fft_stream_push_many(ids5, np.stack([c, p, c], axis=1))
    print("generic", F, N, r.shape, s.shape, m.shape, np.isfinite(r).mean())
    be.close()
# push_many, N = 240: chain in tensor memory, several rounds per slot
frames, _ = synth.camera_streams(8, 6, 240, 240, seed=3, max_speed=4.0)
frames = np.tile(frames, (40, 1, 1, 1))
be = api.Backend(240, 240, 80.0, n_streams=320, max_batch=320 * 6)
m = be.fft_stream_push_many(np.arange(320, dtype=np.int32), frames)
m2 = be.fft_stream_push_many(np.arange(320, dtype=np.int32), frames[:, :2])
print("push_many 240", m.shape, np.isfinite(m).mean(), m2.shape)
be.close()
print("done")
