"""Shifts of one batch of synthetic pairs into an .npy file (A/B of two builds for bitwise equality).
usage: MOF_B200_LIB=... python tools/pg_dump.py N F pairs out.npy"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

from bench import device_batch, make_inputs_host
from mrs_optic_flow_b200 import api

N, F, n, path = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), sys.argv[4]
dev = torch.device("cuda", 0)
hc, hp = make_inputs_host(32, F, seed=1)
dc, dp = device_batch(torch, hc, hp, n, dev)
be = api.Backend(F, N, 80.0)
out = torch.empty((n, be.S, 2), dtype=torch.float64, device=dev)
be.fft_pairs(dc, dp, out=out)
torch.cuda.synchronize()
np.save(path, out.cpu().numpy())
