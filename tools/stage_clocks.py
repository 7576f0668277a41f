"""Per-stage cycle accounting of the phase-correlation kernel: builds a -DMOF_STAGE_CLOCKS variant of
the backend into a scratch directory, runs the N=240 batched workload and prints, per warp of CTA 0,
the cycles spent between stage boundaries.  usage (GPU box): python tools/stage_clocks.py [N] [F]"""
import ctypes
import os
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
N = int(sys.argv[1]) if len(sys.argv) > 1 and sys.argv[1] != "build" else 240
F = int(sys.argv[2]) if len(sys.argv) > 2 else 240
tmp = tempfile.mkdtemp()
src = os.path.join(ROOT, "mrs_optic_flow_b200", "csrc")
so = os.path.join(tmp, "libmof_b200.so")
files = ["pc_pfa_kernel.cu", "bm_kernels.cu", "lp_kernels.cu", "mof_capi.cu"]
extra = [f for f in os.listdir(src) if f.endswith(".cu") and f not in files]
prebuilt = os.path.join(ROOT, "tools", "_build", "libmof_clk.so")   # python tools/stage_clocks.py build (no GPU needed)
if len(sys.argv) > 1 and sys.argv[1] == "build":
    os.makedirs(os.path.dirname(prebuilt), exist_ok=True)
    so = prebuilt
if os.path.exists(prebuilt) and so != prebuilt:
    so = prebuilt
else:
  subprocess.check_call(["nvcc", "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-DMOF_STAGE_CLOCKS",
                       "-Xcompiler", "-fPIC", "-shared", "-o", so] + [os.path.join(src, f) for f in files + extra])
if len(sys.argv) > 1 and sys.argv[1] == "build":
    sys.exit(0)
import numpy as np
import torch
from bench import device_batch, make_inputs_host
from mrs_optic_flow_b200 import _lib

_lib.LIB_PATH = so   # the instrumented variant instead of the in-tree library
from mrs_optic_flow_b200 import api

L = _lib.load()
dev = torch.device("cuda", 0)
secs_per_pair = (F // N) ** 2
resident = {240: 148, 120: 148 * 3, 60: 148 * 15}[N]
n = max(1, resident * 16 // secs_per_pair)
hc, hp = make_inputs_host(32, F, seed=1)
dc, dp = device_batch(torch, hc, hp, n, dev)
be = api.Backend(F, N, 80.0)
out = torch.empty((n, be.S, 2), dtype=torch.float64, device=dev)
buf = (ctypes.c_ulonglong * (32 * 24))()
be.fft_pairs(dc, dp, out=out)
torch.cuda.synchronize()
L.mof_debug_stage_clocks(buf, 1)
be.fft_pairs(dc, dp, out=out)
torch.cuda.synchronize()
assert L.mof_debug_stage_clocks(buf, 0) == 0
clk = np.array(buf[:], dtype=np.float64).reshape(32, 24)
names = ["loop/tail", "FX15", "FX16", "bar R", "prefetch", "FY16", "FY15 store", "FY15 corr", "IY16", "bar C",
         "prefetch2", "IX15", "IX16", "search", "bar B1", "px ld+stg", "px words"]
secs = 16 if N == 240 else None
print(f"N={N} F={F}: cycles per stage, CTA 0, summed over its sections (per warp); last column = total")
print("(FX15 = butterflies only; its pixel fetch + staging and word pick-up are the last two columns)")
print("warp " + " ".join(f"{x[:9]:>9s}" for x in names[:17]) + "     total")
for w in range(32):
    if clk[w].sum() == 0:
        continue
    print(f"{w:4d} " + " ".join(f"{v:9.0f}" for v in clk[w, :17]) + f" {clk[w].sum():9.0f}")
be.close()
shutil.rmtree(tmp, ignore_errors=True)
