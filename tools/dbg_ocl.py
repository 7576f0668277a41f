"""GPU-side debugging aid: where does ocl_compat differ from the oracle's restatement?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
from helpers import expand_bank
from mrs_optic_flow_b200 import api, synth
from oracle import oracle

for F, N, n in ((240, 60, 24), (240, 120, 500), (240, 240, 460)):
    bank_c, bank_p, _ = synth.shifted_pairs(min(n, 32), F, N, seed=1300 + N, max_shift=min(N / 4.0, 40.0))
    both = expand_bank(np.stack([bank_c, bank_p], axis=1), n, N, seed=5)
    curr, prev = np.ascontiguousarray(both[:, 0]), np.ascontiguousarray(both[:, 1])
    be = api.Backend(F, N, 80.0, max_batch=64, ocl_compat=True)
    got = be.fft_pairs(curr, prev)
    be.close()
    exp = oracle.fft_pairs(curr, prev, N, max_px_speed=80.0, nthreads=os.cpu_count(), ocl_compat=True)
    nm = np.argwhere(np.isnan(got).any(axis=2) != np.isnan(exp).any(axis=2))
    err = np.nan_to_num(np.abs(got - exp).max(axis=2))
    big = np.argwhere(err >= 0.01)
    print(f"F={F} N={N} n={n}: NaN mismatches {len(nm)}, >=0.01: {len(big)}, nan got {np.isnan(got).any(axis=2).sum()} exp {np.isnan(exp).any(axis=2).sum()}")
    for p, s_ in list(nm[:6]) + list(big[:6]):
        print("   pair", p, "sec", s_, "cuda", got[p, s_], "oracle", exp[p, s_])
