"""The header-only C++ adapter (include/mof_opticflowcalc.hpp) compiled against stand-in OpenCV /
OpticFlowCalc headers (tests/stubs/opencv_stub) and linked to libmof_b200.so."""
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "stubs", "adapter_check")


def _build():
    libdir = os.path.join(ROOT, "mrs_optic_flow_b200")
    if not os.path.exists(os.path.join(libdir, "libmof_b200.so")):
        subprocess.check_call(["make", "-C", os.path.join(libdir, "csrc"), "-j4", "-s"])
    subprocess.check_call(
        ["g++", "-std=c++17", "-Wall", "-Werror", "-I", os.path.join(ROOT, "tests", "stubs", "opencv_stub"),
         "-I", os.path.join(ROOT, "include"), os.path.join(ROOT, "tests", "stubs", "adapter_check.cpp"),
         "-o", EXE, "-L", libdir, "-lmof_b200", f"-Wl,-rpath,{libdir}"])


def test_adapter_compiles_and_fails_loudly_without_gpu():
    import torch

    _build()
    if torch.cuda.is_available():
        pytest.skip("GPU present: covered by the gpu test")
    r = subprocess.run([EXE], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "no CPU fallback" in r.stdout


def _lcg_frame():
    """The frame adapter_check.cpp generates (LCG bytes), and its (+3, -2) circular shift."""
    import numpy as np

    s, vals = 12345, []
    for _ in range(240 * 240):
        s = (s * 1664525 + 1013904223) & 0xFFFFFFFF
        vals.append(s >> 24)
    a = np.array(vals, np.uint8).reshape(240, 240)
    return a, np.roll(a, (-2, 3), axis=(0, 1))


@pytest.mark.gpu
def test_adapter_runs_through_plugin_interface(oracle_mod):
    import numpy as np

    _build()
    r = subprocess.run([EXE, "gpu"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    lines = r.stdout.splitlines()
    assert lines[0] == "sections 4"
    first = np.array([[float(v) for v in ln.split()[1:]] for ln in lines if ln.startswith("first")])
    second = np.array([[float(v) for v in ln.split()[1:]] for ln in lines if ln.startswith("second")])
    a, b = _lcg_frame()
    # first call: prev := curr (src/FftMethod.cpp:1791-1793); second: what the reference CPU path gives
    exp1 = oracle_mod.fft_pairs(a[None], a[None], 120, max_px_speed=80.0)[0]
    exp2 = oracle_mod.fft_pairs(b[None], a[None], 120, max_px_speed=80.0)[0]
    assert np.abs(first - exp1).max() < 0.01 and np.abs(second - exp2).max() < 0.01
    assert np.abs(second - np.array([3.0, -2.0])).max() < 0.2  # content moved by (+3, -2) px
    # the black frame given to setImPrev before the first call was overwritten (zeros, not NaN); a later
    # setImPrev(b) is used: third = a against b, fourth = b against a again
    assert np.abs(first).max() < 1e-5
    third = np.array([[float(v) for v in ln.split()[1:]] for ln in lines if ln.startswith("third")])
    fourth = np.array([[float(v) for v in ln.split()[1:]] for ln in lines if ln.startswith("fourth")])
    exp3 = oracle_mod.fft_pairs(a[None], b[None], 120, max_px_speed=80.0)[0]
    assert np.abs(third - exp3).max() < 0.01 and np.abs(fourth - exp2).max() < 0.01
    assert "scalerot first (1.000, 0.000)" in r.stdout
    assert "longrange 0" in lines
    lr = np.array([[float(v) for v in ln.split()[1:]] for ln in lines if ln.startswith("lr60")])
    exp_lr = oracle_mod.fft_pairs_longrange(b[None], a[None], 60, max_px_speed=80.0)[0]
    assert lr.shape == exp_lr.shape == (1, 2)
    assert np.array_equal(np.isnan(lr), np.isnan(exp_lr)) and np.nanmax(np.abs(lr - exp_lr), initial=0.0) < 0.01
    # constructor fallback (one 240 px section) and a section size outside the prime-factor plan (3 x 3 of 80 px)
    fb = np.array([[float(v) for v in ln.split()[1:]] for ln in lines if ln.startswith("fallback240")])
    g80 = np.array([[float(v) for v in ln.split()[1:]] for ln in lines if ln.startswith("generic80")])
    assert fb.shape == (1, 2) and g80.shape == (9, 2)
    assert np.abs(fb - oracle_mod.fft_pairs(b[None], a[None], 100, max_px_speed=80.0)[0]).max() < 0.01
    assert np.abs(g80 - oracle_mod.fft_pairs(b[None], a[None], 80, max_px_speed=80.0)[0]).max() < 0.01
    # BlockMethodCuda: vote (stride-60 SAD search, radius 8, no gate) then Refine, against the oracle
    bm = [float(v) for v in next(ln for ln in lines if ln.startswith("blockmethod")).split()[1:]]
    black = np.zeros_like(a)
    v1 = oracle_mod.blockmethod(a, black, 60, 8)[2]
    v2 = oracle_mod.blockmethod(b, a, 60, 8)[2]
    assert (int(v2[0]), int(v2[1])) == (-3, 2)   # where the block was in imPrev (src/BlockMethod.cpp:52-54)
    e1 = oracle_mod.refine(a, [[int(v1[0]), int(v1[1])]])[0]
    e2 = oracle_mod.refine(b, [[int(v2[0]), int(v2[1])]])[0]
    assert bm == [e1[0], e1[1], e2[0], e2[1]]
