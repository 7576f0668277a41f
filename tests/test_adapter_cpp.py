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


@pytest.mark.gpu
def test_adapter_runs_through_plugin_interface():
    _build()
    r = subprocess.run([EXE, "gpu"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    # first call: zero shifts; second call: content moved by (+3, -2) px in every section
    line = r.stdout.splitlines()[0]
    assert "sections 4" in line
    nums = [float(x) for x in re.findall(r"-?\d+\.\d+", line)]
    assert abs(nums[0]) < 1e-4 and abs(nums[1]) < 1e-4
    assert abs(nums[2] - 3.0) < 0.01 and abs(nums[3] + 2.0) < 0.01
    assert "scalerot first (1.000, 0.000)" in r.stdout
