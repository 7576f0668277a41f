#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_generic_sizes.py -m gpu -q > gpurun_out/r3b_pytest.log 2>&1
grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/r3b_pytest.log | cut -c1-200
bash tools/gpu_r3a.sh
