#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_generic_sizes.py tests/test_gpu_parity.py -m gpu -q -k "scalerot or logpolar" > gpurun_out/r3j_pytest.log 2>&1
grep -E "^(FAILED|ERROR)|passed|failed|^E " gpurun_out/r3j_pytest.log | head -20 | cut -c1-250
