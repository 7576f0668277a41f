#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_round2.py tests/test_adapter_cpp.py -m gpu -q -k "multi_context or adapter" > gpurun_out/r3k_pytest.log 2>&1
grep -E "^(FAILED|ERROR)|passed|failed|^E " gpurun_out/r3k_pytest.log | head -20 | cut -c1-250
