#!/bin/bash
# last check of the round: full GPU suite and smoke on HEAD
mkdir -p gpurun_out
timeout 900 python -m pytest tests -x -q -m gpu --timeout 600 > gpurun_out/r5n_pytest.log 2>&1
grep -E "passed|failed" gpurun_out/r5n_pytest.log | tail -n 1; grep "^FAILED" gpurun_out/r5n_pytest.log | head -5 | cut -c1-200
timeout 300 python __graft_entry__.py smoke > gpurun_out/r5n_smoke.log 2>&1; tail -n 1 gpurun_out/r5n_smoke.log | cut -c1-300
