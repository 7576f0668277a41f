#!/bin/bash
# L1 loads for the next image at N = 120 (default on): N=120 tests, then cfg5 streams bench new vs old
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -x -q -m gpu --timeout 600 -k "120 or stream or cfg5 or push or raw or seq or steady" > gpurun_out/r4j_pytest.log 2>&1
grep -E "passed|failed" gpurun_out/r4j_pytest.log | tail -n 1; grep "^FAILED\|Error" gpurun_out/r4j_pytest.log | head -8 | cut -c1-200
for i in 1 2; do
echo "new: $(timeout 300 python bench.py --workload streams --steps 5 --warmup 3 2>/dev/null | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["value"], d.get("e2e",{}).get("value"))')"
echo "old: $(MOF_B200_LIB=$PWD/ab_variant.so timeout 300 python bench.py --workload streams --steps 5 --warmup 3 2>/dev/null | python -c 'import json,sys; d=json.loads(sys.stdin.read()); print(d["value"], d.get("e2e",{}).get("value"))')"
done | tee gpurun_out/r4j_ab_streams.txt
