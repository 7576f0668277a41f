#!/bin/bash
# round-2 second GPU pass: parity suite with the new features, A/B of kernel variants, streams bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -s > gpurun_out/r2b_pytest.log 2>&1
echo "pytest rc $?" >> gpurun_out/r2b_pytest.log
{
for cfg in "240 240 4736" "240 240 444" "120 480 2368" "60 480 2368" "60 240 4736"; do
  echo "A (in-tree):  $(timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
  echo "B (cp.async): $(MOF_B200_LIB=$PWD/ab_ldgsts.so timeout 120 python tools/prof_fft.py $cfg 8 2>&1 | tail -1)"
done
} > gpurun_out/r2b_ab.txt 2>&1
timeout 300 python tools/stage_clocks.py 240 240 > gpurun_out/r2b_stage_clocks_pc240.txt 2>&1
timeout 600 python bench.py --workload streams --steps 5 --warmup 3 > gpurun_out/r2b_bench_streams_1gpu.json 2> gpurun_out/r2b_bench_streams.err
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/r2b_bench_1gpu.json 2> gpurun_out/r2b_bench.err
tail -3 gpurun_out/r2b_pytest.log; cat gpurun_out/r2b_ab.txt
