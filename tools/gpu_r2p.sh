#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_steady_state.py -m gpu -q --timeout 900 -k "240" > gpurun_out/r2p_pytest.log 2>&1
{
for rep in 1 2; do
  echo "A (column stages interleaved on half of the warps): $(timeout 120 python tools/prof_fft.py 240 240 8192 10 2>&1 | tail -1)"
  echo "B (all warps in the same stage):                    $(MOF_B200_LIB=$PWD/ab_nomix.so timeout 120 python tools/prof_fft.py 240 240 8192 10 2>&1 | tail -1)"
done
} > gpurun_out/r2p_ab.txt 2>&1
timeout 300 python tools/stage_clocks.py 240 240 > gpurun_out/r2p_stage_clocks_pc240.txt 2>&1
tail -n 1 gpurun_out/r2p_pytest.log; cat gpurun_out/r2p_ab.txt; sed -n 4,9p gpurun_out/r2p_stage_clocks_pc240.txt | cut -c1-200
