#!/bin/bash
mkdir -p gpurun_out
# (compute-sanitizer is closed on this pool: the script runs plain; bounds are argued in fe_kernels.cu / DESIGN §4.4)
timeout 300 python tools/san_new.py > gpurun_out/r3m_plain.log 2>&1
tail -n 4 gpurun_out/r3m_plain.log | cut -c1-200
