#!/bin/bash
# scratch (stream) variants without the OCL-compatible epilogue / window option compiled in: A = in-tree, B = -DMOF_NO_EXTRAS
mkdir -p gpurun_out
for i in 1 2; do
echo "A (in-tree):"; timeout 300 python tools/prof_streams.py 2>&1 | tail -4
echo "B (no extras):"; MOF_B200_LIB=$PWD/ab_variant.so timeout 300 python tools/prof_streams.py 2>&1 | tail -4
done | tee gpurun_out/r4q_ab_no_extras.txt
