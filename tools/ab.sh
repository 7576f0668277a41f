#!/bin/bash
# A/B of a kernel variant: builds the backend with extra nvcc flags ($1, e.g. -DMOF_NO_STAGGER) into
# gpurun_out/ab/ HERE (no GPU needed); on the GPU box run tools/ab_run.sh
set -e
cd "$(dirname "$0")/../mrs_optic_flow_b200/csrc"
mkdir -p ../../gpurun_out/ab
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Xcompiler -fPIC $1 -shared \
  -o ../../${AB_OUT:-ab_variant.so} pc_pfa_kernel.cu bm_kernels.cu lp_kernels.cu fe_kernels.cu rf_kernels.cu pg_kernels.cu mof_capi.cu mof_multi.cu
