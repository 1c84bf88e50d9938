#!/bin/bash
# Build tuning variants of the library side by side (A/B runs on the GPU box):
#   scripts/ab_build.sh NAME "-DLS_WARPS_N=12 ..."   ->  nautilus_b200/lib/variants/NAME.so
set -e
cd "$(dirname "$0")/.."
mkdir -p nautilus_b200/lib/variants
S=nautilus_b200/csrc
nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo --fmad=false -Xcompiler -fPIC -shared -DNAUTILUS_CSM_TUNING $2 \
  $S/csm_tables.cu $S/csm_search.cu $S/csm_capi.cu $S/csm_comm.cu $S/csm_image_ops.cu $S/nn_kernels.cu $S/nn_capi.cu $S/csm_uncertainty.cu $S/frontend.cu $S/residuals.cu \
  -o nautilus_b200/lib/variants/$1.so -ldl
