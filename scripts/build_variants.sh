#!/bin/bash
# Build-time variants of the fused kernel for A/B runs on the GPU box (RADOP_LIB=<path> python bench.py ...).
cd "$(dirname "$0")/../cosmo_pol_b200/csrc" || exit 1
F="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared -Xcompiler -fPIC"
rm -f ../libradop_*.so
nvcc $F -DFUSED_NOBARRIER=0 -o ../libradop_bar64.so radop_api.cu &
nvcc $F -DFUSED_MAX_WARPS=9 -DFUSED_MIN_CTAS=4 -DFUSED_NOBARRIER=0 -o ../libradop_bar56.so radop_api.cu &
nvcc $F -DFUSED_MAX_WARPS=9 -DFUSED_MIN_CTAS=4 -o ../libradop_r56.so radop_api.cu &
nvcc $F -DFUSED_MAX_WARPS=9 -DFUSED_MIN_CTAS=5 -o ../libradop_r40.so radop_api.cu &
wait
nvcc $F -DFUSED_MAX_WARPS=9 -DFUSED_MIN_CTAS=4 -DFUSED_UNROLL_COLS=1 -o ../libradop_r56u.so radop_api.cu &
nvcc $F -DFUSED_UNROLL_COLS=1 -o ../libradop_r64u.so radop_api.cu &
wait
ls -la ../libradop*.so
