#!/bin/bash
# Build-time variants of the fused kernel for A/B runs on the GPU box (RADOP_LIB=<path> python bench.py ...).
# The default build (libradop.so, __graft_entry__.build()) is 9 warps x 4 CTAs (56 registers) with the CTA barrier.
cd "$(dirname "$0")/../cosmo_pol_b200/csrc" || exit 1
F="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared -Xcompiler -fPIC"
rm -f ../libradop_*.so
nvcc $F -DFUSED_MAX_WARPS=9 -DFUSED_MIN_CTAS=3 -o ../libradop_bar72.so radop_api.cu &
nvcc $F -DFUSED_MAX_WARPS=10 -DFUSED_MIN_CTAS=3 -o ../libradop_bar64.so radop_api.cu &
nvcc $F -DFUSED_MAX_WARPS=9 -DFUSED_MIN_CTAS=5 -o ../libradop_bar40.so radop_api.cu &
nvcc $F -DFUSED_MAX_WARPS=9 -DFUSED_MIN_CTAS=4 -DFUSED_UNROLL_COLS=1 -o ../libradop_bar56u.so radop_api.cu &
wait
ls -la ../libradop*.so
