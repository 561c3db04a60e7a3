#!/bin/bash
# Build-time variants of the fused kernel for A/B runs on the GPU box (RADOP_LIB=<path> python bench.py ...).
# The default build (libradop.so, __graft_entry__.build()) is 9 warps x 4 CTAs (56 registers), unrolled columns, one row buffer.
cd "$(dirname "$0")/../cosmo_pol_b200/csrc" || exit 1
F="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared -Xcompiler -fPIC"
rm -f ../libradop_*.so
nvcc $F -DFUSED_DBUF=1 -o ../libradop_b56u_dbuf.so radop_api.cu &
nvcc $F -DFUSED_MIN_CTAS=3 -o ../libradop_b72u.so radop_api.cu &
nvcc $F -DFUSED_MIN_CTAS=3 -DFUSED_DBUF=1 -o ../libradop_b72u_dbuf.so radop_api.cu &
nvcc $F -DFUSED_UNROLL_COLS=0 -DFUSED_DBUF=1 -o ../libradop_b56_dbuf.so radop_api.cu &
wait
ls -la ../libradop*.so
