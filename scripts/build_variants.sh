#!/bin/bash
# Build-time variants of the fused kernel for A/B runs on the GPU box:  RADOP_LIB=$PWD/cosmo_pol_b200/libradop_<tag>.so python bench.py ...
# The default build (libradop.so, __graft_entry__.build()) is 9 warps x 3 CTAs (72 registers), unrolled columns, two row buffers.
# Measured variants: profiles/r02_fused_variants.md.  The variant libraries travel to the GPU box with the snapshot; remove them
# after an ABI change (rm cosmo_pol_b200/libradop_*.so) -- _abi.py checks the struct sizes of whatever it loads.
cd "$(dirname "$0")/../cosmo_pol_b200/csrc" || exit 1
F="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -shared -Xcompiler -fPIC"
rm -f ../libradop_*.so
nvcc $F -DFUSED_MIN_CTAS=4 -o ../libradop_r56.so radop_api.cu &          # 56 registers, 4 CTAs per SM
nvcc $F -DFUSED_MIN_CTAS=2 -o ../libradop_r96.so radop_api.cu &          # 96 registers, 2 CTAs per SM
nvcc $F -DFUSED_DBUF=0 -o ../libradop_1buf.so radop_api.cu &             # one row buffer, two barriers per piece
nvcc $F -DFUSED_UNROLL_COLS=0 -o ../libradop_rolled.so radop_api.cu &    # rolled loop over the four neighbour columns
wait
ls -la ../libradop*.so
