#!/bin/bash
# full ncu capture of one kernel of the bench step: bash scripts/ncu_kernel.sh <kernel regex> <out name>
k=${1:-k_integrate}; o=${2:-prof}; skip=${3:-3}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu"
mkdir -p gpurun_out
timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:$k -s $skip -c 1 \
    -f -o gpurun_out/$o $CMD > gpurun_out/ncu_$o.log 2>&1
echo "ncu rc=$?"
