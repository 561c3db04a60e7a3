#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "antenna or cfg5 or spectrum" > gpurun_out/i_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/i_tests.log
bash scripts/gpu_round2_multi.sh "1 2"
tail -30 gpurun_out/sharded_test_2gpu.log | grep -v "^$" | tail -12
