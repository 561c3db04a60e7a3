#!/bin/bash
# GPU-box check: parity tests, bench, ncu launch list + one full capture of the top kernel.
# usage (under gpurun): bash scripts/gpu_check.sh [tests|bench|ncu|all]
what=${1:-all}
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/smi.txt 2>&1
if [[ $what == all || $what == tests ]]; then
  timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "tests rc=$?"
  tail -3 gpurun_out/gpu_tests.log
fi
if [[ $what == all || $what == bench ]]; then
  timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
  tail -c 3000 gpurun_out/bench.json; tail -5 gpurun_out/bench.err
fi
if [[ $what == all || $what == ncu ]]; then
  CMD="python bench.py --steps 2 --warmup 3 --no-cpu"
  timeout 600 $CMD > gpurun_out/plain.log 2>&1 &&
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv \
      --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
  echo "ncu list rc=$?"
  timeout 600 $CMD > gpurun_out/plain2.log 2>&1 &&
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_interp_psd -s 4 -c 1 \
      -f -o gpurun_out/prof_k1 $CMD > gpurun_out/ncu_full.log 2>&1
  echo "ncu full rc=$?"
fi
if [[ $what == all || $what == ncu ]]; then
  timeout 600 $CMD > gpurun_out/plain3.log 2>&1 &&
  timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_integrate_tab -s 3 -c 1 \
      -f -o gpurun_out/prof_k2t $CMD > gpurun_out/ncu_full2.log 2>&1
  echo "ncu full (k_integrate_tab) rc=$?"
fi
