#!/bin/bash
# whole GPU suite, bench, ncu launch list and a full capture of k_gate_fused
mkdir -p gpurun_out
timeout 2400 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/gpu_tests.log 2>&1; echo "tests rc=$?"
tail -15 gpurun_out/gpu_tests.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_n1.json').read().strip().splitlines()[-1])
print('value %.3e ms %.3f e2e ms %.3f'%(d['value'], d['ms_per_step'], d['e2e']['ms_per_step']), d['stages_ms'])
PY
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-direct"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_gate_fused -s 4 -c 1 -f -o gpurun_out/prof_fused $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
timeout 900 python scripts/lut_stress.py > gpurun_out/lut_stress.json 2> gpurun_out/lut_stress.err; echo "stress rc=$?"; tail -3 gpurun_out/lut_stress.err
