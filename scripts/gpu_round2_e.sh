#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "fused or melting or scan_level or table_vs" > gpurun_out/fused_tests.log 2>&1; echo "fused tests rc=$?"
tail -5 gpurun_out/fused_tests.log
for lib in cosmo_pol_b200/libradop*.so; do
  RADOP_LIB=$PWD/$lib timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --no-direct > gpurun_out/bench_v.json 2> gpurun_out/bench_v.err || tail -3 gpurun_out/bench_v.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/bench_v.json').read().strip().splitlines()[-1])
print('$lib value %.3e ms %.3f e2e ms %.3f fused %.3f'%(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d['stages_ms']['interp_psd']))
PY
done
timeout 2400 python -m pytest tests -m gpu -q --durations=8 > gpurun_out/gpu_tests.log 2>&1; echo "tests rc=$?"
tail -15 gpurun_out/gpu_tests.log
CMD="python bench.py --steps 2 --warmup 3 --no-cpu --no-direct"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:k_gate_fused -s 4 -c 1 -f -o gpurun_out/prof_fused $CMD > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_spectrum$ -s 1 -c 1 -f -o gpurun_out/prof_spectrum python scripts/bench_configs.py cfg5s --steps 2 > gpurun_out/ncu_spec.log 2>&1
echo "ncu spectrum rc=$?"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_default.json 2> gpurun_out/bench_default.err; echo "bench rc=$?"; cat gpurun_out/bench_default.json
timeout 900 python scripts/margins.py > gpurun_out/margins.json 2> gpurun_out/margins.err; echo "margins rc=$?"
