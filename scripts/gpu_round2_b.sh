#!/bin/bash
# fused kernel bring-up: the new parity test, the whole GPU suite, bench at N=1 with W variants
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "fused" > gpurun_out/fused_tests.log 2>&1; echo "fused tests rc=$?"
tail -25 gpurun_out/fused_tests.log
timeout 1500 python -m pytest tests -m gpu -x -q --durations=5 > gpurun_out/gpu_tests.log 2>&1; echo "tests rc=$?"
tail -12 gpurun_out/gpu_tests.log
for w in 0 9; do
RADOP_FUSED_WARPS=$w timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --no-direct > gpurun_out/bench_w$w.json 2> gpurun_out/bench_w$w.err; echo "bench w=$w rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/bench_w$w.json').read().strip().splitlines()[-1])
print('w=$w value %.3e ms %.3f e2e ms %.3f'%(d['value'], d['ms_per_step'], d['e2e']['ms_per_step']), d['stages_ms'])
PY
tail -3 gpurun_out/bench_w$w.err
done
