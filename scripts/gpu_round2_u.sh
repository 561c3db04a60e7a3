#!/bin/bash
mkdir -p gpurun_out
for lib in cosmo_pol_b200/libradop.so cosmo_pol_b200/libradop_w2.so cosmo_pol_b200/libradop.so cosmo_pol_b200/libradop_w2.so; do
  RADOP_LIB=$PWD/$lib timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --no-direct > gpurun_out/bench_v.json 2> gpurun_out/bench_v.err || tail -3 gpurun_out/bench_v.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/bench_v.json').read().strip().splitlines()[-1])
print('$lib value %.3e ms %.3f e2e ms %.3f (rep %.3f) fused %.3f'%(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e']['ms_per_step_repeated_scan'], d['stages_ms']['interp_psd']))
PY
done
RADOP_LIB=$PWD/cosmo_pol_b200/libradop_w2.so timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "guessed or non_monotone or indices or fused or full_volume or observables or goldens" > gpurun_out/u_tests.log 2>&1; echo "w2 tests rc=$?"; tail -3 gpurun_out/u_tests.log
