#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -3 gpurun_out/noalloc_alone.log
timeout 900 python scripts/cfg5_diag.py > gpurun_out/cfg5_diag.log 2>&1; echo "cfg5 diag rc=$?"; tail -40 gpurun_out/cfg5_diag.log
for lib in cosmo_pol_b200/libradop.so cosmo_pol_b200/libradop_b*.so; do
  RADOP_LIB=$PWD/$lib timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --no-direct > gpurun_out/bench_v.json 2> gpurun_out/bench_v.err || tail -3 gpurun_out/bench_v.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/bench_v.json').read().strip().splitlines()[-1])
print('$lib value %.3e ms %.3f e2e ms %.3f fused %.3f'%(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d['stages_ms']['interp_psd']))
PY
done
timeout 2400 python -m pytest tests -m gpu -q --durations=5 > gpurun_out/gpu_tests.log 2>&1; echo "tests rc=$?"
tail -12 gpurun_out/gpu_tests.log
