#!/bin/bash
mkdir -p gpurun_out
for lib in cosmo_pol_b200/libradop.so cosmo_pol_b200/libradop_c2.so cosmo_pol_b200/libradop.so cosmo_pol_b200/libradop_c2.so; do
  RADOP_LIB=$PWD/$lib timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --no-direct > gpurun_out/bench_v.json 2> gpurun_out/bench_v.err || tail -3 gpurun_out/bench_v.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/bench_v.json').read().strip().splitlines()[-1])
print('$lib value %.3e ms %.3f e2e ms %.3f (rep %.3f) fused %.3f'%(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e']['ms_per_step_repeated_scan'], d['stages_ms']['interp_psd']))
PY
done
timeout 600 python scripts/bench_configs.py cfg5s cfg5 --steps 5 > gpurun_out/cfg5_configs.json 2> gpurun_out/cfg5_configs.err; echo "configs rc=$?"
python - <<'PY'
import json
for l in open('gpurun_out/cfg5_configs.json'):
    try: d = json.loads(l)
    except Exception: continue
    print(d['config'], 'device %.2f ms e2e %.2f ms' % (d['ms_per_scan_device'], d['ms_per_scan_e2e']), d['stages_ms'])
PY
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "spectrum or cfg5" > gpurun_out/p_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/p_tests.log
