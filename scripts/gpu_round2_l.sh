#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "spectrum or cfg5" > gpurun_out/l_tests.log 2>&1; echo "spectrum tests rc=$?"; tail -3 gpurun_out/l_tests.log
timeout 600 python scripts/bench_configs.py cfg5s cfg5 --steps 5 > gpurun_out/cfg5_configs.json 2> gpurun_out/cfg5_configs.err; echo "configs rc=$?"
python - <<'PY'
import json
for l in open('gpurun_out/cfg5_configs.json'):
    try: d = json.loads(l)
    except Exception: continue
    print(d['config'], 'device %.2f ms e2e %.2f ms' % (d['ms_per_scan_device'], d['ms_per_scan_e2e']), d['stages_ms'], d.get('roofline_spectrum', {}).get('frac_of_issue_floor'))
PY
for lib in cosmo_pol_b200/libradop.so cosmo_pol_b200/libradop_pair.so cosmo_pol_b200/libradop.so cosmo_pol_b200/libradop_pair.so; do
  RADOP_LIB=$PWD/$lib timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --no-direct > gpurun_out/bench_v.json 2> gpurun_out/bench_v.err || tail -3 gpurun_out/bench_v.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/bench_v.json').read().strip().splitlines()[-1])
print('$lib value %.3e ms %.3f e2e ms %.3f (rep %.3f) fused %.3f'%(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e']['ms_per_step_repeated_scan'], d['stages_ms']['interp_psd']))
PY
done
RADOP_LIB=$PWD/cosmo_pol_b200/libradop_pair.so timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "fused or observables or goldens or full_size_volume or table_vs" > gpurun_out/l_tests2.log 2>&1; echo "pair-lib tests rc=$?"; tail -3 gpurun_out/l_tests2.log
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_spectrum$ -s 1 -c 1 -f -o gpurun_out/prof_spectrum python scripts/bench_configs.py cfg5s --steps 2 > gpurun_out/ncu_spec.log 2>&1
echo "ncu spectrum rc=$?"
