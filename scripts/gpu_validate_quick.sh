#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -x -q -k "scan_level or fused or full_size or edge or no_alloc or melting" > gpurun_out/t_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/t_tests.log
timeout 600 python bench.py > gpurun_out/bench_t.json 2> gpurun_out/bench_t.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_t.json').read().strip().splitlines()[-1])
print('bench value %.3e ms %.3f e2e ms %.3f (repeated %.3f) fused %.3f launches %d'%(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d['e2e']['ms_per_step_repeated_scan'], d['stages_ms']['interp_psd'], d['gpu_launches']))
PY
timeout 900 python scripts/margins.py > gpurun_out/margins.json 2> gpurun_out/margins.err; echo "margins rc=$?"
timeout 600 python scripts/bench_configs.py cfg1 cfg2 cfg4 cfg5 --steps 5 > gpurun_out/other_configs.json 2> gpurun_out/other_configs.err; echo "configs rc=$?"
python - <<'PY'
import json
for l in open('gpurun_out/other_configs.json'):
    try: d = json.loads(l)
    except Exception: continue
    print(d['config'], 'device %.2f ms e2e %.2f ms' % (d['ms_per_scan_device'], d['ms_per_scan_e2e']), d.get('roofline_spectrum', {}).get('frac_of_issue_floor'))
PY
