#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "spectrum or cfg5 or fused" > gpurun_out/spec_tests.log 2>&1; echo "spectrum tests rc=$?"; tail -4 gpurun_out/spec_tests.log
timeout 900 python scripts/bench_configs.py cfg5s cfg5 cfg1 cfg2 cfg4 --steps 5 > gpurun_out/other_configs.json 2> gpurun_out/other_configs.err; echo "configs rc=$?"
python - <<'PY'
import json
for l in open('gpurun_out/other_configs.json'):
    try: d = json.loads(l)
    except Exception: continue
    print(d['config'], 'device %.2f ms e2e %.2f ms' % (d['ms_per_scan_device'], d['ms_per_scan_e2e']), d['stages_ms'])
PY
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu --no-direct > gpurun_out/bench_g.json 2> gpurun_out/bench_g.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_g.json').read().strip().splitlines()[-1])
print('bench value %.3e ms %.3f e2e ms %.3f fused %.3f roof %s'%(d['value'], d['ms_per_step'], d['e2e']['ms_per_step'], d['stages_ms']['interp_psd'], {k: d['roofline'][k] for k in ('bound','kernel','frac')}))
PY
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_spectrum$ -s 1 -c 1 -f -o gpurun_out/prof_spectrum python scripts/bench_configs.py cfg5s --steps 2 > gpurun_out/ncu_spec.log 2>&1
echo "ncu spectrum rc=$?"
