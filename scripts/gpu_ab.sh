#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "spectrum or cfg5" > gpurun_out/ab_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/ab_tests.log
timeout 600 python scripts/bench_configs.py cfg5s cfg5 --steps 5 2>/dev/null > gpurun_out/cfg5_configs.json; python - <<'PY'
import json
for l in open('gpurun_out/cfg5_configs.json'):
    try: d = json.loads(l)
    except Exception: continue
    print(d['config'], 'device %.2f ms e2e %.2f ms' % (d['ms_per_scan_device'], d['ms_per_scan_e2e']), d['stages_ms'], d.get('roofline_spectrum', {}).get('frac_of_issue_floor'))
PY
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_spectrum$ -s 1 -c 1 -f -o gpurun_out/prof_spectrum python scripts/bench_configs.py cfg5s --steps 2 > gpurun_out/ncu_spec.log 2>&1
echo "ncu spectrum rc=$?"
