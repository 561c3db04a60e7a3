#!/bin/bash
mkdir -p gpurun_out
for lib in cosmo_pol_b200/libradop_nopf.so cosmo_pol_b200/libradop.so cosmo_pol_b200/libradop_nopf.so cosmo_pol_b200/libradop.so; do
  RADOP_LIB=$PWD/$lib timeout 600 python scripts/bench_configs.py cfg5s --steps 5 2>/dev/null | python -c "
import json,sys
for l in sys.stdin:
    try: d=json.loads(l)
    except Exception: continue
    print('$lib', d['config'], 'device %.2f ms'%d['ms_per_scan_device'], 'spectrum %.2f'%d['stages_ms']['ms_spectrum'])"
done
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "spectrum or cfg5" > gpurun_out/ab_tests.log 2>&1; echo "tests rc=$?"; tail -3 gpurun_out/ab_tests.log
